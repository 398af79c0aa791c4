"""image_codecs_b200 -- B200-native JPEG decode behind the reference's stbi_* API.

Thin ctypes binding over the C-ABI library ``libimage_codecs_b200.so`` (host C marker parser +
hand-written sm_100a CUDA kernels, sources in ``csrc/``).  Function names, argument meaning and
error behaviour mirror the reference's public entry points (convert.c:188-266,
image_api.c:74-131); the batched decoder is new (BASELINE.json north_star).

There is no CPU decode path: if the library is missing, or no CUDA device is usable, loads raise.
"""
from .api import (  # noqa: F401
    BatchDecoder,
    MultiDecoder,
    LibraryMissing,
    cuda_device_count,
    encode_batch,
    failure_reason,
    inspect,
    lib,
    lib_path,
    set_flip_vertically_on_load,
    stbi_info_from_memory,
    stbi_load,
    stbi_load_from_memory,
    write_jpg_to_bytes,
)

__all__ = [
    "BatchDecoder", "MultiDecoder", "LibraryMissing", "cuda_device_count", "encode_batch", "write_jpg_to_bytes", "failure_reason", "inspect", "lib", "lib_path",
    "set_flip_vertically_on_load", "stbi_info_from_memory", "stbi_load", "stbi_load_from_memory",
]
