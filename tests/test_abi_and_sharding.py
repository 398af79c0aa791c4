"""CPU tests: the C-ABI library loads and exports every symbol include/*.h declares; the image
sharding used for N > 1 GPUs, exercised with a world_size-2 gloo group."""
import ctypes as C
import os
import re
import socket
import subprocess
import sys
import textwrap

import pytest

import image_codecs_b200 as ic
from image_codecs_b200 import shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for hdr in ("image_api.h", "jpeg_cuda_abi.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        src = re.sub(r"#define[^\n]*", "", src)
        for m in re.finditer(r"\b(?:STBIDEF|JCU_API)\b[^;{]*?\b(\w+)\s*\(", src):
            names.add(m.group(1))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) >= 40, names
    lib = C.CDLL(ic.lib_path())
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_abi_struct_sizes_match_python_binding():
    from image_codecs_b200 import api

    assert C.sizeof(api.Params) == 13 * 4
    assert C.sizeof(api.Timings) == 6 * 4 + 3 * 8 + 6 * 4


def test_shard_helpers():
    assert [shard.contiguous_shard(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    sizes = [100, 1, 1, 1, 50, 50, 3, 97]
    parts = shard.balanced_shards(sizes, 3)
    assert sorted(i for p in parts for i in p) == list(range(len(sizes)))
    loads = [sum(sizes[i] for i in p) for p in parts]
    assert max(loads) - min(loads) <= max(sizes)
    assert shard.balanced_shards([], 2) == [[], []]
    with pytest.raises(ValueError):
        shard.contiguous_shard(4, 4, 4)


_WORKER = """
import os, sys, json
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
import image_codecs_b200 as ic
from image_codecs_b200 import shard
from tests import streams as S
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
cases = S.small_matrix()[:24]
datas = [S.stream(*c) for c in cases]
mine = shard.shard_for_rank([len(d) for d in datas], rank, world)
# the per-rank "work" on a CPU box is the host half of the path: full marker parse of the shard
px = 0
for i in mine:
    info = ic.inspect(datas[i], 3)
    assert info is not None
    px += info["width"] * info["height"]
t = torch.tensor([px, len(mine), sum(len(datas[i]) for i in mine)], dtype=torch.int64)
g = [torch.zeros_like(t) for _ in range(world)]
dist.all_gather(g, t)           # bookkeeping only (as bench.py's max/sum over ranks): no data-path collective
seen = [None] * world
dist.all_gather_object(seen, mine)
if rank == 0:
    flat = sorted(i for s in seen for i in s)
    assert flat == list(range(len(datas))), flat
    total_px = sum(int(x[0]) for x in g)
    assert total_px == sum(c[1] * c[2] for c in cases)
    loads = [int(x[2]) for x in g]
    assert max(loads) - min(loads) <= max(len(d) for d in datas)
    print("SHARD_OK", json.dumps(loads))
dist.destroy_process_group()
"""


def test_world_size_2_gloo_sharding(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(textwrap.dedent(_WORKER.format(root=ROOT)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "SHARD_OK" in outs[0]
