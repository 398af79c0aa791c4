/*
 * jpeg_coef_layout.cuh -- layout of the coefficient slab in HBM (this repo's own; the reference keeps
 * `short data[64]` per block on the stack, jpeg.c:1163, or in per-component coeff planes, jpeg.c:1649).
 *
 * One block = 64 int16 in natural (row-major) order = 128 bytes = eight 16-byte chunks (chunk r =
 * coefficient row r).  Chunk r of the block at device address A is stored at chunk position
 * r ^ ((A >> 7) & 7).  The permutation stays inside the block's own 128 bytes, so block writes and
 * bulk (TMA) block copies remain contiguous, while a warp whose lanes each read "their" block from a
 * shared-memory copy (one 16-byte chunk per lane per instruction) hits eight distinct bank groups
 * instead of one: the XOR swizzle is baked into HBM so a plain 1-D bulk copy needs no tensor map.
 */
#pragma once
#include <stdint.h>

__host__ __device__ __forceinline__ unsigned coef_key(const void *block_addr)
{
	return (unsigned)(((uintptr_t)block_addr >> 7) & 7u);
}
/* element index inside the stored block of natural-order coefficient `nat` (0..63) */
__host__ __device__ __forceinline__ int coef_pos(int nat, unsigned key) { return (((nat >> 3) ^ (int)key) << 3) | (nat & 7); }
