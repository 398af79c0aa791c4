/*
 * jpeg_fused1x1.cuh -- fused output stage for files whose decoded components are all sampled 1x1 (sm_100a):
 * 4:4:4 YCbCr (NC = 3) and single-component decodes (NC = 1: grayscale files, and YCbCr files asked for one
 * or two channels, jpeg.c:2246).  Strip-persistent like jpeg_fused420.cuh, without its chroma machinery:
 *
 *   dequantise -> integer 8x8 IDCT (stbi__idct_block, jpeg.c:615) -> [YCbCr->RGB, stbi__YCbCr_to_RGB_row,
 *   jpeg.c:1976 | grey replicate, jpeg.c:2409-2431] -> packed stores,
 *
 * reading each coefficient block once and writing each pixel once.  A CTA owns a vertical strip of F1_BS(NC)
 * blocks inside a band of block rows and walks down it one block row (8 pixel rows) per iteration:
 *
 *   TMA    one thread arms an mbarrier and issues NC bulk copies (one contiguous block row per component);
 *          the copies of row r+1 are issued as soon as the IDCT phase of row r has drained the staging buffer.
 *   IDCT   one thread = one 8x8 block, 64 coefficients in registers; samples go to NC planes of 8 rows.
 *   pixel  one task = 8 pixels of one row.
 *
 * resample_row_1 (jpeg.c:1765) is the identity, so there is no halo of any kind.
 * Algorithmic HBM bytes: 128 B x coefficient blocks + n B x pixels.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"
#include "jpeg_fused420.cuh"

#define F1_THREADS 128
#define F1_ROWS_PER_BAND 8 /* block rows a CTA walks */
#define F1_BS(NC) ((NC) == 3 ? 40 : 128) /* blocks per strip row: NC * F1_BS <= F1_THREADS */
#define F1_OFF_S(NC) ((NC)*F1_BS(NC) * 128)
#define F1_OFF_Q(NC) (F1_OFF_S(NC) + (NC)*8 * F1_BS(NC) * 8)
#define F1_OFF_BAR(NC) (F1_OFF_Q(NC) + 3 * 64 * 2)
#define F1_SMEM_BYTES(NC) (F1_OFF_BAR(NC) + 16)

/* one pixel from the three sample bytes (selected by SEL from the plane words): stbi__YCbCr_to_RGB_row
   (jpeg.c:1981-1990) with the common factor 2^8 divided out and the -128 of both chroma channels folded
   into the constants of the luma term */
template <int SEL>
__device__ __forceinline__ void f1_colour(uint32_t yw, uint32_t cbw, uint32_t crw, int &r, int &g, int &b)
{
	const int y = s420_byte<0x4440 + SEL>(yw);
	const int cb = s420_byte<0x4440 + SEL>(cbw);
	const int cr = s420_byte<0x4440 + SEL>(crw);
	/* y + (t >> 12) is one LEA.HI.SX32; the -128 of both chroma samples lives in the constants */
	r = y + (s420_mad(cr, 5743, 2048 - 128 * 5743) >> 12);
	b = y + (s420_mad(cb, 7258, 2048 - 128 * 7258) >> 12);
	/* Cr's 128 * 2925 = 1462 * 256 + 128: the multiple of 256 moves inside the mask, the 128 is OR-ed into the
	   cleared bits by the same LOP3 */
	int m, k128;
	asm("mov.u32 %0, 128;" : "=r"(k128));
	asm("lop3.b32 %0, %1, 0xffffff00, %2, 0xEA;" : "=r"(m) : "r"(s420_mad(cb, -1410, 128 * 1410 + 2048 + 1462 * 256)), "r"(k128));
	g = y + (s420_mad(cr, -2925, m) >> 12);
}

template <int N, int NC, bool DEQ>
__global__ void __launch_bounds__(F1_THREADS, 4)
k_fused1x1_strip(const int32_t *__restrict__ list, const jcu_image *__restrict__ images, const int16_t *__restrict__ coef,
					  uint8_t *__restrict__ out)
{
	constexpr int BS = F1_BS(NC);
	extern __shared__ __align__(128) unsigned char f1_smem[];
	uint4 *sCoef = reinterpret_cast<uint4 *>(f1_smem);
	uint8_t *sS = f1_smem + F1_OFF_S(NC); /* [NC][8][BS * 8] */
	uint16_t(*sQ)[64] = reinterpret_cast<uint16_t(*)[64]>(f1_smem + F1_OFF_Q(NC));
	uint64_t *bar = reinterpret_cast<uint64_t *>(f1_smem + F1_OFF_BAR(NC));

	const jcu_image *im = &images[list[blockIdx.y]];
	const int strips = im->tiles_x;
	if ((int)blockIdx.x >= strips * im->tiles_y)
		return;
	const int strip = blockIdx.x % strips, band = blockIdx.x / strips;
	const int tid = threadIdx.x;
	const int W = im->width, H = im->height;
	const int bx0 = strip * BS, nb = min(BS, im->comp[0].bw - bx0);
	const int by0 = band * F1_ROWS_PER_BAND, by1 = min(by0 + F1_ROWS_PER_BAND, (H + 7) >> 3);

	if (tid == 0)
		s420_mbar_init(bar, 1);
	if (DEQ)
		for (int i = tid; i < NC * 64; i += F1_THREADS)
			sQ[i >> 6][i & 63] = im->qt_out[i >> 6][i & 63];
	__syncthreads();

	auto issue = [&](int by) {
		const unsigned bytes = (unsigned)(nb * 128);
		s420_mbar_expect_tx(bar, (unsigned)NC * bytes);
#pragma unroll
		for (int c = 0; c < NC; ++c)
			s420_bulk_g2s(reinterpret_cast<unsigned char *>(sCoef) + c * BS * 128,
							  coef + im->comp[c].coef_off + ((size_t)by * im->comp[c].bw + bx0) * 64, bytes, bar);
	};
	if (tid == 0 && by0 < by1)
		issue(by0);
	unsigned parity = 0;

	/* this thread's IDCT task, the same every iteration */
	const int ci = NC == 3 ? tid / BS : 0, lb = tid - ci * BS;
	const bool valid = ci < NC && lb < nb;
	const int cbw_i = im->comp[valid ? ci : 0].bw;
	const int16_t *gbase = coef + im->comp[valid ? ci : 0].coef_off + (size_t)(bx0 + lb) * 64;

	const bool vec_ok = (W & 7) == 0; /* every row starts 8-byte aligned for N = 1, 3; N = 2, 4 need less */
	const int G = nb;						 /* 8-pixel groups per row of the strip */
	const int pitch = im->flip ? -W * N : W * N; /* flip on load as a store epilogue, as in jpeg_fused420.cuh */
	uint8_t *obase = out + im->out_off + (im->flip ? (size_t)(im->height - 1) * (size_t)(W * N) : 0);

#pragma unroll 1
	for (int by = by0; by < by1; ++by)
	{
		s420_mbar_wait(bar, parity);
		parity ^= 1u;
		if (valid)
		{
			int v[64];
			uint32_t rows[16];
			const int16_t *g = gbase + (size_t)by * cbw_i * 64;
			s420_load_block<DEQ>(&sCoef[(ci * BS + lb) * 8], coef_key(g), sQ[ci], v);
			s420_idct(v, rows);
			uint8_t *d = sS + (ci * 8) * (BS * 8) + lb * 8;
#pragma unroll
			for (int rr = 0; rr < 8; ++rr)
				*reinterpret_cast<uint2 *>(d + rr * (BS * 8)) = make_uint2(rows[2 * rr], rows[2 * rr + 1]);
		}
		__syncthreads();
		if (tid == 0 && by + 1 < by1)
			issue(by + 1);

		/* ---- pixel phase: 8 rows x G groups of 8 pixels ---- */
#pragma unroll 1
		for (int t = tid; t < 8 * G; t += F1_THREADS)
		{
			const int ly = t / G, g = t - ly * G;
			const int gy = by * 8 + ly, gx0 = (bx0 + g) * 8;
			if (gy >= H || gx0 >= W)
				continue;
			const int count = min(8, W - gx0);
			const bool whole = vec_ok && count == 8;
			uint8_t *o = obase + (long long)gy * pitch + gx0 * N;
			const uint2 yv = *reinterpret_cast<const uint2 *>(sS + ly * (BS * 8) + g * 8);
			if (NC == 3)
			{
				const uint2 bv = *reinterpret_cast<const uint2 *>(sS + (8 + ly) * (BS * 8) + g * 8);
				const uint2 rv = *reinterpret_cast<const uint2 *>(sS + (16 + ly) * (BS * 8) + g * 8);
				int R[8], Gc[8], B[8];
				f1_colour<0>(yv.x, bv.x, rv.x, R[0], Gc[0], B[0]);
				f1_colour<1>(yv.x, bv.x, rv.x, R[1], Gc[1], B[1]);
				f1_colour<2>(yv.x, bv.x, rv.x, R[2], Gc[2], B[2]);
				f1_colour<3>(yv.x, bv.x, rv.x, R[3], Gc[3], B[3]);
				f1_colour<0>(yv.y, bv.y, rv.y, R[4], Gc[4], B[4]);
				f1_colour<1>(yv.y, bv.y, rv.y, R[5], Gc[5], B[5]);
				f1_colour<2>(yv.y, bv.y, rv.y, R[6], Gc[6], B[6]);
				f1_colour<3>(yv.y, bv.y, rv.y, R[7], Gc[7], B[7]);
				if (whole)
				{
					if (N == 3)
					{
						uint32_t w[6];
#pragma unroll
						for (int q = 0; q < 2; ++q)
						{
							const int a = 4 * q;
							w[3 * q] = s420_pack4(R[a], Gc[a], B[a], R[a + 1]);
							w[3 * q + 1] = s420_pack4(Gc[a + 1], B[a + 1], R[a + 2], Gc[a + 2]);
							w[3 * q + 2] = s420_pack4(B[a + 2], R[a + 3], Gc[a + 3], B[a + 3]);
						}
						uint2 *ow = reinterpret_cast<uint2 *>(o);
						ow[0] = make_uint2(w[0], w[1]);
						ow[1] = make_uint2(w[2], w[3]);
						ow[2] = make_uint2(w[4], w[5]);
					}
					else
					{
#pragma unroll
						for (int q = 0; q < 2; ++q)
							reinterpret_cast<uint4 *>(o)[q] = make_uint4(
								 s420_pack4(R[4 * q], Gc[4 * q], B[4 * q], 255), s420_pack4(R[4 * q + 1], Gc[4 * q + 1], B[4 * q + 1], 255),
								 s420_pack4(R[4 * q + 2], Gc[4 * q + 2], B[4 * q + 2], 255),
								 s420_pack4(R[4 * q + 3], Gc[4 * q + 3], B[4 * q + 3], 255));
					}
				}
				else
				{
#pragma unroll
					for (int p = 0; p < 8; ++p)
						if (p < count)
						{
							o[N * p] = (uint8_t)clamp_u8(R[p]);
							o[N * p + 1] = (uint8_t)clamp_u8(Gc[p]);
							o[N * p + 2] = (uint8_t)clamp_u8(B[p]);
							if (N == 4)
								o[N * p + 3] = 255;
						}
				}
			}
			else if (whole)
			{
				/* grey -> 1..4 channels (jpeg.c:2409-2431): replicate, alpha = 255 */
				if (N == 1)
					*reinterpret_cast<uint2 *>(o) = yv;
				else if (N == 2)
					*reinterpret_cast<uint4 *>(o) = make_uint4(__byte_perm(yv.x, 0xffu, 0x4140), __byte_perm(yv.x, 0xffu, 0x4342),
																			 __byte_perm(yv.y, 0xffu, 0x4140), __byte_perm(yv.y, 0xffu, 0x4342));
				else if (N == 3)
				{
					uint2 *ow = reinterpret_cast<uint2 *>(o);
					ow[0] = make_uint2(__byte_perm(yv.x, 0, 0x1000), __byte_perm(yv.x, 0, 0x2211));
					ow[1] = make_uint2(__byte_perm(yv.x, 0, 0x3332), __byte_perm(yv.y, 0, 0x1000));
					ow[2] = make_uint2(__byte_perm(yv.y, 0, 0x2211), __byte_perm(yv.y, 0, 0x3332));
				}
				else
				{
					reinterpret_cast<uint4 *>(o)[0] = make_uint4(__byte_perm(yv.x, 0xffu, 0x4000), __byte_perm(yv.x, 0xffu, 0x4111),
																				__byte_perm(yv.x, 0xffu, 0x4222), __byte_perm(yv.x, 0xffu, 0x4333));
					reinterpret_cast<uint4 *>(o)[1] = make_uint4(__byte_perm(yv.y, 0xffu, 0x4000), __byte_perm(yv.y, 0xffu, 0x4111),
																				__byte_perm(yv.y, 0xffu, 0x4222), __byte_perm(yv.y, 0xffu, 0x4333));
				}
			}
			else
			{
				const uint32_t yw[2] = {yv.x, yv.y};
#pragma unroll
				for (int p = 0; p < 8; ++p)
					if (p < count)
					{
						const uint8_t y = (uint8_t)(yw[p >> 2] >> (8 * (p & 3)));
						o[N * p] = y;
						if (N >= 3)
						{
							o[N * p + 1] = y;
							o[N * p + 2] = y;
						}
						if (N == 2 || N == 4)
							o[N * p + N - 1] = 255;
					}
			}
		}
		__syncthreads();
	}
}

/* host-side launch: grid = (max strips*bands per image, images on the path) */
template <int NC>
static inline cudaError_t fused1x1_launch(int max_tiles, const int32_t *list, int nlist, const jcu_image *images, const int16_t *coef,
														uint8_t *out, cudaStream_t st, int n_out, int deq)
{
	if (max_tiles <= 0 || nlist <= 0)
		return cudaSuccess;
	void (*k)(const int32_t *, const jcu_image *, const int16_t *, uint8_t *);
	if (NC == 3)
		k = n_out == 4 ? (deq ? k_fused1x1_strip<4, NC, true> : k_fused1x1_strip<4, NC, false>)
							: (deq ? k_fused1x1_strip<3, NC, true> : k_fused1x1_strip<3, NC, false>);
	else
		k = n_out == 1	  ? (deq ? k_fused1x1_strip<1, NC, true> : k_fused1x1_strip<1, NC, false>)
			 : n_out == 2 ? (deq ? k_fused1x1_strip<2, NC, true> : k_fused1x1_strip<2, NC, false>)
			 : n_out == 3 ? (deq ? k_fused1x1_strip<3, NC, true> : k_fused1x1_strip<3, NC, false>)
							  : (deq ? k_fused1x1_strip<4, NC, true> : k_fused1x1_strip<4, NC, false>);
	k<<<dim3((unsigned)max_tiles, (unsigned)nlist), F1_THREADS, F1_SMEM_BYTES(NC), st>>>(list, images, coef, out);
	return cudaGetLastError();
}
