/*
 * jpeg_fused422.cuh -- fused output stage for 4:2:2 YCbCr (luma 2x1, chroma 1x1; what most cameras write), sm_100a.
 *
 * One pass over HBM, like jpeg_fused420.cuh / jpeg_fused1x1.cuh: dequantise -> integer 8x8 IDCT
 * (stbi__idct_block, jpeg.c:615) -> horizontal chroma upsample (stbi__resample_row_h_2, jpeg.c:1784-1812) ->
 * YCbCr->RGB (stbi__YCbCr_to_RGB_row, jpeg.c:1976) -> packed RGB / RGBA stores.  It replaces the generic two-pass
 * route (k_generic_idct + k_generic_color through u8 sample planes) for this sampling.
 *
 * A CTA owns a vertical strip of F2_MS MCUs (16 x 8 pixels each) inside a band of MCU rows and walks down it one MCU
 * row (8 pixel rows) per iteration: three bulk copies (one contiguous luma block row; the Cb and the Cr block row
 * with one halo block per side), one block per thread IDCT into byte planes, then 8-pixel tasks.  h_2 has no
 * vertical reach, so there is no row halo.
 *
 * The resampler, as the reference writes it (w = w_lores):
 *     out[0] = in[0]                      out[2i]   = (3 in[i] + in[i-1] + 2) >> 2      (0 < i < w-1)
 *     out[1] = (3 in[0] + in[1] + 2) >> 2 out[2i+1] = (3 in[i] + in[i+1] + 2) >> 2
 *     out[2w-2] = (3 in[w-2] + in[w-1] + 2) >> 2   <- weights the sample BEFORE the last one (jpeg.c:1805)
 *     out[2w-1] = in[w-1];   w == 1: out[0] = out[1] = in[0]
 * With the column before the first and behind the last real sample replicated, the general pair of formulas gives
 * every output except out[2w-2], which is patched.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"
#include "jpeg_fused420.cuh"

#define F2_THREADS 128
#define F2_MS 30					/* MCUs per strip: 60 luma + 2 x 32 chroma blocks = 124 IDCT tasks */
#define F2_MC (F2_MS + 2)		/* chroma block positions incl. one halo each side */
#define F2_ROWS_PER_BAND 16	/* MCU rows a CTA walks */
#define F2_SLOTS (2 * F2_MS + 2 * F2_MC)
#define F2_YW (F2_MS * 16)
#define F2_CW (F2_MC * 8)
#define F2_OFF_Y (F2_SLOTS * 128)
#define F2_OFF_CB (F2_OFF_Y + 8 * F2_YW)
#define F2_OFF_CR (F2_OFF_CB + 8 * F2_CW)
#define F2_OFF_Q (F2_OFF_CR + 8 * F2_CW)
#define F2_OFF_BAR (F2_OFF_Q + 3 * 64 * 2)
#define F2_SMEM_BYTES (F2_OFF_BAR + 16)

/* one pixel, integer samples: stbi__YCbCr_to_RGB_row (jpeg.c:1981-1990) with the common factor 2^8 divided out and the
   -128 of both chroma samples folded into the constants (derivation: f1_colour in jpeg_fused1x1.cuh) */
__device__ __forceinline__ void f2_colour(int y, int cb, int cr, int &r, int &g, int &b)
{
	r = y + (s420_mad(cr, 5743, 2048 - 128 * 5743) >> 12);
	b = y + (s420_mad(cb, 7258, 2048 - 128 * 7258) >> 12);
	const int m = (s420_mad(cb, -1410, 128 * 1410 + 2048 + 1462 * 256) & ~255) | 128;
	g = y + (s420_mad(cr, -2925, m) >> 12);
}

/* eight upsampled chroma samples from in[-1], in[0..3] (bytes of w), in[4] */
__device__ __forceinline__ void f2_upsample(uint32_t prev, uint32_t w, uint32_t next, int (&o)[8])
{
	const int a[6] = {(int)prev, (int)(w & 255u), (int)((w >> 8) & 255u), (int)((w >> 16) & 255u), (int)(w >> 24), (int)next};
#pragma unroll
	for (int k = 0; k < 4; ++k)
	{
		const int n = 3 * a[k + 1] + 2;
		o[2 * k] = (n + a[k]) >> 2;
		o[2 * k + 1] = (n + a[k + 2]) >> 2;
	}
}

template <int N, bool DEQ>
__global__ void __launch_bounds__(F2_THREADS, 4)
k_fused422_strip(const int32_t *__restrict__ list, const jcu_image *__restrict__ images, const int16_t *__restrict__ coef,
					  uint8_t *__restrict__ out)
{
	extern __shared__ __align__(128) unsigned char f2_smem[];
	uint4 *sCoef = reinterpret_cast<uint4 *>(f2_smem);
	uint8_t *sY = f2_smem + F2_OFF_Y;	/* [8][F2_YW] */
	uint8_t *sCb = f2_smem + F2_OFF_CB; /* [8][F2_CW]: sample i of the strip's first MCU column sits at 8 + i */
	uint8_t *sCr = f2_smem + F2_OFF_CR;
	uint16_t(*sQ)[64] = reinterpret_cast<uint16_t(*)[64]>(f2_smem + F2_OFF_Q);
	uint64_t *bar = reinterpret_cast<uint64_t *>(f2_smem + F2_OFF_BAR);

	const jcu_image *im = &images[list[blockIdx.y]];
	const int strips = im->tiles_x;
	if ((int)blockIdx.x >= strips * im->tiles_y)
		return;
	const int strip = blockIdx.x % strips, band = blockIdx.x / strips;
	const int tid = threadIdx.x;
	const int W = im->width, H = im->height;
	const int ybw = im->comp[0].bw, cbw = im->comp[1].bw, wl = im->comp[1].w_lores;
	const int m0 = strip * F2_MS, Ms = min(F2_MS, im->mcu_x - m0);
	const int by0 = band * F2_ROWS_PER_BAND, by1 = min(by0 + F2_ROWS_PER_BAND, (H + 7) >> 3);
	const int c0 = max(m0 - 1, 0), c1 = min(m0 + Ms + 1, cbw);
	const int16_t *baseY = coef + im->comp[0].coef_off;
	const int16_t *baseCb = coef + im->comp[1].coef_off;
	const int16_t *baseCr = coef + im->comp[2].coef_off;

	if (tid == 0)
		s420_mbar_init(bar, 1);
	if (DEQ)
		for (int i = tid; i < 192; i += F2_THREADS)
			sQ[i >> 6][i & 63] = im->qt_out[i >> 6][i & 63];
	__syncthreads();

	auto issue = [&](int by) {
		const unsigned ybytes = (unsigned)(2 * Ms * 128), cbytes = (unsigned)((c1 - c0) * 128);
		s420_mbar_expect_tx(bar, ybytes + 2u * cbytes);
		unsigned char *dst = reinterpret_cast<unsigned char *>(sCoef);
		s420_bulk_g2s(dst, baseY + ((size_t)by * ybw + 2 * m0) * 64, ybytes, bar);
		const size_t off = ((size_t)by * cbw + c0) * 64;
		const unsigned d = (unsigned)((2 * F2_MS + (c0 - (m0 - 1))) * 128);
		s420_bulk_g2s(dst + d, baseCb + off, cbytes, bar);
		s420_bulk_g2s(dst + d + F2_MC * 128, baseCr + off, cbytes, bar);
	};
	if (tid == 0 && by0 < by1)
		issue(by0);
	unsigned parity = 0;

	/* this thread's IDCT task, the same every iteration */
	const bool isY = tid < 2 * F2_MS;
	const int ct = tid - 2 * F2_MS, ci = ct >= F2_MC ? 1 : 0, cpos = ct - ci * F2_MC;
	const int cbx = m0 - 1 + cpos;
	const bool valid = isY ? tid < 2 * Ms : (ct < 2 * F2_MC && cbx >= c0 && cbx < c1);
	const int slot = min(tid, F2_SLOTS - 1);

	const bool vec_ok = (N == 4) ? ((W & 3) == 0) : ((W & 7) == 0);
	const int G = 2 * Ms;
	const int pitch = im->flip ? -W * N : W * N;
	uint8_t *obase = out + im->out_off + (im->flip ? (size_t)(H - 1) * (size_t)(W * N) : 0);
	const int xfix = wl >= 2 ? 2 * wl - 2 : -1; /* the one output the general formulas do not give (jpeg.c:1805) */

#pragma unroll 1
	for (int by = by0; by < by1; ++by)
	{
		s420_mbar_wait(bar, parity);
		parity ^= 1u;
		if (valid)
		{
			int v[64];
			uint32_t rows[16];
			const int16_t *g = isY ? baseY + ((size_t)by * ybw + 2 * m0 + tid) * 64 : (ci ? baseCr : baseCb) + ((size_t)by * cbw + cbx) * 64;
			s420_load_block<DEQ>(&sCoef[slot * 8], coef_key(g), sQ[isY ? 0 : 1 + ci], v);
			s420_idct(v, rows);
			uint8_t *d = isY ? sY + tid * 8 : (ci ? sCr : sCb) + cpos * 8;
			const int stride = isY ? F2_YW : F2_CW;
#pragma unroll
			for (int rr = 0; rr < 8; ++rr)
				*reinterpret_cast<uint2 *>(d + rr * stride) = make_uint2(rows[2 * rr], rows[2 * rr + 1]);
			/* replicate the column before the first and behind the last REAL chroma sample of the image row (w_lores,
			   jpeg.c:2276): the same thread wrote the source sample */
			if (!isY && (cbx == 0 || cbx == cbw - 1))
			{
				/* a chroma block spans one MCU, so the last real sample always lies in the row's last block: 1 <= e <= 8,
				   and position e = 8 belongs to a halo block nobody loads */
				const int e = wl - cbx * 8;
#pragma unroll
				for (int rr = 0; rr < 8; ++rr)
				{
					uint8_t *p = d + rr * stride;
					if (cbx == 0)
						p[-1] = p[0];
					if (cbx == cbw - 1 && cpos * 8 + e < F2_CW) /* as a right-halo block its column 8 lies outside the plane row, and nobody reads it */
						p[e] = p[e - 1];
				}
			}
		}
		__syncthreads();
		if (tid == 0 && by + 1 < by1)
			issue(by + 1);

		/* ---- pixel phase: 8 rows x G groups of 8 pixels ---- */
#pragma unroll 1
		for (int t = tid; t < 8 * G; t += F2_THREADS)
		{
			const int ly = t / G, g = t - ly * G;
			const int gy = by * 8 + ly, gx0 = (m0 * 2 + g) * 8;
			if (gy >= H || gx0 >= W)
				continue;
			const int count = min(8, W - gx0);
			const bool whole = vec_ok && count == 8;
			uint8_t *o = obase + (long long)gy * pitch + gx0 * N;
			const uint2 yv = *reinterpret_cast<const uint2 *>(sY + ly * F2_YW + g * 8);
			const int P = 8 + g * 4; /* plane position of the group's first chroma sample */
			const uint32_t *pb = reinterpret_cast<const uint32_t *>(sCb + ly * F2_CW + P);
			const uint32_t *pr = reinterpret_cast<const uint32_t *>(sCr + ly * F2_CW + P);
			int cb[8], cr[8];
			f2_upsample(pb[-1] >> 24, pb[0], pb[1] & 255u, cb);
			f2_upsample(pr[-1] >> 24, pr[0], pr[1] & 255u, cr);
			if (xfix >= gx0 && xfix < gx0 + 8)
			{
				/* out[2w-2] = (3 in[w-2] + in[w-1] + 2) >> 2 */
				const int q = 8 + (wl - 2) - m0 * 8; /* plane position of in[w-2]; it may lie in the left halo block */
				const uint8_t *rb = sCb + ly * F2_CW, *rr_ = sCr + ly * F2_CW;
				cb[xfix - gx0] = (3 * rb[q] + rb[q + 1] + 2) >> 2;
				cr[xfix - gx0] = (3 * rr_[q] + rr_[q + 1] + 2) >> 2;
			}
			int R[8], Gc[8], B[8];
#pragma unroll
			for (int p = 0; p < 8; ++p)
				f2_colour((int)(((p < 4 ? yv.x : yv.y) >> (8 * (p & 3))) & 255u), cb[p], cr[p], R[p], Gc[p], B[p]);
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
		__syncthreads();
	}
}

/* host-side launch: grid = (max strips*bands per image, images on the path) */
static inline cudaError_t fused422_launch(int max_tiles, const int32_t *list, int nlist, const jcu_image *images, const int16_t *coef,
														uint8_t *out, cudaStream_t st, int n_out, int deq)
{
	if (max_tiles <= 0 || nlist <= 0)
		return cudaSuccess;
	void (*k)(const int32_t *, const jcu_image *, const int16_t *, uint8_t *) =
		 n_out == 4 ? (deq ? k_fused422_strip<4, true> : k_fused422_strip<4, false>)
						: (deq ? k_fused422_strip<3, true> : k_fused422_strip<3, false>);
	k<<<dim3((unsigned)max_tiles, (unsigned)nlist), F2_THREADS, F2_SMEM_BYTES, st>>>(list, images, coef, out);
	return cudaGetLastError();
}
