/*
 * jpeg_fused.cuh -- the fused output-stage tile kernels (sm_100a).
 *
 * ONE kernel per headline format does, per output tile and without touching HBM in between:
 *   dequantise (short)(c*q)  ->  integer 8x8 IDCT (stbi__idct_block, jpeg.c:615)
 *   ->  chroma upsample (stbi__resample_row_hv_2, jpeg.c:1816)  ->  YCbCr->RGB (jpeg.c:1976)
 *   ->  packed RGB/RGBA stores.
 * It replaces the reference's per-block IDCT calls (jpeg.c:1178,1217,1342) plus the whole
 * row loop of load_jpeg_image (jpeg.c:2301-2432) for these formats.
 *
 * Layout per CTA (256 threads):
 *   phase 1  one thread = one 8x8 block: 128 B of coefficients -> registers -> dequant ->
 *            column pass -> row pass -> 8 rows of 8 samples into the tile's sample planes in
 *            shared memory.  For 4:2:0 the Cb and Cr blocks of one position sit in adjacent
 *            lanes, which exchange rows by warp shuffle and store (Cb | Cr<<16) pairs, so the
 *            upsampling filter runs on both chroma components in one 32-bit lane.
 *            The chroma planes carry a one-block halo ring: the hv_2 filter interpolates across
 *            block and tile borders (jpeg.c:11), so neighbour chroma blocks are re-derived from
 *            their coefficients (L2 hits) instead of spilling planes to HBM.
 *   phase 2  one thread = 16 consecutive pixels of one row: Y from shared memory, chroma
 *            filtered with clamp-to-edge on the REAL component size (w_lores, cy -- never the MCU
 *            padding), colour conversion, 48 (RGB) or 64 (RGBA) bytes stored as
 *            16-byte vectors.
 * No tensor cores: the IDCT must reproduce the reference's fixed-point rounding.
 *
 * Algorithmic HBM bytes per tile: 128 B x own blocks read + n B x pixels written.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_output.cuh"

#define FUSED_THREADS 256

/* tile sizes in output pixels */
#define F420_TW 128
#define F420_TH 64
#define F444_TW 128
#define F444_TH 32
#define FGRAY_TW 256
#define FGRAY_TH 64

static inline void fused_tile_dims(int path, int *tw, int *th)
{
	switch (path)
	{
	case JCU_PATH_FUSED_420: *tw = F420_TW; *th = F420_TH; break;
	case JCU_PATH_FUSED_444: *tw = F444_TW; *th = F444_TH; break;
	default: *tw = FGRAY_TW; *th = FGRAY_TH; break;
	}
}

/* find the image a tile belongs to: list[] is ordered by tile_first */
__device__ __forceinline__ const jcu_image *fused_locate(const int32_t *__restrict__ list, int nlist,
																			const jcu_image *__restrict__ images, int tile, int &local)
{
	int lo = 0, hi = nlist - 1;
	while (lo < hi)
	{
		int mid = (lo + hi + 1) >> 1;
		if (images[list[mid]].tile_first <= tile)
			lo = mid;
		else
			hi = mid - 1;
	}
	const jcu_image *im = &images[list[lo]];
	local = tile - im->tile_first;
	return im;
}

/* write `count` (<=16) pixels of n channels starting at o; rgb[] holds 16 x (r | g<<8 | b<<16) */
template <int N>
__device__ __forceinline__ void store_pixels(uint8_t *o, const uint32_t (&px)[16], int count, bool vec_ok)
{
	if (N == 4)
	{
		if (vec_ok && count == 16)
		{
#pragma unroll
			for (int q = 0; q < 4; ++q)
				reinterpret_cast<uint4 *>(o)[q] = make_uint4(px[4 * q] | 0xff000000u, px[4 * q + 1] | 0xff000000u,
																			px[4 * q + 2] | 0xff000000u, px[4 * q + 3] | 0xff000000u);
			return;
		}
		for (int i = 0; i < count; ++i)
		{
			o[4 * i] = (uint8_t)px[i];
			o[4 * i + 1] = (uint8_t)(px[i] >> 8);
			o[4 * i + 2] = (uint8_t)(px[i] >> 16);
			o[4 * i + 3] = 255;
		}
	}
	else
	{
		if (vec_ok && count == 16)
		{
#pragma unroll
			for (int q = 0; q < 4; ++q)
			{
				/* 4 pixels -> 3 words: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3 */
				uint32_t a = px[4 * q], b = px[4 * q + 1], c = px[4 * q + 2], d = px[4 * q + 3];
				uint32_t w0 = (a & 0x00ffffffu) | (b << 24);
				uint32_t w1 = ((b >> 8) & 0x0000ffffu) | (c << 16);
				uint32_t w2 = ((c >> 16) & 0x000000ffu) | (d << 8);
				uint32_t *ow = reinterpret_cast<uint32_t *>(o) + 3 * q;
				ow[0] = w0;
				ow[1] = w1;
				ow[2] = w2;
			}
			return;
		}
		for (int i = 0; i < count; ++i)
		{
			o[3 * i] = (uint8_t)px[i];
			o[3 * i + 1] = (uint8_t)(px[i] >> 8);
			o[3 * i + 2] = (uint8_t)(px[i] >> 16);
		}
	}
}

__device__ __forceinline__ uint32_t rgb_pack(int y, int cb, int cr)
{
	int r, g, b;
	ycbcr_to_rgb(y, cb, cr, r, g, b);
	return (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16);
}

/* ===================================================================================== */
/* 4:2:0  (Y 2x2, Cb 1x1, Cr 1x1)  YCbCr -> RGB / RGBA                                    */
/* ===================================================================================== */
#define F420_CBW 10 /* chroma blocks across incl. halo */
#define F420_CBH 6
#define F420_CW (F420_CBW * 8)
#define F420_CH (F420_CBH * 8)

template <int N>
__global__ void __launch_bounds__(FUSED_THREADS, 2)
k_fused_420(const int32_t *__restrict__ list, int nlist, const jcu_image *__restrict__ images,
				const int16_t *__restrict__ coef, uint8_t *__restrict__ out)
{
	__shared__ __align__(16) uint8_t sY[F420_TH][F420_TW];
	__shared__ __align__(16) uint32_t sC[F420_CH][F420_CW]; /* Cb | Cr << 16 */

	int local;
	const jcu_image *im = fused_locate(list, nlist, images, blockIdx.x, local);
	const int tx = local % im->tiles_x, ty = local / im->tiles_x;
	const int W = im->width, H = im->height;
	const int tid = threadIdx.x;

	/* ---- phase 1: IDCT ---- */
	{
		int v[64];
		uint32_t rows[16];
		if (tid < 128)
		{
			/* luma: 16 x 8 blocks */
			const jcu_comp *cp = &im->comp[0];
			const int lbx = tid & 15, lby = tid >> 4;
			const int bx = tx * 16 + lbx, by = ty * 8 + lby;
			if (bx < cp->bw && by < cp->bh)
			{
				load_dequant(coef + cp->coef_off + ((size_t)by * cp->bw + bx) * 64, im->qt[0], v);
				idct_block_regs(v, rows);
#pragma unroll
				for (int r = 0; r < 8; ++r)
					*reinterpret_cast<uint2 *>(&sY[lby * 8 + r][lbx * 8]) = make_uint2(rows[2 * r], rows[2 * r + 1]);
			}
		}
		else
		{
			/* chroma: lane pairs (Cb, Cr) of one block position, halo ring included */
			const int item = tid - 128;
			const int ci = 1 + (item & 1), pos = item >> 1;
			const int lbx = pos % F420_CBW, lby = pos / F420_CBW;
			const jcu_comp *cp = &im->comp[ci];
			const int bx = tx * 8 - 1 + lbx, by = ty * 4 - 1 + lby;
			const bool have = item < 2 * F420_CBW * F420_CBH && bx >= 0 && by >= 0 && bx < cp->bw && by < cp->bh;
			if (have)
			{
				load_dequant(coef + cp->coef_off + ((size_t)by * cp->bw + bx) * 64, im->qt[ci], v);
				idct_block_regs(v, rows);
			}
			else
			{
#pragma unroll
				for (int i = 0; i < 16; ++i)
					rows[i] = 0;
			}
			/* exchange with the partner lane; even lane (Cb) packs rows 0-3, odd lane (Cr) rows 4-7 */
			const bool is_cr = (item & 1) != 0;
#pragma unroll
			for (int r = 0; r < 8; ++r)
			{
				uint32_t plo = __shfl_xor_sync(0xffffffffu, rows[2 * r], 1);
				uint32_t phi = __shfl_xor_sync(0xffffffffu, rows[2 * r + 1], 1);
				const bool mine = is_cr ? (r >= 4) : (r < 4);
				if (mine && have)
				{
					uint32_t cb_lo = is_cr ? plo : rows[2 * r], cb_hi = is_cr ? phi : rows[2 * r + 1];
					uint32_t cr_lo = is_cr ? rows[2 * r] : plo, cr_hi = is_cr ? rows[2 * r + 1] : phi;
					uint4 a, b;
					a.x = __byte_perm(cb_lo, cr_lo, 0x0400);
					a.y = __byte_perm(cb_lo, cr_lo, 0x0501);
					a.z = __byte_perm(cb_lo, cr_lo, 0x0602);
					a.w = __byte_perm(cb_lo, cr_lo, 0x0703);
					b.x = __byte_perm(cb_hi, cr_hi, 0x0400);
					b.y = __byte_perm(cb_hi, cr_hi, 0x0501);
					b.z = __byte_perm(cb_hi, cr_hi, 0x0602);
					b.w = __byte_perm(cb_hi, cr_hi, 0x0703);
					/* keep only byte 0 and byte 2 of each word: (cb) | (cr << 16) */
					a.x &= 0x00ff00ffu; a.y &= 0x00ff00ffu; a.z &= 0x00ff00ffu; a.w &= 0x00ff00ffu;
					b.x &= 0x00ff00ffu; b.y &= 0x00ff00ffu; b.z &= 0x00ff00ffu; b.w &= 0x00ff00ffu;
					uint4 *dst = reinterpret_cast<uint4 *>(&sC[lby * 8 + r][lbx * 8]);
					dst[0] = a;
					dst[1] = b;
				}
			}
		}
	}
	__syncthreads();

	/* ---- phase 2: upsample + colour + store ---- */
	const jcu_comp *cc = &im->comp[1];
	const int wl = cc->w_lores, cy = cc->cy;
	const int crow0 = ty * 32 - 8, ccol0 = tx * 64 - 8; /* global chroma coords of sC[0][0] */
	const bool vec_ok = (N == 4) ? ((W & 3) == 0) : ((W & 15) == 0);
	uint8_t *obase = out + im->out_off;
	for (int task = tid; task < (F420_TH * F420_TW) / 16; task += FUSED_THREADS)
	{
		const int ly = task >> 3, g = task & 7;
		const int gy = ty * F420_TH + ly, gx0 = tx * F420_TW + g * 16;
		if (gy >= H || gx0 >= W)
			continue;
		const int k = gy >> 1;
		const int nr = min(k, cy - 1) - crow0;
		const int fr = ((gy & 1) ? min(k + 1, cy - 1) : max(k - 1, 0)) - crow0;
		/* t[j] for chroma columns i0-1 .. i0+8, clamped to the real width */
		uint32_t t[10];
		const int i0 = gx0 >> 1;
#pragma unroll
		for (int j = 0; j < 10; ++j)
		{
			int ci = min(max(i0 - 1 + j, 0), wl - 1) - ccol0;
			t[j] = 3u * sC[nr][ci] + sC[fr][ci];
		}
		const uint4 yv = *reinterpret_cast<const uint4 *>(&sY[ly][g * 16]);
		const uint32_t yw[4] = {yv.x, yv.y, yv.z, yv.w};
		uint32_t px[16];
#pragma unroll
		for (int j = 0; j < 8; ++j)
		{
			/* even pixel 2i uses t[i-1], odd pixel 2i+1 uses t[i+1]  (jpeg.c:1832-1833) */
			uint32_t ce = ((3u * t[j + 1] + t[j] + 0x00080008u) >> 4) & 0x00ff00ffu;
			uint32_t co = ((3u * t[j + 1] + t[j + 2] + 0x00080008u) >> 4) & 0x00ff00ffu;
			int y0 = (int)((yw[j >> 1] >> (16 * (j & 1))) & 255u);
			int y1 = (int)((yw[j >> 1] >> (16 * (j & 1) + 8)) & 255u);
			px[2 * j] = rgb_pack(y0, (int)(ce & 255u), (int)(ce >> 16));
			px[2 * j + 1] = rgb_pack(y1, (int)(co & 255u), (int)(co >> 16));
		}
		const int count = min(16, W - gx0);
		store_pixels<N>(obase + ((size_t)gy * W + gx0) * N, px, count, vec_ok);
	}
}

/* ===================================================================================== */
/* 4:4:4  (all 1x1)  YCbCr -> RGB / RGBA                                                  */
/* ===================================================================================== */
template <int N>
__global__ void __launch_bounds__(FUSED_THREADS, 2)
k_fused_444(const int32_t *__restrict__ list, int nlist, const jcu_image *__restrict__ images,
				const int16_t *__restrict__ coef, uint8_t *__restrict__ out)
{
	__shared__ __align__(16) uint8_t sP[3][F444_TH][F444_TW];

	int local;
	const jcu_image *im = fused_locate(list, nlist, images, blockIdx.x, local);
	const int tx = local % im->tiles_x, ty = local / im->tiles_x;
	const int W = im->width, H = im->height;
	const int tid = threadIdx.x;

	if (tid < 192)
	{
		int v[64];
		uint32_t rows[16];
		const int ci = tid >> 6, b = tid & 63;
		const int lbx = b & 15, lby = b >> 4;
		const jcu_comp *cp = &im->comp[ci];
		const int bx = tx * 16 + lbx, by = ty * 4 + lby;
		if (bx < cp->bw && by < cp->bh)
		{
			load_dequant(coef + cp->coef_off + ((size_t)by * cp->bw + bx) * 64, im->qt[ci], v);
			idct_block_regs(v, rows);
#pragma unroll
			for (int r = 0; r < 8; ++r)
				*reinterpret_cast<uint2 *>(&sP[ci][lby * 8 + r][lbx * 8]) = make_uint2(rows[2 * r], rows[2 * r + 1]);
		}
	}
	__syncthreads();

	const bool vec_ok = (N == 4) ? ((W & 3) == 0) : ((W & 15) == 0);
	uint8_t *obase = out + im->out_off;
	for (int task = tid; task < (F444_TH * F444_TW) / 16; task += FUSED_THREADS)
	{
		const int ly = task >> 3, g = task & 7;
		const int gy = ty * F444_TH + ly, gx0 = tx * F444_TW + g * 16;
		if (gy >= H || gx0 >= W)
			continue;
		const uint4 yv = *reinterpret_cast<const uint4 *>(&sP[0][ly][g * 16]);
		const uint4 bv = *reinterpret_cast<const uint4 *>(&sP[1][ly][g * 16]);
		const uint4 rv = *reinterpret_cast<const uint4 *>(&sP[2][ly][g * 16]);
		const uint32_t yw[4] = {yv.x, yv.y, yv.z, yv.w}, bw_[4] = {bv.x, bv.y, bv.z, bv.w}, rw[4] = {rv.x, rv.y, rv.z, rv.w};
		uint32_t px[16];
#pragma unroll
		for (int p = 0; p < 16; ++p)
		{
			const int sh = 8 * (p & 3);
			px[p] = rgb_pack((int)((yw[p >> 2] >> sh) & 255u), (int)((bw_[p >> 2] >> sh) & 255u), (int)((rw[p >> 2] >> sh) & 255u));
		}
		const int count = min(16, W - gx0);
		store_pixels<N>(obase + ((size_t)gy * W + gx0) * N, px, count, vec_ok);
	}
}

/* ===================================================================================== */
/* component 0 only (gray files; YCbCr files asked for 1 or 2 channels) -> 1..4 channels  */
/* ===================================================================================== */
template <int N>
__global__ void __launch_bounds__(FUSED_THREADS, 2)
k_fused_gray(const int32_t *__restrict__ list, int nlist, const jcu_image *__restrict__ images,
				 const int16_t *__restrict__ coef, uint8_t *__restrict__ out)
{
	__shared__ __align__(16) uint8_t sY[FGRAY_TH][FGRAY_TW];

	int local;
	const jcu_image *im = fused_locate(list, nlist, images, blockIdx.x, local);
	const int tx = local % im->tiles_x, ty = local / im->tiles_x;
	const int W = im->width, H = im->height;
	const int tid = threadIdx.x;
	{
		int v[64];
		uint32_t rows[16];
		const jcu_comp *cp = &im->comp[0];
		const int lbx = tid & 31, lby = tid >> 5;
		const int bx = tx * 32 + lbx, by = ty * 8 + lby;
		if (bx < cp->bw && by < cp->bh)
		{
			load_dequant(coef + cp->coef_off + ((size_t)by * cp->bw + bx) * 64, im->qt[0], v);
			idct_block_regs(v, rows);
#pragma unroll
			for (int r = 0; r < 8; ++r)
				*reinterpret_cast<uint2 *>(&sY[lby * 8 + r][lbx * 8]) = make_uint2(rows[2 * r], rows[2 * r + 1]);
		}
	}
	__syncthreads();

	const bool vec_ok = (N == 1 || N == 3) ? ((W & 15) == 0) : (N == 2 ? ((W & 7) == 0) : ((W & 3) == 0));
	uint8_t *obase = out + im->out_off;
	for (int task = tid; task < (FGRAY_TH * FGRAY_TW) / 16; task += FUSED_THREADS)
	{
		const int ly = task >> 4, g = task & 15;
		const int gy = ty * FGRAY_TH + ly, gx0 = tx * FGRAY_TW + g * 16;
		if (gy >= H || gx0 >= W)
			continue;
		const uint4 yv = *reinterpret_cast<const uint4 *>(&sY[ly][g * 16]);
		const uint32_t yw[4] = {yv.x, yv.y, yv.z, yv.w};
		const int count = min(16, W - gx0);
		uint8_t *o = obase + ((size_t)gy * W + gx0) * N;
		if (N >= 3)
		{
			uint32_t px[16];
#pragma unroll
			for (int p = 0; p < 16; ++p)
				px[p] = ((yw[p >> 2] >> (8 * (p & 3))) & 255u) * 0x00010101u;
			store_pixels<(N >= 3 ? N : 3)>(o, px, count, vec_ok);
		}
		else if (N == 1)
		{
			if (vec_ok && count == 16)
				*reinterpret_cast<uint4 *>(o) = yv;
			else
				for (int p = 0; p < count; ++p)
					o[p] = (uint8_t)(yw[p >> 2] >> (8 * (p & 3)));
		}
		else
		{
			if (vec_ok && count == 16)
			{
#pragma unroll
				for (int q = 0; q < 2; ++q)
				{
					uint4 w;
					uint32_t a = yw[2 * q], b = yw[2 * q + 1];
					w.x = (a & 255u) | 0xff00u | ((a & 0xff00u) << 8) | 0xff000000u;
					w.y = ((a >> 16) & 255u) | 0xff00u | ((a >> 24) << 16) | 0xff000000u;
					w.z = (b & 255u) | 0xff00u | ((b & 0xff00u) << 8) | 0xff000000u;
					w.w = ((b >> 16) & 255u) | 0xff00u | ((b >> 24) << 16) | 0xff000000u;
					reinterpret_cast<uint4 *>(o)[q] = w;
				}
			}
			else
				for (int p = 0; p < count; ++p)
				{
					o[2 * p] = (uint8_t)(yw[p >> 2] >> (8 * (p & 3)));
					o[2 * p + 1] = 255;
				}
		}
	}
}

/* host-side dispatch: one launch per path, grid = tiles of all images on that path.
   n_out is uniform per batch (req_comp applies to the whole call) except that gray files
   and colour files can differ; the list is therefore split by the caller per (path, n). */
static inline cudaError_t fused_launch(int path, int tiles, const int32_t *list, int nlist, const jcu_image *images,
													const int16_t *coef, uint8_t *out, cudaStream_t st, int n_out)
{
	if (tiles <= 0)
		return cudaSuccess;
	dim3 grid((unsigned)tiles), block(FUSED_THREADS);
	switch (path)
	{
	case JCU_PATH_FUSED_420:
		if (n_out == 4)
			k_fused_420<4><<<grid, block, 0, st>>>(list, nlist, images, coef, out);
		else
			k_fused_420<3><<<grid, block, 0, st>>>(list, nlist, images, coef, out);
		break;
	case JCU_PATH_FUSED_444:
		if (n_out == 4)
			k_fused_444<4><<<grid, block, 0, st>>>(list, nlist, images, coef, out);
		else
			k_fused_444<3><<<grid, block, 0, st>>>(list, nlist, images, coef, out);
		break;
	default:
		switch (n_out)
		{
		case 1: k_fused_gray<1><<<grid, block, 0, st>>>(list, nlist, images, coef, out); break;
		case 2: k_fused_gray<2><<<grid, block, 0, st>>>(list, nlist, images, coef, out); break;
		case 3: k_fused_gray<3><<<grid, block, 0, st>>>(list, nlist, images, coef, out); break;
		default: k_fused_gray<4><<<grid, block, 0, st>>>(list, nlist, images, coef, out); break;
		}
		break;
	}
	return cudaGetLastError();
}
