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
#define F420_NCH (F420_CBW * F420_CBH)		/* chroma block positions per component */
#define F420_SLOTS (128 + 2 * F420_NCH)	/* staged coefficient blocks per tile */
#define F420_SMEM_COEF (F420_SLOTS * 128)
#define F420_SMEM_Y (F420_TH * F420_TW)
#define F420_SMEM_C (F420_CH * F420_CW * 4)
#define F420_SMEM_BYTES (F420_SMEM_COEF + F420_SMEM_Y + F420_SMEM_C + 3 * 128)

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc)
{
	unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

/* saturate two s32 to u8 and pack below the low half of c: result bytes = [sat(lo), sat(hi), c.b0, c.b1] */
__device__ __forceinline__ uint32_t pack_sat_u8(int hi, int lo, uint32_t c)
{
	uint32_t d;
	asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(lo), "r"(c));
	return d;
}
/* four s32 -> four saturated bytes [b0, b1, b2, b3] */
__device__ __forceinline__ uint32_t pack4_sat_u8(int b0, int b1, int b2, int b3)
{
	return pack_sat_u8(b1, b0, pack_sat_u8(b3, b2, 0u));
}

/* 64 coefficients of one staged block (swizzled 16-byte chunks) -> sign-extended ints; DEQ applies
   (short)(c*q) from the tile's quant table in shared memory (progressive; jpeg.c:1323) */
template <bool DEQ>
__device__ __forceinline__ void load_block_smem(const uint4 *__restrict__ slot, int swz, const uint16_t *__restrict__ q,
																int (&v)[64])
{
#pragma unroll
	for (int j = 0; j < 8; ++j)
	{
		const uint4 w = slot[j ^ swz];
		const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
		for (int i = 0; i < 4; ++i)
		{
			int c0 = (int)(int16_t)(ws[i] & 0xffffu);
			int c1 = ((int)ws[i]) >> 16;
			if (DEQ)
			{
				c0 = (int)(int16_t)(c0 * (int)q[j * 8 + 2 * i]);
				c1 = (int)(int16_t)(c1 * (int)q[j * 8 + 2 * i + 1]);
			}
			v[j * 8 + 2 * i] = c0;
			v[j * 8 + 2 * i + 1] = c1;
		}
	}
}

/* IDCT of 64 ints in v -> 8 rows of 8 saturated samples, two words per row */
__device__ __forceinline__ void idct_block_packed(int (&v)[64], uint32_t (&rows)[16])
{
	int o[8];
#pragma unroll
	for (int c = 0; c < 8; ++c)
	{
		idct_1d(v[c], v[c + 8], v[c + 16], v[c + 24], v[c + 32], v[c + 40], v[c + 48], v[c + 56], 512, o);
#pragma unroll
		for (int r = 0; r < 8; ++r)
			v[c + 8 * r] = o[r] >> 10;
	}
#pragma unroll
	for (int r = 0; r < 8; ++r)
	{
		idct_1d(v[8 * r], v[8 * r + 1], v[8 * r + 2], v[8 * r + 3], v[8 * r + 4], v[8 * r + 5], v[8 * r + 6],
				  v[8 * r + 7], 65536 + (128 << 17), o);
		rows[2 * r] = pack4_sat_u8(o[0] >> 17, o[1] >> 17, o[2] >> 17, o[3] >> 17);
		rows[2 * r + 1] = pack4_sat_u8(o[4] >> 17, o[5] >> 17, o[6] >> 17, o[7] >> 17);
	}
}

/*
 * grid: (max tiles per image, images on this path).  DEQ = coefficients are raw (progressive);
 * otherwise the entropy stage already stored (short)(c*q) as the reference does for baseline
 * (jpeg.c:325,345,365).
 */
template <int N, bool DEQ>
__global__ void __launch_bounds__(FUSED_THREADS, 2)
k_fused_420(const int32_t *__restrict__ list, const jcu_image *__restrict__ images,
				const int16_t *__restrict__ coef, uint8_t *__restrict__ out)
{
	extern __shared__ __align__(16) unsigned char f420_smem[];
	uint4 *sCoef = reinterpret_cast<uint4 *>(f420_smem);																		  /* [F420_SLOTS * 8] */
	uint8_t(*sY)[F420_TW] = reinterpret_cast<uint8_t(*)[F420_TW]>(f420_smem + F420_SMEM_COEF);				  /* [F420_TH] */
	uint32_t(*sC)[F420_CW] = reinterpret_cast<uint32_t(*)[F420_CW]>(f420_smem + F420_SMEM_COEF + F420_SMEM_Y); /* 16*(Cb|Cr<<16) */
	uint16_t(*sQ)[64] = reinterpret_cast<uint16_t(*)[64]>(f420_smem + F420_SMEM_COEF + F420_SMEM_Y + F420_SMEM_C);

	const jcu_image *im = &images[list[blockIdx.y]];
	const int tiles_x = im->tiles_x;
	if ((int)blockIdx.x >= tiles_x * im->tiles_y)
		return;
	const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
	const int W = im->width, H = im->height;
	const int tid = threadIdx.x;
	const int ybw = im->comp[0].bw, ybh = im->comp[0].bh, cbw = im->comp[1].bw, cbh = im->comp[1].bh;

	/* ---- phase 0: stage the tile's coefficient blocks (own + chroma halo ring) ---- */
	{
		const int16_t *base0 = coef + im->comp[0].coef_off;
		const int16_t *base1 = coef + im->comp[1].coef_off;
		const int16_t *base2 = coef + im->comp[2].coef_off;
#pragma unroll
		for (int it = 0; it < (F420_SLOTS * 8 + FUSED_THREADS - 1) / FUSED_THREADS; ++it)
		{
			const int id = tid + it * FUSED_THREADS;
			const int slot = id >> 3, j = id & 7;
			if (slot < F420_SLOTS)
			{
				const int16_t *src = nullptr;
				if (slot < 128)
				{
					const int bx = tx * 16 + (slot & 15), by = ty * 8 + (slot >> 4);
					if (bx < ybw && by < ybh)
						src = base0 + ((size_t)by * ybw + bx) * 64;
				}
				else
				{
					const int idx = slot - 128;
					const int ci = idx >= F420_NCH, pos = idx - ci * F420_NCH;
					const int bx = tx * 8 - 1 + pos % F420_CBW, by = ty * 4 - 1 + pos / F420_CBW;
					if (bx >= 0 && by >= 0 && bx < cbw && by < cbh)
						src = (ci ? base2 : base1) + ((size_t)by * cbw + bx) * 64;
				}
				if (src)
					cp_async16(&sCoef[slot * 8 + (j ^ (slot & 7))], src + ((j ^ (int)coef_key(src)) * 8));
			}
		}
		if (DEQ)
			for (int i = tid; i < 192; i += FUSED_THREADS)
				sQ[i >> 6][i & 63] = im->qt_out[i >> 6][i & 63];
		cp_async_wait_all();
	}
	__syncthreads();

	/* ---- phase 1: one thread = one block: dequant, column pass, row pass ---- */
	{
		int v[64];
		uint32_t rows[16];
		if (tid < 128)
		{
			const int lbx = tid & 15, lby = tid >> 4;
			if (tx * 16 + lbx < ybw && ty * 8 + lby < ybh)
			{
				load_block_smem<DEQ>(&sCoef[tid * 8], tid & 7, sQ[0], v);
				idct_block_packed(v, rows);
#pragma unroll
				for (int r = 0; r < 8; ++r)
					*reinterpret_cast<uint2 *>(&sY[lby * 8 + r][lbx * 8]) = make_uint2(rows[2 * r], rows[2 * r + 1]);
			}
		}
		else
		{
			/* lanes (2p, 2p+1) = (Cb, Cr) of chroma position p; they swap rows by shuffle and store
			   16*(Cb | Cr<<16) so the upsampling filter handles both components in one register */
			const int item = tid - 128;
			const int ci = item & 1, pos = item >> 1;
			const int slot = 128 + ci * F420_NCH + pos;
			const int lbx = pos % F420_CBW, lby = pos / F420_CBW;
			const int bx = tx * 8 - 1 + lbx, by = ty * 4 - 1 + lby;
			const bool have = pos < F420_NCH && bx >= 0 && by >= 0 && bx < cbw && by < cbh;
			if (have)
			{
				load_block_smem<DEQ>(&sCoef[slot * 8], slot & 7, sQ[1 + ci], v);
				idct_block_packed(v, rows);
			}
			else
			{
#pragma unroll
				for (int i = 0; i < 16; ++i)
					rows[i] = 0;
			}
			const bool is_cr = ci != 0;
#pragma unroll
			for (int r = 0; r < 8; ++r)
			{
				const uint32_t plo = __shfl_xor_sync(0xffffffffu, rows[2 * r], 1);
				const uint32_t phi = __shfl_xor_sync(0xffffffffu, rows[2 * r + 1], 1);
				const bool mine = is_cr ? (r >= 4) : (r < 4);
				if (mine && have)
				{
					const uint32_t cb_lo = is_cr ? plo : rows[2 * r], cb_hi = is_cr ? phi : rows[2 * r + 1];
					const uint32_t cr_lo = is_cr ? rows[2 * r] : plo, cr_hi = is_cr ? rows[2 * r + 1] : phi;
					uint4 a, b;
					a.x = (__byte_perm(cb_lo, cr_lo, 0x0400) & 0x00ff00ffu) << 4;
					a.y = (__byte_perm(cb_lo, cr_lo, 0x0501) & 0x00ff00ffu) << 4;
					a.z = (__byte_perm(cb_lo, cr_lo, 0x0602) & 0x00ff00ffu) << 4;
					a.w = (__byte_perm(cb_lo, cr_lo, 0x0703) & 0x00ff00ffu) << 4;
					b.x = (__byte_perm(cb_hi, cr_hi, 0x0400) & 0x00ff00ffu) << 4;
					b.y = (__byte_perm(cb_hi, cr_hi, 0x0501) & 0x00ff00ffu) << 4;
					b.z = (__byte_perm(cb_hi, cr_hi, 0x0602) & 0x00ff00ffu) << 4;
					b.w = (__byte_perm(cb_hi, cr_hi, 0x0703) & 0x00ff00ffu) << 4;
					uint4 *dst = reinterpret_cast<uint4 *>(&sC[lby * 8 + r][lbx * 8]);
					dst[0] = a;
					dst[1] = b;
				}
			}
		}
	}
	__syncthreads();

	/* ---- clamp-to-edge on the REAL chroma width (w_lores, jpeg.c:2276): replicate one column
	   past each image edge so phase 2 needs no per-sample clamps ---- */
	const int wl = im->comp[1].w_lores, cy = im->comp[1].cy;
	const int crow0 = ty * 32 - 8, ccol0 = tx * 64 - 8; /* global chroma coords of sC[0][0] */
	{
		const int last = wl - 1 - ccol0; /* local column of the last real chroma sample */
		const bool left = tx == 0, right = last + 1 < F420_CW && last >= 0 && last < 8 + 64;
		if (left || right)
		{
			if (tid < F420_CH)
			{
				if (left)
					sC[tid][7] = sC[tid][8];
				if (right)
					sC[tid][last + 1] = sC[tid][last];
			}
			__syncthreads();
		}
	}

	/* ---- phase 2: one task = 16 pixels of one row ---- */
	const bool vec_ok = (N == 4) ? ((W & 3) == 0) : ((W & 15) == 0);
	uint8_t *obase = out + im->out_off;
#pragma unroll 1
	for (int task = tid; task < (F420_TH * F420_TW) / 16; task += FUSED_THREADS)
	{
		const int ly = task >> 3, g = task & 7;
		const int gy = ty * F420_TH + ly, gx0 = tx * F420_TW + g * 16;
		if (gy >= H || gx0 >= W)
			continue;
		const int k = gy >> 1;
		const int nr = min(k, cy - 1) - crow0;
		const int fr = ((gy & 1) ? min(k + 1, cy - 1) : max(k - 1, 0)) - crow0;
		/* t[j] = 16 * (3*near + far) for chroma columns i0-1 .. i0+8 (both components packed) */
		uint32_t t[10];
		{
			const uint32_t *np = &sC[nr][7 + g * 8], *fp = &sC[fr][7 + g * 8];
			const uint4 n0 = *reinterpret_cast<const uint4 *>(np + 1), n1 = *reinterpret_cast<const uint4 *>(np + 5);
			const uint4 f0 = *reinterpret_cast<const uint4 *>(fp + 1), f1 = *reinterpret_cast<const uint4 *>(fp + 5);
			t[0] = 3u * np[0] + fp[0];
			t[1] = 3u * n0.x + f0.x;
			t[2] = 3u * n0.y + f0.y;
			t[3] = 3u * n0.z + f0.z;
			t[4] = 3u * n0.w + f0.w;
			t[5] = 3u * n1.x + f1.x;
			t[6] = 3u * n1.y + f1.y;
			t[7] = 3u * n1.z + f1.z;
			t[8] = 3u * n1.w + f1.w;
			t[9] = 3u * np[9] + fp[9];
		}
		const uint4 yv = *reinterpret_cast<const uint4 *>(&sY[ly][g * 16]);
		const uint32_t yw[4] = {yv.x, yv.y, yv.z, yv.w};
		int R[16], G[16], B[16];
#pragma unroll
		for (int p = 0; p < 16; ++p)
		{
			const int j = p >> 1;
			/* (3*t[i] + t[i-1 or i+1] + 8) >> 4 (jpeg.c:1832-1833), here scaled by 16 so that the two
			   results sit in bytes 1 and 3; xor 0x80 turns them into signed (c - 128) */
			const uint32_t c16 = (3u * t[j + 1] + ((p & 1) ? t[j + 2] : t[j]) + 0x00800080u) ^ 0x80008000u;
			const int cb = (int)(int8_t)(c16 >> 8);
			const int cr = ((int)c16) >> 24;
			const int y = (int)((yw[p >> 2] >> (8 * (p & 3))) & 255u);
			/* stbi__YCbCr_to_RGB_row (jpeg.c:1981-1990) with the common factor 2^8 divided out */
			const int yf = y * 4096 + 2048;
			R[p] = (yf + cr * 5743) >> 12;
			G[p] = (yf + cr * -2925 + ((cb * -1410) & ~255)) >> 12;
			B[p] = (yf + cb * 7258) >> 12;
		}
		const int count = min(16, W - gx0);
		uint8_t *o = obase + ((size_t)gy * W + gx0) * N;
		if (vec_ok && count == 16)
		{
			if (N == 3)
			{
#pragma unroll
				for (int q = 0; q < 4; ++q)
				{
					const int a = 4 * q;
					uint32_t *ow = reinterpret_cast<uint32_t *>(o) + 3 * q;
					ow[0] = pack4_sat_u8(R[a], G[a], B[a], R[a + 1]);
					ow[1] = pack4_sat_u8(G[a + 1], B[a + 1], R[a + 2], G[a + 2]);
					ow[2] = pack4_sat_u8(B[a + 2], R[a + 3], G[a + 3], B[a + 3]);
				}
			}
			else
			{
#pragma unroll
				for (int q = 0; q < 4; ++q)
					reinterpret_cast<uint4 *>(o)[q] =
						 make_uint4(pack4_sat_u8(R[4 * q], G[4 * q], B[4 * q], 255), pack4_sat_u8(R[4 * q + 1], G[4 * q + 1], B[4 * q + 1], 255),
										pack4_sat_u8(R[4 * q + 2], G[4 * q + 2], B[4 * q + 2], 255),
										pack4_sat_u8(R[4 * q + 3], G[4 * q + 3], B[4 * q + 3], 255));
			}
		}
		else
		{
#pragma unroll
			for (int p = 0; p < 16; ++p)
				if (p < count)
				{
					o[N * p] = (uint8_t)clamp_u8(R[p]);
					o[N * p + 1] = (uint8_t)clamp_u8(G[p]);
					o[N * p + 2] = (uint8_t)clamp_u8(B[p]);
					if (N == 4)
						o[N * p + 3] = 255;
				}
		}
	}
}

/* ===================================================================================== */
/* 4:4:4  (all 1x1)  YCbCr -> RGB / RGBA                                                  */
/* ===================================================================================== */
template <int N>
__global__ void __launch_bounds__(FUSED_THREADS, 2)
k_fused_444(const int32_t *__restrict__ list, const jcu_image *__restrict__ images,
				const int16_t *__restrict__ coef, uint8_t *__restrict__ out)
{
	__shared__ __align__(16) uint8_t sP[3][F444_TH][F444_TW];

	const jcu_image *im = &images[list[blockIdx.y]];
	if ((int)blockIdx.x >= im->tiles_x * im->tiles_y)
		return;
	const int tx = blockIdx.x % im->tiles_x, ty = blockIdx.x / im->tiles_x;
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
			load_dequant(coef + cp->coef_off + ((size_t)by * cp->bw + bx) * 64, im->qt_out[ci], v);
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
k_fused_gray(const int32_t *__restrict__ list, const jcu_image *__restrict__ images,
				 const int16_t *__restrict__ coef, uint8_t *__restrict__ out)
{
	__shared__ __align__(16) uint8_t sY[FGRAY_TH][FGRAY_TW];

	const jcu_image *im = &images[list[blockIdx.y]];
	if ((int)blockIdx.x >= im->tiles_x * im->tiles_y)
		return;
	const int tx = blockIdx.x % im->tiles_x, ty = blockIdx.x / im->tiles_x;
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
			load_dequant(coef + cp->coef_off + ((size_t)by * cp->bw + bx) * 64, im->qt_out[0], v);
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

/* host-side dispatch: one launch per path; grid = (max tiles per image, images on the path).
   n_out is uniform per path within one call (req_comp applies to the whole call). */
static inline cudaError_t fused_launch(int path, int max_tiles, const int32_t *list, int nlist, const jcu_image *images,
													const int16_t *coef, uint8_t *out, cudaStream_t st, int n_out, int deq)
{
	if (max_tiles <= 0 || nlist <= 0)
		return cudaSuccess;
	dim3 grid((unsigned)max_tiles, (unsigned)nlist), block(FUSED_THREADS);
	switch (path)
	{
	case JCU_PATH_FUSED_420:
	{
		void (*k)(const int32_t *, const jcu_image *, const int16_t *, uint8_t *) =
			 n_out == 4 ? (deq ? k_fused_420<4, true> : k_fused_420<4, false>) : (deq ? k_fused_420<3, true> : k_fused_420<3, false>);
		cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, F420_SMEM_BYTES);
		if (e != cudaSuccess)
			return e;
		k<<<grid, block, F420_SMEM_BYTES, st>>>(list, images, coef, out);
		break;
	}
	case JCU_PATH_FUSED_444:
		if (n_out == 4)
			k_fused_444<4><<<grid, block, 0, st>>>(list, images, coef, out);
		else
			k_fused_444<3><<<grid, block, 0, st>>>(list, images, coef, out);
		break;
	default:
		switch (n_out)
		{
		case 1: k_fused_gray<1><<<grid, block, 0, st>>>(list, images, coef, out); break;
		case 2: k_fused_gray<2><<<grid, block, 0, st>>>(list, images, coef, out); break;
		case 3: k_fused_gray<3><<<grid, block, 0, st>>>(list, images, coef, out); break;
		default: k_fused_gray<4><<<grid, block, 0, st>>>(list, images, coef, out); break;
		}
		break;
	}
	return cudaGetLastError();
}
