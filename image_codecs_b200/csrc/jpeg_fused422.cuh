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
 * with one halo block per side), one block per thread IDCT into a luma byte plane and a plane of packed (Cb, Cr)
 * words, then 8-pixel tasks.  h_2 has no vertical reach, so there is no row halo.
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
#define F2_MIN_CTAS 5
#define F2_MS 30					/* MCUs per strip: 60 luma + 2 x 32 chroma blocks = 124 IDCT tasks */
#define F2_MC (F2_MS + 2)		/* chroma block positions incl. one halo each side */
#define F2_ROWS_PER_BAND 16	/* MCU rows a CTA walks */
#define F2_SLOTS (2 * F2_MS + 2 * F2_MC)
#define F2_YW (F2_MS * 16)
#define F2_CW (F2_MC * 8)
#define F2_OFF_Y (F2_SLOTS * 128)
#define F2_OFF_C (F2_OFF_Y + 8 * F2_YW)		 /* [8][F2_CW] words: one per chroma sample, both planes */
#define F2_OFF_Q (F2_OFF_C + 8 * F2_CW * 4)
#define F2_OFF_BAR (F2_OFF_Q + 3 * 64 * 2)
#define F2_SMEM_BYTES (F2_OFF_BAR + 16)

/*
 * Chroma lives in shared memory as ONE word per sample position:  64*Cb + 32 | (64*(Cr-128) + 32) << 16.
 * The resampler's  (3 in[i] + in[i+-1] + 2) >> 2  is then a single multiply-add on both planes at once,
 * 3 * w[i] + w[i+-1]: the two +32 add up to the rounding term 2 * 64, the Cb field stays below 2^16
 * (64 * 1020 + 128), and the result carries Cb in byte 1 and Cr - 128 (signed) in byte 3 -- exactly what
 * s420_colour() of the 4:2:0 kernel takes.  (The first version kept two byte planes and filtered each with scalar
 * code: 56.5 instead of 47 instructions per pixel, 117 registers, local-memory traffic for the out[2w-2] patch.)
 */
__device__ __forceinline__ uint4 f2_chroma_words(uint32_t cbw, uint32_t crw)
{
	const uint32_t K = 0xE0200020u; /* 32 | (32 - 64 * 128) << 16 */
	const uint32_t ev = __byte_perm(cbw, crw, 0x6240); /* Cb0 Cr0 Cb2 Cr2 */
	const uint32_t od = __byte_perm(cbw, crw, 0x7351); /* Cb1 Cr1 Cb3 Cr3 */
	uint4 r;
	r.x = s420_spread(ev, 0x4140u) * 64u + K;
	r.y = s420_spread(od, 0x4140u) * 64u + K;
	r.z = s420_spread(ev, 0x4342u) * 64u + K;
	r.w = s420_spread(od, 0x4342u) * 64u + K;
	return r;
}

template <int N, bool DEQ>
__global__ void __launch_bounds__(F2_THREADS, F2_MIN_CTAS)
k_fused422_strip(const int32_t *__restrict__ list, const jcu_image *__restrict__ images, const int16_t *__restrict__ coef,
					  uint8_t *__restrict__ out)
{
	extern __shared__ __align__(128) unsigned char f2_smem[];
	uint4 *sCoef = reinterpret_cast<uint4 *>(f2_smem);
	uint8_t *sY = f2_smem + F2_OFF_Y;												 /* [8][F2_YW] */
	uint32_t *sC = reinterpret_cast<uint32_t *>(f2_smem + F2_OFF_C);		 /* [8][F2_CW]: sample i of the strip's first MCU column sits at 8 + i */
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

	/* this thread's IDCT task, the same every iteration: 2 * F2_MS luma blocks, then (Cb, Cr) of one chroma position in
	   neighbouring lanes (2 * F2_MS is even, so a pair never straddles two warps) */
	const int nY = 2 * F2_MS;
	const bool isY = tid < nY;
	const int ct = tid - nY, cpos = ct >> 1, ci = ct & 1;
	const int cbx = m0 - 1 + cpos;
	const bool yvalid = tid < 2 * Ms;
	const bool cvalid = !isY && cpos < F2_MC && cbx >= c0 && cbx < c1;
	const int slot = isY ? tid : nY + ci * F2_MC + min(cpos, F2_MC - 1);
	const bool chroma_warp = (tid | 31) >= nY; /* warp-uniform: this warp holds chroma lanes */

	const bool vec_ok = (N == 4) ? ((W & 3) == 0) : ((W & 7) == 0);
	const bool full_w = vec_ok && (m0 + Ms) * 16 <= W; /* CTA-uniform: every 8-pixel group of the strip is whole */
	const int G = 2 * Ms;										/* 8-pixel groups per row of the strip */
	const int pitch = im->flip ? -W * N : W * N;
	uint8_t *obase = out + im->out_off + (im->flip ? (size_t)(H - 1) * (size_t)(W * N) : 0);
	uint8_t *ostrip = obase + m0 * 16 * N;
	const int xfix = wl >= 2 ? 2 * wl - 2 : -1; /* the one output the general formulas do not give (jpeg.c:1805) */
	const int qfix = 8 + (wl - 2) - m0 * 8;		  /* plane position of in[w-2]; it may lie in the left halo block */
	const int ly_first = tid / G, g_first = tid - ly_first * G;
	const int dly = F2_THREADS / G, dg = F2_THREADS - dly * G;

#pragma unroll 1
	for (int by = by0; by < by1; ++by)
	{
		s420_mbar_wait(bar, parity);
		parity ^= 1u;
		/* ---- IDCT phase.  Lanes without a block transform whatever their slot holds and store nothing, so that the
		   chroma pairs can exchange rows with warp shuffles ---- */
		if (__any_sync(0xffffffffu, yvalid || cvalid))
		{
			int v[64];
			uint32_t rows[16];
			const int16_t *g = isY ? baseY + ((size_t)by * ybw + 2 * m0 + tid) * 64 : (ci ? baseCr : baseCb) + ((size_t)by * cbw + cbx) * 64;
			s420_load_block<DEQ>(&sCoef[slot * 8], coef_key(g), sQ[isY ? 0 : 1 + ci], v);
			s420_idct(v, rows);
			if (yvalid)
			{
				uint8_t *d = sY + tid * 8;
#pragma unroll
				for (int rr = 0; rr < 8; ++rr)
					*reinterpret_cast<uint2 *>(d + rr * F2_YW) = make_uint2(rows[2 * rr], rows[2 * rr + 1]);
			}
			if (chroma_warp)
			{
				/* lanes (2p, 2p+1) = (Cb, Cr) of one position: the Cb lane keeps rows 0-3, the Cr lane rows 4-7 */
				uint32_t *dbase = sC + (ci ? 4 : 0) * F2_CW + cpos * 8;
#pragma unroll
				for (int rr = 0; rr < 4; ++rr)
				{
					const uint32_t send_lo = ci ? rows[2 * rr] : rows[2 * (rr + 4)];
					const uint32_t send_hi = ci ? rows[2 * rr + 1] : rows[2 * (rr + 4) + 1];
					const uint32_t got_lo = __shfl_xor_sync(0xffffffffu, send_lo, 1);
					const uint32_t got_hi = __shfl_xor_sync(0xffffffffu, send_hi, 1);
					const uint32_t cb_lo = ci ? got_lo : rows[2 * rr], cb_hi = ci ? got_hi : rows[2 * rr + 1];
					const uint32_t cr_lo = ci ? rows[2 * (rr + 4)] : got_lo, cr_hi = ci ? rows[2 * (rr + 4) + 1] : got_hi;
					if (cvalid)
					{
						uint4 *d = reinterpret_cast<uint4 *>(dbase + rr * F2_CW);
						d[0] = f2_chroma_words(cb_lo, cr_lo);
						d[1] = f2_chroma_words(cb_hi, cr_hi);
					}
				}
				/* replicate the column before the first and behind the last REAL chroma sample of the image row (w_lores,
				   jpeg.c:2276): the same thread wrote the source sample.  A chroma block spans one MCU, so the last real
				   sample always lies in the row's last block (1 <= e <= 8); as a right-halo block its column 8 lies outside
				   the plane row, and nobody reads it */
				if (cvalid && (cbx == 0 || cbx == cbw - 1))
				{
					const int e = wl - cbx * 8;
#pragma unroll
					for (int rr = 0; rr < 4; ++rr)
					{
						uint32_t *d = dbase + rr * F2_CW;
						if (cbx == 0)
							d[-1] = d[0];
						if (cbx == cbw - 1 && cpos * 8 + e < F2_CW)
							d[e] = d[e - 1];
					}
				}
			}
		}
		__syncthreads();
		if (tid == 0 && by + 1 < by1)
			issue(by + 1);
		/* ---- pixel phase: 8 rows x G groups of 8 pixels; (ly, g) stepped without a division ---- */
		{
			int ly = ly_first, g = g_first;
#pragma unroll 1
			while (ly < 8)
			{
				const int gy = by * 8 + ly, gx0 = m0 * 16 + g * 8;
				if (gy < H && (full_w || gx0 < W))
				{
					const uint32_t *np = sC + ly * F2_CW + 8 + g * 4; /* first chroma sample of the group; 16-byte aligned */
					const uint4 n4 = *reinterpret_cast<const uint4 *>(np);
					const uint32_t wp = np[-1], wn = np[4];
					uint32_t c[8];
					c[0] = 3u * n4.x + wp;
					c[1] = 3u * n4.x + n4.y;
					c[2] = 3u * n4.y + n4.x;
					c[3] = 3u * n4.y + n4.z;
					c[4] = 3u * n4.z + n4.y;
					c[5] = 3u * n4.z + n4.w;
					c[6] = 3u * n4.w + n4.z;
					c[7] = 3u * n4.w + wn;
					const unsigned patch = (unsigned)(xfix - gx0);
					if (patch < 8u)
					{
						/* out[2w-2] = (3 in[w-2] + in[w-1] + 2) >> 2: an even output of this group */
						const uint32_t *rw = sC + ly * F2_CW + qfix;
						const uint32_t fix = 3u * rw[0] + rw[1];
#pragma unroll
						for (int p = 0; p < 8; p += 2)
							if (patch == (unsigned)p)
								c[p] = fix;
					}
					const uint2 yv = *reinterpret_cast<const uint2 *>(sY + ly * F2_YW + g * 8);
					int R[8], Gc[8], B[8];
					s420_colour<0>(c[0], yv.x, R[0], Gc[0], B[0]);
					s420_colour<1>(c[1], yv.x, R[1], Gc[1], B[1]);
					s420_colour<2>(c[2], yv.x, R[2], Gc[2], B[2]);
					s420_colour<3>(c[3], yv.x, R[3], Gc[3], B[3]);
					s420_colour<0>(c[4], yv.y, R[4], Gc[4], B[4]);
					s420_colour<1>(c[5], yv.y, R[5], Gc[5], B[5]);
					s420_colour<2>(c[6], yv.y, R[6], Gc[6], B[6]);
					s420_colour<3>(c[7], yv.y, R[7], Gc[7], B[7]);
					uint8_t *o = ostrip + ((long long)gy * pitch + g * 8 * N);
					if (full_w || (vec_ok && gx0 + 8 <= W))
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
						const int count = min(8, W - gx0);
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
				g += dg;
				ly += dly;
				if (g >= G)
				{
					g -= G;
					++ly;
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
