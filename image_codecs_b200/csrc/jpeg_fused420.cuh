/*
 * jpeg_fused420.cuh -- fused output stage for 4:2:0 YCbCr (sm_100a), strip-persistent edition.
 *
 * One kernel does dequantise -> integer 8x8 IDCT (stbi__idct_block, jpeg.c:615) -> h2v2 chroma
 * upsample (stbi__resample_row_hv_2, jpeg.c:1816, row selection jpeg.c:2301-2318) -> YCbCr->RGB
 * (stbi__YCbCr_to_RGB_row, jpeg.c:1976) -> packed RGB/RGBA stores, touching HBM only to read each
 * coefficient block once and write each pixel once.
 *
 * Work decomposition.  A CTA owns a vertical STRIP of one image (<= S420_MS MCUs wide) inside a
 * BAND of MCU rows and walks down it one MCU row per iteration:
 *
 *   TMA    one thread arms an mbarrier and issues <= 4 bulk copies (cp.async.bulk, SASS UBLKCP): the
 *          two luma block rows of MCU row r and the Cb / Cr block rows of MCU row r+1 (plus one
 *          halo block each side).  Block rows are contiguous in the slab, so a row is ONE copy.
 *   IDCT   one thread = one 8x8 block, all 64 coefficients in registers (column pass, row pass);
 *          luma of row r goes to the sY sample plane, chroma of row r+1 to a 2-slot ring.  Cb and
 *          Cr of one position sit in adjacent lanes and swap half their rows by warp shuffle, so
 *          the ring holds ONE word per chroma sample: 16*Cb+8 | (16*(Cr-128)+8) << 16.
 *   pixel  one task = 8 pixels of one row: the h2v2 filter runs on both chroma fields of the word at
 *          once (3*near+far, then 3*t[i]+t[i+-1]; the +8 per sample becomes the filter's rounding
 *          term and the x16 puts the result in bytes 1 and 3), colour, saturating pack, store.
 *
 * The h2v2 filter reaches one chroma row above and below an MCU row (jpeg.c:11).  Walking DOWN the
 * image removes that halo from the hot loop: chroma of row r+1 is computed one iteration early
 * (its first row is the lower neighbour of row r), and the last row of row r-1 is kept in a line
 * buffer.  Only the band's first row pays a prologue (two chroma-only iterations) and each strip
 * pays one halo block column per side; nothing is spilled to HBM and no block is re-read from HBM
 * except those halo columns (L2 hits).  Image edges clamp on the REAL component size (w_lores, cy --
 * never the MCU padding) by replicating one column in the ring.
 *
 * No tensor cores: the IDCT must reproduce the reference's fixed-point rounding (north_star).
 * Algorithmic HBM bytes: 128 B x coefficient blocks + n B x pixels.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"
#include "jpeg_output.cuh"

#ifndef S420_MIN_CTAS
#define S420_MIN_CTAS (640 / S420_THREADS) /* resident CTAs per SM asked of the register allocator: 5 x 128 threads = 96 registers, no spills, 45 KB of shared memory each */
#endif
#ifndef S420_THREADS
#define S420_THREADS 128
#endif
#define S420_MS (5 * S420_THREADS / 32)									  /* MCUs per strip (640 px) */
#define S420_MC (S420_MS + 2)						  /* chroma block positions incl. one halo each side */
#define S420_SLOTS (4 * S420_MS + 2 * S420_MC) /* staged coefficient blocks per iteration (244) */
#define S420_YW (S420_MS * 16)					  /* luma sample row, bytes */
#define S420_CW (8 * S420_MC + 4)								  /* chroma ring row stride in words (>= 8 * S420_MC, == 20 mod 32) */
#define S420_OFF_Y (S420_SLOTS * 128)
#define S420_OFF_C (S420_OFF_Y + 2 * 16 * S420_YW)			/* two luma planes: IDCT of row r+1 may start while row r is still read */
#define S420_OFF_P (S420_OFF_C + 3 * 8 * S420_CW * 4) /* three chroma slots: MCU rows r, r+1 (bottom halo), r+2 (being written) */
#define S420_OFF_Q (S420_OFF_P + 2 * S420_CW * 4)		/* two line buffers: last chroma row of MCU row r-1 (top halo) */
#define S420_OFF_BAR(DEQ) (S420_OFF_Q + ((DEQ) ? 3 * 64 * 2 : 0)) /* the quantiser tables are staged for progressive files only */
#define S420_SMEM_BYTES(DEQ) (S420_OFF_BAR(DEQ) + 16)

/* ---- mbarrier + bulk async copy (TMA, non-tensor form) ---- */
__device__ __forceinline__ uint32_t s420_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void s420_mbar_init(uint64_t *bar, unsigned count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s420_smem_u32(bar)), "r"(count) : "memory");
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void s420_mbar_expect_tx(uint64_t *bar, unsigned bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s420_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void s420_mbar_wait(uint64_t *bar, unsigned parity)
{
	asm volatile(
		 "{\n"
		 ".reg .pred p;\n"
		 "WAIT_%=:\n"
		 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		 "@p bra DONE_%=;\n"
		 "bra WAIT_%=;\n"
		 "DONE_%=:\n"
		 "}\n" ::"r"(s420_smem_u32(bar)),
		 "r"(parity)
		 : "memory");
}
/* global -> shared bulk copy, completion counted in bytes on the mbarrier; size multiple of 16 */
__device__ __forceinline__ void s420_bulk_g2s(void *dst_smem, const void *src_gmem, unsigned bytes, uint64_t *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s420_smem_u32(dst_smem)),
					 "l"(src_gmem), "r"(bytes), "r"(s420_smem_u32(bar))
					 : "memory");
}

/* saturate two s32 to u8 and pack below the low half of c: result bytes = [sat(lo), sat(hi), c.b0, c.b1] */
__device__ __forceinline__ uint32_t s420_pack2(int hi, int lo, uint32_t c)
{
	uint32_t d;
	asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(lo), "r"(c));
	return d;
}
__device__ __forceinline__ uint32_t s420_pack4(int b0, int b1, int b2, int b3) { return s420_pack2(b1, b0, s420_pack2(b3, b2, 0u)); }

/* 64 coefficients of one staged block -> sign-extended ints.  `key` undoes the slab's chunk
   permutation (the copy in shared memory is byte-identical to HBM); DEQ applies (short)(c*q) for
   progressive files (stbi__jpeg_dequantize, jpeg.c:1323) */
template <bool DEQ>
__device__ __forceinline__ void s420_load_block(const uint4 *__restrict__ slot, unsigned key, const uint16_t *__restrict__ q,
																int (&v)[64])
{
#pragma unroll
	for (int j = 0; j < 8; ++j)
	{
		const uint4 w = slot[j ^ key];
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

/* column pass + row pass (jpeg.c:615-679) on registers -> 8 rows of 8 saturated samples, two words per row */
#ifndef S420_FMA_COL
#define S420_FMA_COL true /* measured: 3.541 -> 3.524 ms per 1024 x 1080p */
#endif
#ifndef S420_FMA_ROW
#define S420_FMA_ROW false
#endif
#ifndef S420_COL_MODE
#define S420_COL_MODE 0 /* idct_1d MODE of the column pass */
#endif
#ifndef S420_ROW_MODE
#define S420_ROW_MODE 0
#endif
__device__ __forceinline__ void s420_idct(int (&v)[64], uint32_t (&rows)[16])
{
	int o[8];
	int(&u)[64] = v; /* a second register array for the column pass's results compiles to the same code */
#pragma unroll
	for (int c = 0; c < 8; ++c)
	{
		idct_1d<S420_FMA_COL, S420_COL_MODE>(v[c], v[c + 8], v[c + 16], v[c + 24], v[c + 32], v[c + 40], v[c + 48], v[c + 56], 512, o);
#pragma unroll
		for (int r = 0; r < 8; ++r)
			u[c + 8 * r] = o[r] >> 10;
	}
#pragma unroll
	for (int r = 0; r < 8; ++r)
	{
		idct_1d<S420_FMA_ROW, S420_ROW_MODE>(u[8 * r], u[8 * r + 1], u[8 * r + 2], u[8 * r + 3], u[8 * r + 4], u[8 * r + 5], u[8 * r + 6], u[8 * r + 7],
				  65536 + (128 << 17), o);
		rows[2 * r] = s420_pack4(o[0] >> 17, o[1] >> 17, o[2] >> 17, o[3] >> 17);
		rows[2 * r + 1] = s420_pack4(o[4] >> 17, o[5] >> 17, o[6] >> 17, o[7] >> 17);
	}
}

/* four chroma samples (bytes of cbw / crw) -> four ring words 16*Cb+8 | (16*(Cr-128)+8) << 16.
   Two byte permutes interleave the even / odd samples of both planes, one more per sample spreads a (Cb, Cr) pair
   over the two 16-bit fields with zeros above (the zero register is the second source), one multiply-add scales and
   biases both fields: 2.5 instructions per sample. */
__device__ __forceinline__ uint32_t s420_spread(uint32_t pairs, uint32_t sel)
{
	uint32_t d;
	asm("prmt.b32 %0, %1, 0, %2;" : "=r"(d) : "r"(pairs), "r"(sel));
	return d;
}
__device__ __forceinline__ uint4 s420_chroma_words(uint32_t cbw, uint32_t crw)
{
	const uint32_t K = 0xF8080008u; /* 8 | (8 - 16*128) << 16 */
	const uint32_t ev = __byte_perm(cbw, crw, 0x6240); /* Cb0 Cr0 Cb2 Cr2 */
	const uint32_t od = __byte_perm(cbw, crw, 0x7351); /* Cb1 Cr1 Cb3 Cr3 */
	uint4 r;
	r.x = s420_spread(ev, 0x4140u) * 16u + K;
	r.y = s420_spread(od, 0x4140u) * 16u + K;
	r.z = s420_spread(ev, 0x4342u) * 16u + K;
	r.w = s420_spread(od, 0x4342u) * 16u + K;
	return r;
}

/* ---- pinned instruction selection for the pixel phase: NVVM otherwise re-associates the colour sums and
 *      re-adds the rounding constant once per channel ---- */
__device__ __forceinline__ int s420_mad(int a, int b, int c)
{
	int d;
	asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
	return d;
}
template <int SEL>
__device__ __forceinline__ int s420_byte(uint32_t w) /* PRMT: selected byte -> bits 0-7, zeros above */
{
	int d;
	asm("prmt.b32 %0, %1, 0, %2;" : "=r"(d) : "r"(w), "n"(SEL));
	return d;
}

/* one pixel: chroma word c16 (Cb in byte 1, Cr-128 signed in byte 3), luma byte -> r, g, b (not yet clamped).
   stbi__YCbCr_to_RGB_row (jpeg.c:1981-1990) with the common factor 2^8 divided out; Cb's -128 folded into constants */
template <int YSEL>
__device__ __forceinline__ void s420_colour(uint32_t c16, uint32_t yw, int &r, int &g, int &b)
{
	const int cr = ((int)c16) >> 24;
	const int cb = s420_byte<0x4441>(c16);
	const int y = s420_byte<0x4440 + YSEL>(yw);
	/* y + (t >> 12) is one LEA.HI.SX32; 2048 is a multiple of 256, so it moves inside the masked term */
	r = y + (s420_mad(cr, 5743, 2048) >> 12);
	b = y + (s420_mad(cb, 7258, 2048 - 128 * 7258) >> 12);
	const int m = s420_mad(cb, -1410, 128 * 1410 + 2048) & ~255;
	g = y + (s420_mad(cr, -2925, m) >> 12);
}

/*
 * grid: (max strips*bands per image, images on this path).  DEQ = coefficients are raw (progressive);
 * otherwise the entropy stage already stored (short)(c*q) as the reference does for baseline.
 */
template <int N, bool DEQ>
__global__ void __launch_bounds__(S420_THREADS, S420_MIN_CTAS)
k_fused420_strip(const int32_t *__restrict__ list, const jcu_image *__restrict__ images, const int16_t *__restrict__ coef,
					  uint8_t *__restrict__ out, int rows_per_band)
{
	extern __shared__ __align__(128) unsigned char s420_smem[];
	uint4 *sCoef = reinterpret_cast<uint4 *>(s420_smem);
	uint8_t *sY = s420_smem + S420_OFF_Y;											  /* [2][16][S420_YW] */
	uint32_t *sC = reinterpret_cast<uint32_t *>(s420_smem + S420_OFF_C);	  /* [3][8][S420_CW]: chroma of MCU row c lives in slot c % 3 */
	uint32_t *sP = reinterpret_cast<uint32_t *>(s420_smem + S420_OFF_P);	  /* [2][S420_CW], directly behind sC */
	uint16_t(*sQ)[64] = reinterpret_cast<uint16_t(*)[64]>(s420_smem + S420_OFF_Q);
	uint64_t *bar = reinterpret_cast<uint64_t *>(s420_smem + S420_OFF_BAR(DEQ));
	__shared__ int4 sRow[2][16]; /* per pixel row of MCU row r (buffer r & 1): word offsets (from sC) of its near / far chroma
											 rows, byte offset of its luma row in sY, byte offset of its output row from the image's first */

	const jcu_image *im = &images[list[blockIdx.y]];
	const int strips = im->tiles_x;
	if ((int)blockIdx.x >= strips * im->tiles_y)
		return;
	const int strip = blockIdx.x % strips, band = blockIdx.x / strips;
	const int tid = threadIdx.x;
	const int W = im->width, H = im->height;
	const int mcu_y = im->mcu_y;
	const int ybw = im->comp[0].bw, cbw = im->comp[1].bw;
	const int wl = im->comp[1].w_lores, cy = im->comp[1].cy;
	const int m0 = strip * S420_MS, Ms = min(S420_MS, im->mcu_x - m0), Mc = Ms + 2;
	const int r0 = band * rows_per_band, r1 = min(r0 + rows_per_band, mcu_y);
	const int c0 = max(m0 - 1, 0), c1 = min(m0 + Ms + 1, cbw);
	const int rc_last = min(r1, mcu_y - 1); /* last chroma MCU row this band touches */
	const int16_t *baseY = coef + im->comp[0].coef_off;
	const int16_t *baseCb = coef + im->comp[1].coef_off;
	const int16_t *baseCr = coef + im->comp[2].coef_off;

	if (tid == 0)
		s420_mbar_init(bar, 1);
	if (DEQ)
		for (int i = tid; i < 192; i += S420_THREADS)
			sQ[i >> 6][i & 63] = im->qt_out[i >> 6][i & 63];
	__syncthreads();

	/* thread 0: arm the barrier and issue the bulk copies of iteration r */
	auto issue = [&](int r) {
		const bool doY = r >= r0, doC = (r + 1) <= rc_last;
		const unsigned ybytes = (unsigned)(2 * Ms * 128), cbytes = (unsigned)((c1 - c0) * 128);
		s420_mbar_expect_tx(bar, (doY ? 2u * ybytes : 0u) + (doC ? 2u * cbytes : 0u));
		unsigned char *dst = reinterpret_cast<unsigned char *>(sCoef);
		if (doY)
		{
			const int16_t *s0 = baseY + ((size_t)(2 * r) * ybw + 2 * m0) * 64;
			s420_bulk_g2s(dst, s0, ybytes, bar);
			s420_bulk_g2s(dst + ybytes, s0 + (size_t)ybw * 64, ybytes, bar);
		}
		if (doC)
		{
			const size_t off = ((size_t)(r + 1) * cbw + c0) * 64;
			const unsigned d = (unsigned)((4 * Ms + (c0 - (m0 - 1))) * 128);
			s420_bulk_g2s(dst + d, baseCb + off, cbytes, bar);
			s420_bulk_g2s(dst + d + Mc * 128, baseCr + off, cbytes, bar);
		}
	};
	const int r_start = max(r0 - 2, -1);
	if (tid == 0)
		issue(r_start);
	unsigned parity = 0;

	/* this thread's IDCT task, the same every iteration */
	const int nY = 4 * Ms;
	const bool isY = tid < nY;
	const int ct = tid - nY, cpos = ct >> 1, ci = ct & 1;
	const int cbx = m0 - 1 + cpos;
	const bool cvalid = !isY && ct < 2 * Mc && cbx >= 0 && cbx < cbw;
	const int slot = isY ? tid : (nY + ci * Mc + cpos);
	const int slot_safe = min(slot, S420_SLOTS - 1); /* lanes beyond the strip's blocks read a valid slot and discard the result */
	const int yrow = (isY && tid >= 2 * Ms) ? 1 : 0, lbx = tid - yrow * 2 * Ms;
	const bool chroma_warp = (tid | 31) >= nY; /* warp-uniform: this warp holds chroma lanes */

	const bool vec_ok = (N == 4) ? ((W & 3) == 0) : ((W & 7) == 0);
	const bool full_w = vec_ok && (m0 + Ms) * 16 <= W; /* CTA-uniform: every 8-pixel group of the strip is whole */
	/* vertical flip on load (stbi__vertical_flip convert.c:37) is a store epilogue: the row pitch
	   turns negative and the base moves to the last row */
	const int pitch = im->flip ? -W * N : W * N;
	uint8_t *obase = out + im->out_off + (im->flip ? (size_t)(H - 1) * (size_t)(W * N) : 0);
	const int G = 2 * Ms; /* 8-pixel groups per row of the strip */
	const int ly_first = tid / G, g_first = tid - ly_first * G;
	const int dly = S420_THREADS / G, dg = S420_THREADS - dly * G;

#pragma unroll 1
	for (int r = r_start; r < r1; ++r)
	{
		const bool doY = r >= r0, doC = (r + 1) <= rc_last;
		s420_mbar_wait(bar, parity);
		parity ^= 1u;

		/* chroma slots of MCU rows r-1, r, r+1 (r >= -1) */
		const int sl_prev = (r + 2) % 3, sl_cur = (r + 3) % 3, sl_next = (r + 4) % 3;
		/* The slot of row r-1 is the one the IDCT of iteration r+1 overwrites (with row r+2) while slower warps may
		   still be in the pixel phase of row r: its last row, the top halo of row r, moves to a line buffer first.
		   Row r-1 has been complete since the barrier of iteration r-2. */
		if (doY && r > 0)
			for (int c = tid; c < S420_CW; c += S420_THREADS)
				sP[(r & 1) * S420_CW + c] = sC[(sl_prev * 8 + 7) * S420_CW + c];
		/* row selection of the h2v2 resampler (jpeg.c:2301-2318) for the 16 pixel rows of MCU row r */
		if (tid < 16 && doY)
		{
			const int ly = tid, lk = ly >> 1, k = r * 8 + lk;
			const int nr = (sl_cur * 8 + lk) * S420_CW;
			int fr;
			if (ly & 1)
				fr = (k + 1 > cy - 1) ? nr : (lk < 7 ? nr + S420_CW : (sl_next * 8) * S420_CW);
			else
				fr = (lk > 0) ? nr - S420_CW : (r == 0 ? nr : (24 + (r & 1)) * S420_CW); /* line buffer sP follows sC */
			sRow[r & 1][ly] = make_int4(nr + 8, (r * 16 + ly < H) ? fr + 8 : -1, ((r & 1) * 16 + ly) * S420_YW, (r * 16 + ly) * pitch);
		}
		/* ---- IDCT phase: luma of MCU row r, chroma of MCU row r+1 ---- */
		{
			int v[64];
			uint32_t rows[16];
			/* warp-uniform: a warp none of whose lanes has a block this iteration (luma warps in the chroma-only
			   prologue) skips the phase; inside, lanes without a block transform whatever their slot holds and store
			   nothing, so no value has to be merged from a second path */
			if (!__any_sync(0xffffffffu, (isY && doY) || (cvalid && doC)))
				goto idct_done;
#ifdef S420_SKIP_IDCT /* timing experiment only */
			if (r != 12345)
				goto idct_done;
#endif
			{
				const int16_t *g = isY ? baseY + ((size_t)(2 * r + yrow) * ybw + 2 * m0 + lbx) * 64
											  : (ci ? baseCr : baseCb) + ((size_t)(r + 1) * cbw + cbx) * 64;
				s420_load_block<DEQ>(&sCoef[slot_safe * 8], coef_key(g), sQ[isY ? 0 : 1 + ci], v);
				s420_idct(v, rows);
			}
			if (isY && doY)
			{
				uint8_t *d = sY + ((r & 1) * 16 + yrow * 8) * S420_YW + lbx * 8;
#pragma unroll
				for (int rr = 0; rr < 8; ++rr)
					*reinterpret_cast<uint2 *>(d + rr * S420_YW) = make_uint2(rows[2 * rr], rows[2 * rr + 1]);
			}
			if (chroma_warp)
			{
				/* lanes (2p, 2p+1) = (Cb, Cr) of one position: the Cb lane keeps rows 0-3, the Cr lane rows 4-7 */
				const bool store = cvalid && doC;
				uint32_t *dbase = sC + (sl_next * 8 + (ci ? 4 : 0)) * S420_CW + cpos * 8;
#pragma unroll
				for (int rr = 0; rr < 4; ++rr)
				{
					const uint32_t send_lo = ci ? rows[2 * rr] : rows[2 * (rr + 4)];
					const uint32_t send_hi = ci ? rows[2 * rr + 1] : rows[2 * (rr + 4) + 1];
					const uint32_t got_lo = __shfl_xor_sync(0xffffffffu, send_lo, 1);
					const uint32_t got_hi = __shfl_xor_sync(0xffffffffu, send_hi, 1);
					const uint32_t cb_lo = ci ? got_lo : rows[2 * rr], cb_hi = ci ? got_hi : rows[2 * rr + 1];
					const uint32_t cr_lo = ci ? rows[2 * (rr + 4)] : got_lo, cr_hi = ci ? rows[2 * (rr + 4) + 1] : got_hi;
					if (store)
					{
						uint4 *d = reinterpret_cast<uint4 *>(dbase + rr * S420_CW);
						d[0] = s420_chroma_words(cb_lo, cr_lo);
						d[1] = s420_chroma_words(cb_hi, cr_hi);
					}
				}
				/* clamp-to-edge on the real chroma width (w_lores, jpeg.c:2276): replicate one column past
				   each image edge; the same thread wrote the source sample, so no barrier is needed */
				if (store && (cbx == 0 || cbx == cbw - 1))
				{
					const int e = wl - cbx * 8; /* samples of the last block that are real: 1..8 */
#pragma unroll
					for (int rr = 0; rr < 4; ++rr)
					{
						uint32_t *d = dbase + rr * S420_CW;
						if (cbx == 0)
							d[-1] = d[0];
						if (cbx == cbw - 1)
							d[e] = d[e - 1];
					}
				}
			}
		}
	idct_done:
		__syncthreads();
		if (tid == 0 && r + 1 < r1)
			issue(r + 1);

		/* ---- pixel phase: MCU row r, one task = 8 pixels of one row ---- */
#ifdef S420_SKIP_PIXEL /* timing experiment only */
		if (doY && r == 12345)
#else
		if (doY)
#endif
		{
			/* everything that depends on the row alone comes out of the row table with one 16-byte load; the column
			   terms are three multiples of g */
			uint8_t *ostrip = obase + m0 * 16 * N;
			const int4 *rowtab = sRow[r & 1];
			int ly = ly_first, g = g_first;
#pragma unroll 1
			while (ly < 16)
			{
				const int4 rw = rowtab[ly]; /* x, y: word offsets of the near / far chroma rows (+ 8: first column of the strip); y < 0: row beyond the image */
				const int gx0 = m0 * 16 + g * 8;
				if (rw.y >= 0 && (full_w || gx0 < W))
				{
					const uint32_t *np = sC + rw.x + g * 4, *fp = sC + rw.y + g * 4;
					uint32_t t[6];
					{
						const uint4 n4 = *reinterpret_cast<const uint4 *>(np), f4 = *reinterpret_cast<const uint4 *>(fp);
						t[0] = 3u * np[-1] + fp[-1];
						t[1] = 3u * n4.x + f4.x;
						t[2] = 3u * n4.y + f4.y;
						t[3] = 3u * n4.z + f4.z;
						t[4] = 3u * n4.w + f4.w;
						t[5] = 3u * np[4] + fp[4];
					}
					const uint2 yv = *reinterpret_cast<const uint2 *>(sY + rw.z + g * 8);
					int R[8], Gc[8], B[8];
					/* (3*t[i] + t[i-1 or i+1] + 8) >> 4 (jpeg.c:1832-1833) on both chroma fields at once */
					s420_colour<0>(3u * t[1] + t[0], yv.x, R[0], Gc[0], B[0]);
					s420_colour<1>(3u * t[1] + t[2], yv.x, R[1], Gc[1], B[1]);
					s420_colour<2>(3u * t[2] + t[1], yv.x, R[2], Gc[2], B[2]);
					s420_colour<3>(3u * t[2] + t[3], yv.x, R[3], Gc[3], B[3]);
					s420_colour<0>(3u * t[3] + t[2], yv.y, R[4], Gc[4], B[4]);
					s420_colour<1>(3u * t[3] + t[4], yv.y, R[5], Gc[5], B[5]);
					s420_colour<2>(3u * t[4] + t[3], yv.y, R[6], Gc[6], B[6]);
					s420_colour<3>(3u * t[4] + t[5], yv.y, R[7], Gc[7], B[7]);
					uint8_t *o = ostrip + (rw.w + g * 8 * N);
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
							uint2 *ow = reinterpret_cast<uint2 *>(o); /* 24 bytes, 8-byte aligned when W % 8 == 0 */
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
		/* ONE barrier per iteration.  The luma plane, the row table and the top-halo line are double buffered and the
		   chroma ring has three slots, so a warp that is done with the pixels of row r goes straight on to the IDCT of
		   row r+1 (it writes plane (r+1)&1 and chroma slot (r+2)%3, neither of which the pixel phase of row r reads)
		   and only waits, at the barrier above, for everybody's IDCT of row r+1 -- which also proves that nobody still
		   reads row r-1. */
	}
}

/* host-side launch: grid = (max strips*bands per image, images on the path) */
static inline cudaError_t fused420_launch(int max_tiles, const int32_t *list, int nlist, const jcu_image *images, const int16_t *coef,
														uint8_t *out, cudaStream_t st, int n_out, int deq, int rows_per_band)
{
	if (max_tiles <= 0 || nlist <= 0)
		return cudaSuccess;
	void (*k)(const int32_t *, const jcu_image *, const int16_t *, uint8_t *, int) =
		 n_out == 4 ? (deq ? k_fused420_strip<4, true> : k_fused420_strip<4, false>)
						: (deq ? k_fused420_strip<3, true> : k_fused420_strip<3, false>);
	const int smem = deq ? S420_SMEM_BYTES(1) : S420_SMEM_BYTES(0);
	cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
	if (e != cudaSuccess)
		return e;
	k<<<dim3((unsigned)max_tiles, (unsigned)nlist), S420_THREADS, smem, st>>>(list, images, coef, out, rows_per_band);
	return cudaGetLastError();
}
