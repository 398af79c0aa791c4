/*
 * jpeg_entropy_prog.cuh -- warp-cooperative decode of progressive AC scans WITHOUT restart markers (sm_100a).
 *
 * Replaces stbi__jpeg_decode_block_prog_ac (jpeg.c:406-558) driven by the non-interleaved loop of
 * stbi__parse_entropy_coded_data (jpeg.c:1236-1266) for scans that are one long restart segment.  Such a
 * scan is inherently serial: where block b+1 starts in the bitstream is known only after block b has been
 * parsed, and in a refinement scan the parse itself depends on which coefficients of the block are already
 * non-zero (every one of them consumes a correction bit, jpeg.c:489-497, 533-541), so not even a
 * self-synchronising decoder can start in the middle.  What can be removed is everything else that the
 * one-thread-per-scan decoder (jpeg_entropy.cuh) pays per coefficient:
 *
 *   - ONE WARP PER SCAN.  Lane l owns the coefficients at zig-zag positions l and l+32 of the current block
 *     (two 2-byte loads from one 128-byte line, the next block's issued before the current one is parsed,
 *     so the serial chain never waits for HBM).  Two ballots turn them into the block's 64-bit history mask
 *     M (bit z = coefficient at zig-zag position z is non-zero).
 *   - The parse runs warp-uniform on bit masks, never touching the coefficients: a symbol (run r, size s)
 *     covers the positions k..kz, kz = position of the (r+1)-th zero of M at or after k (jpeg.c:528-551);
 *     its correction bits are the popc(M & range) bits that follow in the stream.  An end-of-band run
 *     covers k..se (jpeg.c:476-497, 513).
 *   - Correction bits are applied by all lanes at once: the coefficient at position z takes stream bit
 *     number popc(M & range & below(z)) -- a parallel bit deposit.
 *   - In a FIRST scan (jpeg.c:412-466) blocks inside an end-of-band run consume no bits, so a run advances
 *     the block counter by its whole length in one step.
 *   - The stream is the unstuffed copy made by k_unstuff (jpeg_entropy_baseline.cuh): the bit reader is a
 *     bit position plus three words in registers over a 64-word shared-memory window that the lanes refill
 *     32 words at a time, one global load per lane, half a window ahead.
 *
 * Results are bit-identical to the reference, including its behaviour on damaged streams: positions past
 * the band map to coefficient 63 through the padded zig-zag table (jpeg.c:303), size != 1 in a refinement
 * scan and undecodable codes fail the image ("bad huffman code"), and zero bits are fed past the end.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"
#include "jpeg_entropy.cuh"
#include "jpeg_entropy_baseline.cuh"

#define PC_THREADS 32
#define PC_RING_WORDS 64

struct PcShared
{
	HuffLut lut;
	uint32_t ring[PC_RING_WORDS];
	uint8_t zz[80];
};

/* bit reader over the unstuffed stream: bp = bit position, (cur, nxt, nx2) = words wi, wi+1, wi+2, wi = bp / 32 */
struct PcReader
{
	uint32_t bp, wi;
	uint32_t cur, nxt, nx2;
	uint32_t pend; /* this lane's word of the chunk being prefetched */
	uint32_t nwords;
	const uint32_t *src;
};

__device__ __forceinline__ uint32_t pc_word(const PcReader &br, uint32_t i) { return i < br.nwords ? __ldg(br.src + i) : 0u; }

__device__ __forceinline__ void pc_init(PcReader &br, uint32_t *ring, const uint32_t *src, uint32_t nwords, int lane)
{
	br.src = src;
	br.nwords = nwords;
	br.bp = 0;
	br.wi = 0;
	ring[lane] = pc_word(br, (uint32_t)lane);
	ring[32 + lane] = pc_word(br, 32u + (uint32_t)lane);
	br.pend = 0;
	__syncwarp();
	br.cur = ring[0];
	br.nxt = ring[1];
	br.nx2 = ring[2];
}
/* next 32 bits, and the 32 after them */
__device__ __forceinline__ uint32_t pc_peek(const PcReader &br) { return __funnelshift_l(br.nxt, br.cur, br.bp & 31u); }
__device__ __forceinline__ uint32_t pc_peek2(const PcReader &br) { return __funnelshift_l(br.nx2, br.nxt, br.bp & 31u); }

/* consume n <= 32 bits (warp-uniform).  Word wi+2 enters the window from the ring; chunk c (32 words) lives in
   ring half c & 1.  When the window starts reading chunk c, chunk c-1 is dead: every lane requests its word of
   chunk c+1 and parks it in the ring eight words later, long before the window gets there. */
__device__ __forceinline__ void pc_skip(PcReader &br, uint32_t *ring, uint32_t n, int lane)
{
	br.bp += n;
	if ((br.bp >> 5) != br.wi)
	{
		++br.wi;
		br.cur = br.nxt;
		br.nxt = br.nx2;
		const uint32_t w2 = br.wi + 2u;
		br.nx2 = ring[w2 & (PC_RING_WORDS - 1)];
		if ((w2 & 31u) == 0u)
			br.pend = pc_word(br, w2 + 32u + (uint32_t)lane);
		else if ((w2 & 31u) == 8u && w2 >= 32u)
		{
			ring[((w2 + 24u) & (PC_RING_WORDS - 1)) + (uint32_t)lane] = br.pend; /* half of chunk (w2 / 32) + 1 */
			__syncwarp();
		}
	}
}
__device__ __forceinline__ uint32_t pc_bits(PcReader &br, uint32_t *ring, uint32_t n, int lane) /* 1 <= n <= 16 */
{
	const uint32_t v = pc_peek(br) >> (32u - n);
	pc_skip(br, ring, n, lane);
	return v;
}

/* one Huffman symbol from the next bits: code length, run, size; -1 = no such code.  When the table entry
   already holds the extended value (LUT_RESOLVED) `len` covers the magnitude bits too and `value` is set. */
__device__ __forceinline__ int pc_symbol(const HuffLut *L, uint32_t top, int &r, int &s, int &value, bool &resolved)
{
	const uint32_t e = L->fast[top >> (32 - ENT_LUT_BITS)];
	resolved = false;
	value = 0;
	if ((e & 31u) == 0u)
	{
		const uint32_t t16 = top >> 16;
		int l = ENT_LUT_BITS + 1;
		while (t16 >= L->limit[l])
			++l;
		if (l > 16)
			return -1;
		const int sym = (int)L->vals[(L->base[l] + (int)(t16 >> (16 - l))) & 255];
		r = sym >> 4;
		s = sym & 15;
		return l;
	}
	r = (int)((e >> 5) & 15u);
	if (e & LUT_RESOLVED)
	{
		resolved = true;
		value = (int)e >> 16;
		s = 1; /* callers that care about the exact size look at `value` */
	}
	else
		s = (int)((e >> 9) & 15u);
	return (int)(e & 31u);
}

__device__ __forceinline__ uint64_t pc_below(int z) { return z >= 64 ? ~0ull : ((1ull << z) - 1ull); } /* bits 0..z-1 */

/*
 * grid = CTA work items (one single-segment AC scan each), 32 threads.
 */
__global__ void __launch_bounds__(PC_THREADS)
k_prog_ac_coop(const EntWork *__restrict__ work, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
					const uint32_t *__restrict__ seg_start, const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ uslab,
					const uint32_t *__restrict__ useg_len, int16_t *__restrict__ coef, int32_t *__restrict__ img_err)
{
	__shared__ PcShared sh;
	const int lane = threadIdx.x;
	const EntWork w = work[blockIdx.x];
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	for (int i = lane; i < 80; i += PC_THREADS)
		sh.zz[i] = c_natural[i];
	if (sc->ac_spec[0] < 0)
	{
		if (lane == 0)
			atomicOr(&img_err[sc->image], 1);
		return;
	}
	hufflut_build(&sh.lut, &huff[sc->ac_spec[0]], false, lane, PC_THREADS);
	const HuffLut *L = &sh.lut;

	const bool refine = sc->kind == JCU_SCAN_PROG_AC_REFINE;
	const int ss = sc->ss, se = sc->se, al = sc->al;
	const int total = sc->total_units, upr = sc->units_per_row;
	const jcu_comp *cp = &im->comp[sc->comp[0]];
	int16_t *plane = coef + cp->coef_off;
	const int bw = cp->bw;
	const int nat0 = sh.zz[lane], nat1 = sh.zz[lane + 32];
	const uint64_t band = pc_below(se + 1) & ~pc_below(ss);
	const int bit = (int)(int16_t)(1 << al);

	PcReader br;
	{
		const uint32_t g = sc->seg_first + w.seg_first;
		const uint32_t *src = reinterpret_cast<const uint32_t *>(uslab + entb_useg_off(im->data_off, seg_start[g], g));
		pc_init(br, sh.ring, src, useg_len[g] >> 2, lane);
	}
	auto block_at = [&](int u) -> int16_t * { return plane + ((size_t)(u / upr) * bw + (u % upr)) * 64; };

	bool failed = false;
	if (!refine)
	{
		/* ---- first pass of a band (jpeg.c:412-466): values go straight to the (zeroed) slab ---- */
		int u = 0;
		while (u < total && !failed)
		{
			int16_t *blk = block_at(u);
			const unsigned key = coef_key(blk);
			int k = ss;
			int eob_run = 0;
			do
			{
				int r, s, value;
				bool resolved;
				const int len = pc_symbol(L, pc_peek(br), r, s, value, resolved);
				if (len < 0)
				{
					failed = true;
					break;
				}
				pc_skip(br, sh.ring, (uint32_t)len, lane);
				if (s == 0)
				{
					if (r < 15)
					{
						eob_run = 1 << r;
						if (r)
							eob_run += (int)pc_bits(br, sh.ring, (uint32_t)r, lane);
						--eob_run;
						break;
					}
					k += 16;
				}
				else
				{
					k += r;
					const int zig = sh.zz[k < 79 ? k : 79];
					++k;
					if (!resolved)
					{
						const int v = (int)pc_bits(br, sh.ring, (uint32_t)s, lane);
						value = v < (1 << (s - 1)) ? v + (int)(0xffffffffu << s) + 1 : v;
					}
					if (lane == 0)
						blk[coef_pos(zig, key)] = (int16_t)(value * (1 << al));
				}
			} while (k <= se);
			/* the blocks of an end-of-band run consume no bits (jpeg.c:416-420) */
			u += 1 + eob_run;
		}
	}
	else
	{
		/* ---- refinement of a band (jpeg.c:468-556) ---- */
		int eob_run = 0;
		int c0n = 0, c1n = 0; /* prefetched coefficients of the next block */
		if (total > 0)
		{
			const int16_t *b0 = block_at(0);
			const unsigned key = coef_key(b0);
			c0n = b0[coef_pos(nat0, key)];
			c1n = b0[coef_pos(nat1, key)];
		}
		for (int u = 0; u < total && !failed; ++u)
		{
			int16_t *blk = block_at(u);
			const unsigned key = coef_key(blk);
			int c0 = c0n, c1 = c1n;
			const int o0 = c0, o1 = c1;
			if (u + 1 < total)
			{
				const int16_t *bn = block_at(u + 1);
				const unsigned kn = coef_key(bn);
				c0n = bn[coef_pos(nat0, kn)];
				c1n = bn[coef_pos(nat1, kn)];
			}
			const uint64_t M = (uint64_t)__ballot_sync(0xffffffffu, c0 != 0) | ((uint64_t)__ballot_sync(0xffffffffu, c1 != 0) << 32);

			/* correction bits of the history coefficients in `range`, then consume them */
			auto correct = [&](uint64_t range) {
				const uint64_t R = M & range;
				const int n = __popcll(R);
				if (n == 0)
					return;
				const uint32_t shi = pc_peek(br), slo = pc_peek2(br);
				if ((R >> lane) & 1ull)
				{
					const int rank = __popcll(R & pc_below(lane));
					const uint32_t b = rank < 32 ? (shi >> (31 - rank)) : (slo >> (63 - rank));
					if ((b & 1u) && (c0 & bit) == 0)
						c0 = c0 > 0 ? c0 + bit : c0 - bit;
				}
				if ((R >> (lane + 32)) & 1ull)
				{
					const int rank = __popcll(R & pc_below(lane + 32));
					const uint32_t b = rank < 32 ? (shi >> (31 - rank)) : (slo >> (63 - rank));
					if ((b & 1u) && (c1 & bit) == 0)
						c1 = c1 > 0 ? c1 + bit : c1 - bit;
				}
				if (n > 32)
				{
					pc_skip(br, sh.ring, 32u, lane);
					pc_skip(br, sh.ring, (uint32_t)(n - 32), lane);
				}
				else
					pc_skip(br, sh.ring, (uint32_t)n, lane);
			};

			if (eob_run)
			{
				--eob_run;
				correct(band);
			}
			else
			{
				int k = ss;
				do
				{
					int r, s, value;
					bool resolved;
					const int len = pc_symbol(L, pc_peek(br), r, s, value, resolved);
					if (len < 0)
					{
						failed = true;
						break;
					}
					pc_skip(br, sh.ring, (uint32_t)len, lane);
					int newval = 0;
					bool to_end = false;
					if (s == 0)
					{
						if (r < 15)
						{
							eob_run = (1 << r) - 1;
							if (r)
								eob_run += (int)pc_bits(br, sh.ring, (uint32_t)r, lane);
							to_end = true;
						}
					}
					else
					{
						if (resolved)
						{
							if (value != 1 && value != -1)
							{
								failed = true;
								break;
							}
							newval = value > 0 ? bit : -bit;
						}
						else
						{
							if (s != 1)
							{
								failed = true;
								break;
							}
							newval = pc_bits(br, sh.ring, 1u, lane) ? bit : -bit;
						}
					}
					/* positions k..kz: kz = the (r+1)-th zero-history position from k, else the end of the band */
					int kz = -1;
					if (!to_end)
					{
						uint64_t t = ~M & band & ~pc_below(k);
						for (int i = 0; i < r && t; ++i)
							t &= t - 1ull;
						if (t)
							kz = __ffsll((long long)t) - 1;
					}
					const int kend = kz >= 0 ? kz : se;
					correct(pc_below(kend + 1) & ~pc_below(k) & band);
					if (kz >= 0)
					{
						/* jpeg.c:545: the new coefficient (a zero for ZRL) lands on the zero it stopped at */
						if (kz == lane)
							c0 = newval;
						if (kz == lane + 32)
							c1 = newval;
					}
					k = kend + 1;
				} while (k <= se);
			}
			if (c0 != o0)
				blk[coef_pos(nat0, key)] = (int16_t)c0;
			if (c1 != o1)
				blk[coef_pos(nat1, key)] = (int16_t)c1;
		}
	}
	if (failed && lane == 0)
		atomicOr(&img_err[sc->image], 1);
}
