/*
 * jpeg_selfsync.cuh -- self-synchronising speculative Huffman decode for baseline scans WITHOUT restart
 * markers (sm_100a).  BASELINE.json north_star / config 3.
 *
 * The reference decodes such a scan strictly serially (stbi__parse_entropy_coded_data, jpeg.c:1198-1232:
 * one bit buffer, one set of DC predictors for the whole image).  Here the scan's entropy-coded bytes are cut
 * into fixed SS_SUB_BYTES subsequences and one thread decodes each, exploiting the fact that a Huffman
 * decoder started at a wrong bit position falls into step with the true symbol boundaries after a few
 * symbols (and, because EOB resets the zig-zag index and the block pattern of an MCU is short, with the
 * true block structure soon after):
 *
 *   U unstuff  the scan's bytes with the stuffing resolved (FF 00 -> FF), written by all threads in parallel into the
 *              unstuffed slab: per 256-byte raw chunk a count, a prefix sum, a second pass that stores.  Every later
 *              phase reads bits with a stateless reader (bit position -> two words -> funnel shift).
 *   1 guess    thread i starts at bit 2048 i of the unstuffed stream assuming "start of an MCU" and records its EXIT
 *              state (bit position, block-in-MCU, zig-zag index) at the first symbol start past its boundary.
 *   2 refine   every thread re-decodes its subsequence from its predecessor's exit state (Jacobi rounds,
 *              double-buffered); subsequence 0 starts from the true state, so correctness spreads at least
 *              one subsequence per round and, thanks to self-synchronisation, in practice converges in a few
 *              rounds; a thread whose entry state did not change skips the work.  Rounds stop when no exit
 *              state changed -- all inside ONE cooperative launch (grid barrier per round), no host round trip.
 *   3 scan     exclusive prefix sum of the number of blocks each subsequence completes; the one block per
 *              subsequence boundary that two threads share is cleared (k_ss_zero_shared).
 *   4 write    every thread decodes its subsequence once more from its (now true) entry state and writes
 *              (short)(value * dequant) coefficients (jpeg.c:345,365): blocks it starts leave as whole 128-byte lines,
 *              the shared ones coefficient by coefficient into the cleared block; DC symbols are written as raw
 *              differences.
 *   5 dc       per component: prefix sum of the DC differences in decode order = the reference's running
 *              dc_pred (jpeg.c:323-325), stored as (short)(dc * dequant[0]).
 *
 * An exit state is (bit position, zig-zag index, block-in-MCU): a function of where the decoder stands, not of how it
 * got there, so states of decoders with different histories compare equal once they have fallen into step.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"
#include "jpeg_entropy_baseline.cuh"

#define SS_THREADS 128
#define SS_SLOT_STRIDE 136 /* bytes between the block-assembly slots of neighbouring threads (write pass) */
#ifndef SS_MIN_SUB
#define SS_MIN_SUB 64u /* shortest decoding subsequence (bytes); measured on ONE 512 x 512 file: 16 / 32 / 64 / 128 / 256 -> 0.69 / 0.64 / 0.60 / 0.77 / 0.95 ms */
#endif
#define SS_SUB_BYTES 256 /* raw bytes per unstuffing thread; also the largest decoding subsequence (SsBuffers.sub_bytes) */

/* entry / exit state of a subsequence: everything a decoder needs to resume, and nothing that depends on its history */
struct SsState
{
	uint32_t bp;  /* bit position in the unstuffed stream */
	uint32_t aux; /* zig-zag index k (bits 0-7) | block-in-MCU (bits 8-19) */
};
__device__ __forceinline__ bool ss_same(const SsState &a, const SsState &b) { return a.bp == b.bp && a.aux == b.aux; }
/* exit states cross CTA boundaries INSIDE one launch of k_ss_refine_loop: read them from L2, never from an L1 line that
   an earlier round left behind */
__device__ __forceinline__ SsState ss_load_l2(const SsState *p)
{
	const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(p));
	SsState s;
	s.bp = v.x;
	s.aux = v.y;
	return s;
}

/* one CTA work item: up to SS_THREADS consecutive subsequences of one scan */
struct SsWork
{
	int32_t scan;
	uint32_t first;	  /* index within the scan of this CTA's first subsequence */
	uint32_t count;	  /* subsequences handled by this CTA */
	uint32_t scan_base; /* global index of the scan's subsequence 0 */
	uint32_t scan_nsub; /* subsequences in the scan (by RAW bytes: an upper bound of the unstuffed ones) */
	uint32_t pad;
};

struct SsBuffers
{
	SsState *exit_a, *exit_b; /* ping-pong exit states */
	SsState *entry_used;		  /* entry state the current exit state was computed from */
	uint32_t *nblk;			  /* blocks completed inside each subsequence (phase U: bytes each raw chunk keeps) */
	uint32_t *first_blk;		  /* exclusive prefix of nblk within the scan (phase U: output offset of each raw chunk) */
	int32_t *changed;			  /* [3] rotating flags: round t sets changed[t % 3] when an exit state changed */
	uint32_t *sync;			  /* [0] arrivals at the grid barrier of k_ss_refine_loop, [1] its generation */
	int32_t *result;			  /* [0] ping-pong side that holds the final exit states, [1] rounds run */
	uint32_t *ticket;			  /* [3] rotating work-item counters of k_ss_refine_loop: round t hands out items from ticket[t % 3] */
	uint32_t sub_bytes;		  /* unstuffed bytes per decoding subsequence: 32 .. 256, chosen per batch (jpeg_cuda.cu) */
};

/* write pass: one entry per block of an MCU (k_ss_decode<SS_WRITE>) */
struct __align__(16) SsBlockInfo
{
	long long base;	 /* element offset of the block at unit (0, 0): component plane + position inside the unit */
	int32_t sx, sy;	 /* elements per unit column / unit row */
	uint32_t dc_tab, ac_tab, zq; /* shared-memory addresses: lookup tables, zig-zag/dequantiser table */
	uint32_t pad;
};

/* per self-sync scan */
struct SsScan
{
	int32_t scan;
	uint32_t base; /* global index of its subsequence 0 */
	uint32_t nsub;
	uint32_t pad;
};

/* ------------------------------------------------------------------------------------------------------
 * Phase U: byte stuffing out of the way, in parallel.  Inside [scan start, data_end) of a scan without restart
 * markers every run of 0xFF bytes is followed by 0x00 (anything else would be the marker that ends the scan, and
 * the host parser put data_end there), and stbi__grow_buffer_unsafe (jpeg.c:167-187) turns the whole run + the
 * zero into ONE data byte 0xFF.  So a raw byte survives unless it is a 0x00 behind an 0xFF or an 0xFF in front of
 * another 0xFF -- a rule that only looks one byte to each side.  Thread i counts the survivors of raw bytes
 * [256 i, 256 i + 256), k_ss_scan turns the counts into offsets, and the second pass stores them as big-endian words
 * (byte j of the stream at j ^ 3) in the segment's area of the unstuffed slab (entb_useg_off), zero padded like
 * k_unstuff does, with useg_len in the same format.
 * ------------------------------------------------------------------------------------------------------ */
/* raw bytes [p0, p0 + 16) of the file as four little-endian words; p0 is 16-byte aligned in the (16-byte aligned, padded)
   bitstream slab, so the load never leaves it */
__device__ __forceinline__ uint4 ssu_load16(const uint8_t *__restrict__ base, uint32_t p0)
{
	return __ldg(reinterpret_cast<const uint4 *>(base + p0));
}

/* Output side of the write pass: surviving bytes are collected big-endian in a 64-bit accumulator and leave as whole
   32-bit words (byte j of the stream sits at j ^ 3, i.e. a word is stored as its big-endian value) -- into a
   shared-memory image of the CTA's part of the output, which is contiguous (chunk i + 1 continues where chunk i
   stopped) and goes to global memory as coalesced lines afterwards; per-thread global stores, 256 bytes apart from
   lane to lane, left the kernel at 28 % issue.  The first word of a chunk may have been begun by the previous chunk
   and the last one may be finished by the next: those bytes go one by one. */
struct SsuOut
{
	uint32_t sm;	/* shared-memory address of the word with index jw_base */
	uint32_t jw_base;
	uint64_t acc;	/* pending bytes, first one in the top byte */
	uint32_t nacc; /* 0..3 pending (at the start: bytes of the first word that belong to the previous chunk) */
	uint32_t jw;	/* index of the word being filled */
	uint32_t h0;	/* > 0 until the first word has left: its bytes below h0 are not ours */
};
__device__ __forceinline__ void ssu_sts8(const SsuOut &o, uint32_t jbyte, uint32_t v)
{
	asm volatile("st.shared.u8 [%0], %1;" ::"r"(o.sm + ((jbyte ^ 3u) - 4u * o.jw_base)), "r"(v) : "memory");
}
__device__ __forceinline__ void ssu_put(SsuOut &o, uint32_t x_top, uint32_t k) /* k = 1..4 bytes, left-aligned in x_top */
{
	o.acc |= (uint64_t)x_top << (32u - 8u * o.nacc);
	o.nacc += k;
	if (o.nacc >= 4u)
	{
		const uint32_t word = (uint32_t)(o.acc >> 32);
		if (o.h0)
		{
			for (uint32_t b = o.h0; b < 4u; ++b)
				ssu_sts8(o, 4u * o.jw + b, (word >> (24u - 8u * b)) & 255u);
			o.h0 = 0u;
		}
		else
			asm volatile("st.shared.u32 [%0], %1;" ::"r"(o.sm + 4u * (o.jw - o.jw_base)), "r"(word) : "memory");
		++o.jw;
		o.acc <<= 32;
		o.nacc -= 4u;
	}
}

template <bool WRITE>
__global__ void __launch_bounds__(SS_THREADS)
k_ss_unstuff(const SsWork *__restrict__ work, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
				 const uint32_t *__restrict__ seg_start, const uint8_t *__restrict__ bitstream, uint8_t *__restrict__ uslab,
				 uint32_t *__restrict__ useg_len, SsBuffers sb)
{
	/* WRITE: the CTA's part of the output (at most SS_THREADS x 256 bytes + the word it shares with its predecessor) */
	__shared__ uint32_t s_out[WRITE ? SS_THREADS * (SS_SUB_BYTES / 4) + 4 : 1];
	__shared__ uint32_t s_jend;
	const SsWork w = work[blockIdx.x];
	const bool has_work = threadIdx.x < w.count;
	if (!WRITE && !has_work)
		return;
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	const uint32_t i = w.first + (has_work ? threadIdx.x : 0u), gi = w.scan_base + i;
	const uint8_t *__restrict__ base = bitstream + im->data_off; /* 16-byte aligned */
	const uint32_t scan_start = seg_start[sc->seg_first], end = sc->data_end;
	/* chunks are cut on 256-byte boundaries of the FILE (the first one is shorter), so every chunk is 16 aligned loads */
	const uint32_t a0 = scan_start & ~(SS_SUB_BYTES - 1u);
	const uint32_t lo = max(scan_start, min(end, a0 + i * SS_SUB_BYTES));
	const uint32_t hi = has_work ? min(end, a0 + (i + 1u) * SS_SUB_BYTES) : lo; /* threads without a chunk only help to copy */
	const uint32_t j0 = WRITE ? sb.first_blk[gi] : 0u;
	const uint32_t j_cta = WRITE ? sb.first_blk[w.scan_base + w.first] : 0u; /* first output byte of this CTA */
	uint32_t j = j0; /* output bytes so far (count pass: survivors of this chunk) */
	uint8_t *const dst = WRITE ? uslab + entb_useg_off(im->data_off, scan_start, (uint32_t)sc->seg_first) : nullptr;
	SsuOut o;
	o.sm = (uint32_t)__cvta_generic_to_shared(s_out);
	o.jw_base = j_cta >> 2;
	o.acc = 0ull;
	o.nacc = j0 & 3u;
	o.jw = j0 >> 2;
	o.h0 = o.nacc;
	if (lo < hi)
	{
		uint32_t prev = lo > scan_start ? (uint32_t)__ldg(base + lo - 1u) : 0u;
		for (uint32_t p0 = lo & ~15u; p0 < hi; p0 += 16u)
		{
			const uint4 v = ssu_load16(base, p0);
			/* the byte behind this unit decides whether its last 0xFF is a fill byte; behind the scan it reads as 0 */
			const uint32_t after = p0 + 16u < end ? (uint32_t)__ldg(base + p0 + 16u) : 0u;
			const uint32_t ws[5] = {v.x, v.y, v.z, v.w, after};
#pragma unroll
			for (int q4 = 0; q4 < 4; ++q4)
			{
				const uint32_t wp = p0 + 4u * (uint32_t)q4, wd = ws[q4];
				if (wp >= lo && wp + 4u <= hi)
				{
					/* a whole word of this chunk: per-byte flags in bit 7 of each byte */
					const uint32_t ffm = ((wd & 0x7f7f7f7fu) + 0x01010101u) & wd & 0x80808080u;			/* byte == 0xFF */
					const uint32_t zm = ~(((wd & 0x7f7f7f7fu) + 0x7f7f7f7fu) | wd) & 0x80808080u;			/* byte == 0x00 */
					const uint32_t nb = wp + 4u >= end ? 0u : (ws[q4 + 1] & 255u);
					const uint32_t prev_ff = (ffm << 8) | (prev == 0xffu ? 0x80u : 0u);
					const uint32_t next_ff = (ffm >> 8) | (nb == 0xffu ? 0x80000000u : 0u);
					const uint32_t drop = (zm & prev_ff) | (ffm & next_ff);
					if (drop == 0u)
					{
						if (WRITE)
							ssu_put(o, __byte_perm(wd, 0, 0x0123), 4u);
						j += 4u;
					}
					else
					{
#pragma unroll
						for (int q = 0; q < 4; ++q)
							if (!((drop >> (8 * q + 7)) & 1u))
							{
								if (WRITE)
									ssu_put(o, ((wd >> (8 * q)) & 255u) << 24, 1u);
								++j;
							}
					}
					prev = wd >> 24;
				}
				else
				{
					/* a word that straddles the first or the last byte of the scan: byte by byte */
#pragma unroll
					for (int q = 0; q < 4; ++q)
					{
						const uint32_t p = wp + (uint32_t)q;
						const uint32_t cur = (wd >> (8 * q)) & 255u;
						uint32_t nxt = q < 3 ? (wd >> (8 * (q + 1))) & 255u : ws[q4 + 1] & 255u;
						if (p + 1u >= end)
							nxt = 0u;
						if (p >= lo && p < hi)
						{
							const bool dropped = (cur == 0u && prev == 0xffu) || (cur == 0xffu && nxt == 0xffu);
							if (!dropped)
							{
								if (WRITE)
									ssu_put(o, cur << 24, 1u);
								++j;
							}
							prev = cur;
						}
					}
				}
			}
		}
	}
	if (!WRITE)
		sb.nblk[gi] = j;
	/* bytes of a last word the next chunk completes (and, if no word left at all, not the ones below h0) */
	for (uint32_t b = o.h0; WRITE && b < o.nacc; ++b)
		ssu_sts8(o, 4u * o.jw + b, (uint32_t)(o.acc >> (56u - 8u * b)) & 255u);
	if (WRITE && threadIdx.x + 1u == w.count)
		s_jend = j; /* chunks are in stream order: the last one ends the CTA's output */
	if (WRITE)
		__syncthreads();
	if (WRITE)
	{
		/* shared -> global: whole words as coalesced lines; the bytes of a first / last word that another CTA shares, one by one */
		const uint32_t j_end = s_jend;
		const uint32_t head_end = min(j_end, (j_cta + 3u) & ~3u), tail_begin = max(head_end, j_end & ~3u);
		const uint8_t *sb8 = reinterpret_cast<const uint8_t *>(s_out);
		if (threadIdx.x == 0)
			for (uint32_t q = j_cta; q < head_end && (j_cta & 3u); ++q)
				dst[q ^ 3u] = sb8[(q ^ 3u) - 4u * o.jw_base];
		if (threadIdx.x == 32u % SS_THREADS)
			for (uint32_t q = tail_begin; q < j_end; ++q)
				dst[q ^ 3u] = sb8[(q ^ 3u) - 4u * o.jw_base];
		const uint32_t w0 = ((j_cta & 3u) ? head_end : j_cta) >> 2, w1 = tail_begin >> 2;
		uint32_t *dst32 = reinterpret_cast<uint32_t *>(dst);
		for (uint32_t q = w0 + threadIdx.x; q < w1; q += SS_THREADS)
			dst32[q] = s_out[q - o.jw_base];
	}
	if (WRITE && has_work && hi == end && (lo < end || i == 0))
	{
		/* the chunk that reaches the end of the scan: pad to 16 bytes, add one all-zero unit, publish the length */
		const uint32_t exact = j, padded = (exact + 15u) & ~15u;
		for (uint32_t q = exact; q < padded + 16u; ++q)
			dst[q ^ 3u] = 0u;
		useg_len[sc->seg_first] = padded + (padded - exact);
	}
}

/* ------------------------------------------------------------------------------------------------------
 * Stateless bit reader over the unstuffed slab: the next 32 stream bits at bit position bp are one funnel shift
 * of the two big-endian words around it; words behind the segment read as zero (what the reference feeds behind a
 * marker or the end of the file, jpeg.c:176-184, common.c:39).
 * ------------------------------------------------------------------------------------------------------ */
struct SsReader
{
	const uint32_t *src;
	uint32_t nwords;
	uint32_t cur, nxt, nx2; /* stream words bp / 32, + 1, + 2 of the position last primed / advanced to */
};
__device__ __forceinline__ uint32_t ssr_word(const SsReader &r, uint32_t wi) { return wi < r.nwords ? __ldg(r.src + wi) : 0u; }
__device__ __forceinline__ void ssr_prime(SsReader &r, uint32_t bp)
{
	const uint32_t wi = bp >> 5;
	r.cur = ssr_word(r, wi);
	r.nxt = ssr_word(r, wi + 1u);
	r.nx2 = ssr_word(r, wi + 2u);
}
/* a symbol is < 32 bits: at most one word boundary per step.  The word loaded here is not needed before the NEXT
   boundary, so its latency stays off the symbol-to-symbol chain. */
__device__ __forceinline__ void ssr_advance(SsReader &r, uint32_t &bp, uint32_t n)
{
	const uint32_t nbp = bp + n;
	if ((nbp ^ bp) & 32u)
	{
		r.cur = r.nxt;
		r.nxt = r.nx2;
		r.nx2 = ssr_word(r, (nbp >> 5) + 2u);
	}
	bp = nbp;
}

/* one symbol + magnitude at bp; false = invalid code (16 bits are consumed, as jpeg.c:226-230 does) */
template <int BITS, bool IS_DC>
__device__ __forceinline__ bool ss_symbol(SsReader &r, uint32_t &bp, const EntbLutT<BITS> *__restrict__ L, int &value, int &kadv)
{
	const uint32_t top = __funnelshift_l(r.nxt, r.cur, bp);
	uint32_t e = L->fast[top >> (32 - BITS)];
	if (e == 0)
	{
		const uint32_t t16 = top >> 16;
		int l = BITS + 1;
		while (t16 >= L->limit[l])
			++l;
		if (l > 16)
		{
			ssr_advance(r, bp, 16u);
			kadv = 1;
			value = 0;
			return false;
		}
		e = entb_entry(L->vals[(L->base[l] + (int)(t16 >> (16 - l))) & 255], l, IS_DC);
	}
	const uint32_t len = ENTB_E_LEN(e), size = ENTB_E_SIZE(e);
	kadv = (int)ENTB_E_KADV(e);
	const uint32_t after = top << len;
	const uint32_t s32 = 32u - size;
	const uint32_t v = shr_clamp(after, s32), mask = shr_clamp(0xffffffffu, s32);
	const int t = (int)after >> 31;
	value = (int)v - (int)(mask & ~(uint32_t)t);
	ssr_advance(r, bp, len + size);
	return true;
}

/*
 * One symbol of the SKELETON decode (guess and refine only need where symbols and blocks end, not what they say):
 * DC and AC symbols, and the end of a block, take the same instructions -- the table and the prefix width are
 * selected, the block bookkeeping is arithmetic -- so the 32 subsequences of a warp, which stand at unrelated places of
 * their blocks, do not serialise on these cases.  Only a code longer than the table's prefix branches.
 */
/* per block of an MCU: shared-memory addresses of its DC / AC lookup tables (built once per CTA behind entb_setup) */
template <typename SH>
__device__ __forceinline__ void ss_build_tabs(const SH &sh, uint2 *s_tabs, int tid)
{
	if (tid < sh.nblk)
	{
		const uint32_t ent = sh.blk[tid];
		s_tabs[tid] = make_uint2((uint32_t)__cvta_generic_to_shared(sh.dc[(ent >> 12) & 3u].fast),
										 (uint32_t)__cvta_generic_to_shared(sh.ac[(ent >> 16) & 3u].fast));
	}
	__syncthreads();
}
__device__ __forceinline__ uint2 ss_load_tabs(uint32_t tabs_addr, int b)
{
	uint2 t;
	asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(t.x), "=r"(t.y) : "r"(tabs_addr + 8u * (uint32_t)b));
	return t;
}
template <typename SH>
__device__ __forceinline__ void ss_skeleton_step(const SH &sh, uint32_t tabs_addr, uint2 &tabs, SsReader &r, uint32_t &bp, int &k, int &b,
																 uint32_t &nb, int nblk_mcu)
{
	const bool dc = k == 0;
	const uint32_t top = __funnelshift_l(r.nxt, r.cur, bp);
	const uint32_t idx = dc ? top >> (32 - ENTB_DC_BITS) : top >> (32 - ENTB_AC_BITS);
	uint32_t e = lds_u16((dc ? tabs.x : tabs.y) + (idx << 1));
	uint32_t adv;
	int kadv;
	if (e == 0u)
	{
		const uint32_t ent = sh.blk[b];
		const uint32_t tdc = (ent >> 12) & 3u, tac = (ent >> 16) & 3u;
		const uint32_t t16 = top >> 16;
		const uint32_t *limit = dc ? sh.dc[tdc].limit : sh.ac[tac].limit;
		const int32_t *basep = dc ? sh.dc[tdc].base : sh.ac[tac].base;
		const uint8_t *vals = dc ? sh.dc[tdc].vals : sh.ac[tac].vals;
		int l = (dc ? ENTB_DC_BITS : ENTB_AC_BITS) + 1;
		while (t16 >= limit[l])
			++l;
		if (l > 16)
		{
			adv = 16u; /* invalid code: 16 bits are consumed (jpeg.c:226-230), the block goes on */
			kadv = 1;
		}
		else
		{
			e = entb_entry(vals[(basep[l] + (int)(t16 >> (16 - l))) & 255], l, dc);
			adv = ENTB_E_LEN(e) + ENTB_E_SIZE(e);
			kadv = (int)ENTB_E_KADV(e);
		}
	}
	else
	{
		adv = ENTB_E_LEN(e) + ENTB_E_SIZE(e);
		kadv = (int)ENTB_E_KADV(e);
	}
	ssr_advance(r, bp, adv);
	k = dc ? 1 : k + kadv;
	const bool done = k >= 64;
	k = done ? 0 : k;
	nb += done ? 1u : 0u;
	if (done)
	{
		b = b + 1 == nblk_mcu ? 0 : b + 1;
		tabs = ss_load_tabs(tabs_addr, b); /* needed by the next block's first symbol only */
	}
}

enum
{
	SS_GUESS = 0,
	SS_WRITE = 2
};

/*
 * Phases 1 and 4.  grid = SsWork items; thread t handles subsequence work.first + t of the CTA's scan: the unstuffed
 * bits [2048 i, 2048 (i + 1)).  GUESS starts at the subsequence's first bit assuming "start of an MCU" and records
 * the exit state at the first symbol start past its end.  WRITE decodes from the converged entry state (the exit
 * state of the predecessor on side sb.result[0]) and stores coefficients.
 */
template <int MODE, typename SH>
__global__ void __launch_bounds__(SS_THREADS)
k_ss_decode(const SsWork *__restrict__ work, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
				const uint32_t *__restrict__ seg_start, const jcu_huff_spec *__restrict__ huff,
				const uint8_t *__restrict__ bitstream, const uint8_t *__restrict__ uslab, const uint32_t *__restrict__ useg_len,
				int16_t *__restrict__ coef, int32_t *__restrict__ img_err, SsBuffers sb)
{
	__shared__ SH sh;
	/* WRITE: one block under assembly per thread, natural order; 34 words apart so that the threads of a warp storing
	   the same coefficient hit 16 different banks */
	__shared__ __align__(8) unsigned char s_slot[MODE == SS_WRITE ? SS_THREADS * SS_SLOT_STRIDE : 8];
	const int tid = threadIdx.x;
	const SsWork w = work[blockIdx.x];
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	entb_setup(sh, sc, im, huff, tid, SS_THREADS);
	__shared__ uint2 s_tabs[MODE == SS_GUESS ? ENTB_MAX_MCU_BLOCKS : 1];
	if (MODE == SS_GUESS)
		ss_build_tabs(sh, s_tabs, tid);
	/* WRITE: per block of an MCU, what a thread needs when it starts that block -- the block's address is
	   coef + base + ux * sx + uy * sy (elements), the shared-memory addresses of its tables */
	__shared__ SsBlockInfo s_bi[MODE == SS_WRITE ? ENTB_MAX_MCU_BLOCKS : 1];
	if (MODE == SS_WRITE)
	{
		if (tid < sh.nblk)
		{
			const uint32_t ent = sh.blk[tid];
			const int sl = (int)(ent & 15u);
			SsBlockInfo bi;
			bi.base = (long long)sh.coff[sl] + ((long long)((ent >> 8) & 15u) * sh.bw[sl] + (long long)((ent >> 4) & 15u)) * 64;
			bi.sx = sh.hn[sl] * 64;
			bi.sy = sh.vn[sl] * sh.bw[sl] * 64;
			bi.dc_tab = (uint32_t)__cvta_generic_to_shared(sh.dc[(ent >> 12) & 3u].fast);
			bi.ac_tab = (uint32_t)__cvta_generic_to_shared(sh.ac[(ent >> 16) & 3u].fast);
			bi.zq = (uint32_t)__cvta_generic_to_shared(sh.zq[sl]);
			bi.pad = 0u;
			s_bi[tid] = bi;
		}
		__syncthreads();
	}
	/* the write pass flushes blocks warp-wide, so every thread of the CTA stays until the end; the others leave here */
	const bool has_work = (uint32_t)tid < w.count;
	if (MODE != SS_WRITE && !has_work)
		return;
	const uint32_t slot = (uint32_t)__cvta_generic_to_shared(s_slot) + (uint32_t)tid * SS_SLOT_STRIDE;
	if (MODE == SS_WRITE)
		for (int q = 0; q < 16; ++q)
			asm volatile("st.shared.v2.u32 [%0], {%1, %1};" ::"r"(slot + 8u * (uint32_t)q), "r"(0u) : "memory");

	const uint32_t i = w.first + (uint32_t)(has_work ? tid : 0), gi = w.scan_base + i;
	const uint32_t scan_start = seg_start[sc->seg_first];
	const uint32_t ul = useg_len[sc->seg_first];
	const uint32_t exact = (ul & ~15u) - (ul & 15u); /* bytes of the scan once the stuffing is gone */
	SsReader rd;
	rd.src = reinterpret_cast<const uint32_t *>(uslab + entb_useg_off(im->data_off, scan_start, (uint32_t)sc->seg_first));
	rd.nwords = (ul & ~15u) >> 2;
	const uint32_t sub = sb.sub_bytes;
	const uint32_t nsub_real = max(1u, (exact + sub - 1u) / sub); /* subsequences that hold data */
	const bool last = i + 1u == nsub_real;
	const uint32_t limit = (i + 1u) * (sub * 8u);
	const int nblk_mcu = sh.nblk;

	uint32_t bp = i * (sub * 8u);
	int b = 0, k = 0;
	if (MODE == SS_WRITE && i > 0)
	{
		const SsState *cur = sb.result[0] ? sb.exit_b : sb.exit_a;
		const SsState en = cur[gi - 1];
		bp = en.bp;
		k = (int)(en.aux & 255u);
		b = (int)((en.aux >> 8) & 0xfffu);
	}
	const bool alive = has_work && i < nsub_real;
	if (MODE != SS_WRITE && !alive)
	{
		/* behind the data (the raw scan was longer than the unstuffed one): nothing to decode; pass the state on */
		if (MODE == SS_GUESS)
		{
			SsState ex;
			ex.bp = bp;
			ex.aux = 0u;
			sb.exit_a[gi] = ex;
			sb.nblk[gi] = 0u;
			SsState none;
			none.bp = none.aux = 0xffffffffu;
			sb.entry_used[gi] = none;
		}
		return;
	}

	ssr_prime(rd, bp);
	/* write phase: global block counter and the address of the block in progress */
	const uint32_t total_blocks = (uint32_t)sc->total_units * (uint32_t)nblk_mcu;
	const int upr = sc->units_per_row;
	uint32_t g = (MODE == SS_WRITE && i > 0) ? sb.first_blk[gi] : 0u;
	int16_t *blk = nullptr;
	unsigned key = 0;
	auto locate = [&](uint32_t gb) {
		const uint32_t m = gb / (uint32_t)nblk_mcu;
		const uint32_t ent = sh.blk[gb - m * (uint32_t)nblk_mcu];
		const int s = (int)(ent & 15u);
		const int ux = (int)(m % (uint32_t)upr), uy = (int)(m / (uint32_t)upr);
		const int bx = ux * sh.hn[s] + (int)((ent >> 4) & 15u), by = uy * sh.vn[s] + (int)((ent >> 8) & 15u);
		blk = coef + sh.coff[s] + ((size_t)by * sh.bw[s] + bx) * 64;
		key = coef_key(blk);
	};
	const uint32_t g_entry = g;
	if (MODE == SS_WRITE && g < total_blocks)
		locate(g);

	uint32_t nb = 0;
	bool failed = false, in_slot_at_exit = false;
	if (MODE == SS_GUESS)
	{
		const uint32_t tabs_addr = (uint32_t)__cvta_generic_to_shared(s_tabs);
		uint2 tabs = ss_load_tabs(tabs_addr, b);
		while (bp < limit)
			ss_skeleton_step(sh, tabs_addr, tabs, rd, bp, k, b, nb, nblk_mcu);
	}
	else
	{
		/* One instruction path for DC and AC symbols (table, prefix width, target position and multiplier are selected):
		   the lanes of a warp stand at unrelated places of their blocks.  The end of a block is arithmetic as well -- the
		   counters step under a predicate and the next block's address and tables come from s_bi -- and the completed
		   blocks of a step are written by the WHOLE warp, one 128-byte line per block (round 2 first had each lane flush its
		   own block: 41 % of the warp instructions ran with 2.7 active lanes). */
		int ux = 0, uy = 0;
		if (g < total_blocks)
		{
			const uint32_t m = g / (uint32_t)nblk_mcu;
			ux = (int)(m % (uint32_t)upr);
			uy = (int)(m / (uint32_t)upr);
		}
		uint32_t dc_tab, ac_tab, zq_addr;
		{
			const SsBlockInfo bi = s_bi[b];
			dc_tab = bi.dc_tab;
			ac_tab = bi.ac_tab;
			zq_addr = bi.zq;
		}
		/* A block this thread STARTS is assembled in its shared-memory slot and leaves as one 128-byte line when it is
		   complete (whole sectors: no read-modify-write).  The block it enters in the middle of (the
		   predecessor began it) and the one it leaves unfinished go coefficient by coefficient into a block k_ss_zero_shared cleared, as
		   two threads share them. */
		bool in_slot = k == 0;
		const uint32_t lane = (uint32_t)tid & 31u;
		const uint32_t warp_slots = (uint32_t)__cvta_generic_to_shared(s_slot) + (uint32_t)(tid & ~31) * SS_SLOT_STRIDE + 4u * lane;
		bool active = alive && g < total_blocks && (last || bp < limit);
		while (__any_sync(0xffffffffu, active))
		{
			bool done = false;
			if (active)
			{
				const bool dc = k == 0;
				const uint32_t top = __funnelshift_l(rd.nxt, rd.cur, bp);
				const uint32_t idx = dc ? top >> (32 - ENTB_DC_BITS) : top >> (32 - ENTB_AC_BITS);
				uint32_t e = lds_u16((dc ? dc_tab : ac_tab) + (idx << 1));
				bool ok = true;
				if (e == 0u)
				{
					const uint32_t ent = sh.blk[b];
					const uint32_t tdc = (ent >> 12) & 3u, tac = (ent >> 16) & 3u;
					const uint32_t t16 = top >> 16;
					const uint32_t *lim = dc ? sh.dc[tdc].limit : sh.ac[tac].limit;
					const int32_t *basep = dc ? sh.dc[tdc].base : sh.ac[tac].base;
					const uint8_t *vals = dc ? sh.dc[tdc].vals : sh.ac[tac].vals;
					int l = (dc ? ENTB_DC_BITS : ENTB_AC_BITS) + 1;
					while (t16 >= lim[l])
						++l;
					ok = l <= 16;
					/* invalid code: 16 bits are consumed (jpeg.c:226-230) and the block goes on by one position */
					e = ok ? entb_entry(vals[(basep[l] + (int)(t16 >> (16 - l))) & 255], l, dc) : (1u | (16u << 11));
				}
				const uint32_t len = ENTB_E_LEN(e), size = ENTB_E_SIZE(e);
				const int kadv = (int)ENTB_E_KADV(e);
				const uint32_t after = top << len;
				const uint32_t s32 = 32u - size;
				const uint32_t v = shr_clamp(after, s32), mask = shr_clamp(0xffffffffu, s32);
				const int t = (int)after >> 31;
				const int val = (int)v - (int)(mask & ~(uint32_t)t);
				ssr_advance(rd, bp, len + size);
				failed |= !ok;
				/* DC: the raw difference at position 0 (phase 5 turns it into (short)(dc * q)); AC: (short)(value * q) */
				const uint32_t zqe = lds_u32(zq_addr + 4u * (uint32_t)(dc ? 0 : ((k + kadv - 1) & 127)));
				const int mul = dc ? 1 : (int)(zqe >> 16);
				if (ok && (dc || val))
				{
					const int nat = dc ? 0 : (int)ENTB_ZQ_NAT(zqe);
					if (in_slot)
						sts_u16(slot + 2u * (uint32_t)nat, (uint32_t)(val * mul));
					else
						blk[coef_pos(nat, key)] = (int16_t)(val * mul);
				}
				k = dc ? 1 : k + kadv;
				done = k >= 64;
			}
			/* ---- completed blocks that sit in a slot: lane l moves word l of each (chunk r = l / 4 goes to chunk position
			   r ^ key of the block, jpeg_coef_layout.cuh) and clears it for the owner's next block ---- */
			unsigned fm = __ballot_sync(0xffffffffu, done && in_slot);
			if (fm)
			{
				__syncwarp();
				do
				{
					const int L = __ffs(fm) - 1;
					fm &= fm - 1u;
					uint32_t *d = reinterpret_cast<uint32_t *>(__shfl_sync(0xffffffffu, (unsigned long long)blk, L));
					const uint32_t sa = warp_slots + (uint32_t)L * SS_SLOT_STRIDE;
					const uint32_t wv = lds_u32(sa);
					asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa), "r"(0u) : "memory");
					d[((((lane >> 2) ^ coef_key(d)) << 2) | (lane & 3u))] = wv;
				} while (fm);
				__syncwarp();
			}
			/* ---- the next block starts here ---- */
			{
				g += done ? 1u : 0u;
				const bool wrap_b = b + 1 == nblk_mcu;
				const bool step_x = done && wrap_b;
				b = done ? (wrap_b ? 0 : b + 1) : b;
				ux += step_x ? 1 : 0;
				const bool wrap_x = step_x && ux == upr;
				ux = wrap_x ? 0 : ux;
				uy += wrap_x ? 1 : 0;
				k = done ? 0 : k;
				in_slot = in_slot || done;
				if (done)
				{
					const SsBlockInfo bi = s_bi[b];
					blk = coef + (bi.base + (long long)ux * bi.sx + (long long)uy * bi.sy);
					dc_tab = bi.dc_tab;
					ac_tab = bi.ac_tab;
					zq_addr = bi.zq;
				}
			}
			active = active && g < total_blocks && (last || bp < limit);
		}
		in_slot_at_exit = in_slot;
	}
	if (MODE == SS_WRITE)
	{
		/* left in the middle of a block (the successor finishes it): if this thread started it, what was assembled so far
		   goes into the cleared block coefficient by coefficient, like the parts two threads share */
		if (alive && k != 0 && in_slot_at_exit && g < total_blocks)
		{
			key = coef_key(blk);
			for (int n = 0; n < 64; ++n)
			{
				const uint32_t v = lds_u16(slot + 2u * (uint32_t)n);
				if (v)
					blk[coef_pos(n, key)] = (int16_t)v;
			}
		}
		if (failed)
			report_failure(img_err, sc, 0u, JCU_ERR_BAD_HUFFMAN);
		else if (g_entry < total_blocks && g >= total_blocks)
		{
			/* this thread finished the scan's last block: the left-over rule of jpeg_entropy.cuh for a scan that simply runs
			   out of units, on the unstuffed copy as in k_entropy_baseline.  A stuffed 0xFF (a data byte 0xFF) wholly behind
			   the end position and beyond anything the reference can have buffered trips its marker loop; one inside the
			   32-bit window needs the exact fill level of its bit buffer: replay its reader over the scan (damaged files) */
			const int pos = (int)bp, R = 8 * (int)exact - pos;
			if (R >= 9)
			{
				const uint8_t *ub = reinterpret_cast<const uint8_t *>(rd.src);
				int f = -1;
				for (int q = (pos + 7) >> 3; q < (int)exact && f < 0; ++q)
					if (ub[q ^ 3] == 0xffu)
						f = q;
				if (f >= 0 && (8 * f >= pos + 32 ||
									entb_replay(sh, bitstream + im->data_off, im->data_len, scan_start, sc->total_units, false, sc->data_end) == REPLAY_TRIPS))
					report_failure(img_err, sc, 0u, JCU_ERR_UNKNOWN_MARKER);
			}
		}
		return;
	}
	SsState ex;
	ex.bp = bp;
	ex.aux = (uint32_t)k | ((uint32_t)b << 8);
	sb.nblk[gi] = nb;
	sb.exit_a[gi] = ex;
	SsState none;
	none.bp = none.aux = 0xffffffffu;
	sb.entry_used[gi] = none;
}

/*
 * Phase 2: Jacobi rounds.  One work item = SS_INNER iterations inside a CTA: iteration 0 takes every entry state from
 * the previous round (global memory), the following ones pass exit states from thread t-1 to thread t through shared
 * memory, so a correction travels up to SS_INNER subsequences per round.  A thread whose entry state did not
 * change does not decode again.
 */
#ifndef SS_INNER
#define SS_INNER 16 /* measured on 256 x 1080p 4:4:4: 8 -> 6 rounds 12.06 ms, 16 -> 4 rounds 11.92 ms, 32 -> 3 rounds 11.88 ms */
#endif
template <typename SH>
__device__ __forceinline__ bool ss_refine_item(SH &sh, uint2 *s_tabs, SsState *s_exit, const SsWork w, const jcu_scan *__restrict__ scans,
															  const jcu_image *__restrict__ images, const uint32_t *__restrict__ seg_start,
															  const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ uslab,
															  const uint32_t *__restrict__ useg_len, const SsBuffers &sb, int parity, int32_t *changed_flag,
															  bool rebuild)
{
	const int tid = threadIdx.x;
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	const uint32_t tabs_addr = (uint32_t)__cvta_generic_to_shared(s_tabs);
	const uint32_t i = w.first + (uint32_t)tid, gi = w.scan_base + i;
	const uint32_t ul = useg_len[sc->seg_first];
	const uint32_t exact = (ul & ~15u) - (ul & 15u);
	SsReader rd;
	rd.src = reinterpret_cast<const uint32_t *>(uslab + entb_useg_off(im->data_off, seg_start[sc->seg_first], (uint32_t)sc->seg_first));
	rd.nwords = (ul & ~15u) >> 2;
	const uint32_t nsub_real = max(1u, (exact + sb.sub_bytes - 1u) / sb.sub_bytes);
	const bool valid = (uint32_t)tid < w.count;
	const bool live = valid && i < nsub_real; /* subsequences behind the data only pass their entry state on */
	const uint32_t limit = (i + 1u) * (sb.sub_bytes * 8u);
	const SsState *cur = parity ? sb.exit_b : sb.exit_a;
	SsState *nxt = parity ? sb.exit_a : sb.exit_b;

	SsState mine, used, en0;
	mine.bp = mine.aux = 0;
	used = en0 = mine;
	if (valid)
	{
		mine = ss_load_l2(&cur[gi]);
		used = ss_load_l2(&sb.entry_used[gi]);
		en0 = i > 0 ? ss_load_l2(&cur[gi - 1]) : used;
	}
	/* after the first rounds most work items have nothing to do: find that out BEFORE the scan's tables are built */
	if (!__syncthreads_or(valid && !ss_same(en0, used)))
	{
		if (valid)
			nxt[gi] = mine;
		return false;
	}
	if (rebuild) /* consecutive work items of a CTA often belong to one scan: its tables are built once */
	{
		entb_setup(sh, sc, im, huff, tid, SS_THREADS);
		ss_build_tabs(sh, s_tabs, tid);
	}
	const int nblk_mcu = sh.nblk;
	const SsState first = mine;
	for (int it = 0; it < SS_INNER; ++it)
	{
		bool changed = false;
		if (valid && i > 0 && (it == 0 || tid > 0))
		{
			const SsState en = it == 0 ? en0 : s_exit[tid - 1];
			if (!ss_same(en, used))
			{
				used = en;
				SsState ex = en;
				if (live)
				{
					uint32_t bp = en.bp;
					int k = (int)(en.aux & 255u), b = (int)((en.aux >> 8) & 0xfffu);
					uint32_t nb = 0;
					ssr_prime(rd, bp);
					uint2 tabs = ss_load_tabs(tabs_addr, b);
					while (bp < limit)
						ss_skeleton_step(sh, tabs_addr, tabs, rd, bp, k, b, nb, nblk_mcu);
					ex.bp = bp;
					ex.aux = (uint32_t)k | ((uint32_t)b << 8);
					sb.nblk[gi] = nb;
				}
				changed = !ss_same(ex, mine);
				mine = ex;
			}
		}
		__syncthreads(); /* everyone has read s_exit of the previous iteration */
		s_exit[tid] = mine;
		if (!__syncthreads_or(changed))
			break;
	}
	if (valid)
	{
		sb.entry_used[gi] = used;
		nxt[gi] = mine;
		if (!ss_same(mine, first))
			atomicOr(changed_flag, 1);
	}
	__syncthreads(); /* the next work item of this CTA may rebuild the tables and reuses s_exit */
	return true; /* the tables in shared memory are this scan's */
}

/* barrier over all CTAs of a cooperative launch (they are co-resident): arrivals counter + generation */
__device__ __forceinline__ void ss_grid_barrier(uint32_t *sync, unsigned nctas)
{
	__syncthreads();
	if (threadIdx.x == 0)
	{
		volatile uint32_t *gen = sync + 1;
		const uint32_t g = *gen;
		__threadfence(); /* this CTA's exit states before its arrival */
		if (atomicAdd(sync, 1u) == nctas - 1u)
		{
			sync[0] = 0u;
			__threadfence();
			atomicAdd(sync + 1, 1u);
		}
		else
			while (*gen == g)
				__nanosleep(40);
		__threadfence();
	}
	__syncthreads();
}

/*
 * Phase 2 WITHOUT the host: one cooperative launch runs the Jacobi rounds to convergence.  Every round the CTAs share
 * out the work items, meet at a grid barrier and read the round's "anything changed" flag; the side of the ping-pong
 * that holds the final exit states goes to sb.result[0] for the write pass.  The reference has no counterpart (its
 * decoder is serial, jpeg.c:1198-1232); what this replaces is one kernel launch + one device-to-host flag copy + one
 * stream synchronisation per round.
 */
#ifndef SS_REFINE_MIN_CTAS
#define SS_REFINE_MIN_CTAS 8
#endif
template <typename SH>
__global__ void __launch_bounds__(SS_THREADS, SS_REFINE_MIN_CTAS)
k_ss_refine_loop(const SsWork *__restrict__ work, unsigned nwork, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
					  const uint32_t *__restrict__ seg_start, const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ uslab,
					  const uint32_t *__restrict__ useg_len, SsBuffers sb, unsigned max_rounds)
{
	__shared__ SH sh;
	__shared__ uint2 s_tabs[ENTB_MAX_MCU_BLOCKS];
	__shared__ SsState s_exit[SS_THREADS];
	__shared__ unsigned s_ticket;
	int parity = 0, have_scan = -1;
	unsigned round = 0;
#ifdef SS_TRACE /* timing experiment only: per round, CTA 0's time in its items and at the barrier */
	unsigned long long t_prev = 0;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_prev));
#endif
	for (; round < max_rounds; ++round)
	{
		int32_t *flag = sb.changed + round % 3u;
		/* Work items are handed out by ticket: an item takes one to sixteen decoding passes, so a fixed deal left most
		   CTAs waiting at the barrier for the unlucky ones (measured in the first round: 1.3 ms of work, 1.15 ms of
		   waiting).  Neighbouring items (one image) still spread over the SMs. */
		for (;;)
		{
			__syncthreads(); /* everybody has read the previous ticket */
			if (threadIdx.x == 0)
				s_ticket = atomicAdd(sb.ticket + round % 3u, 1u);
			__syncthreads();
			const unsigned it = s_ticket;
			if (it >= nwork)
				break;
			const SsWork w = work[it];
			if (ss_refine_item(sh, s_tabs, s_exit, w, scans, images, seg_start, huff, uslab, useg_len, sb, parity, flag, w.scan != have_scan))
				have_scan = w.scan;
		}
		parity ^= 1;
#ifdef SS_TRACE
		unsigned long long t_work = 0, t_bar = 0;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_work));
#endif
		ss_grid_barrier(sb.sync, gridDim.x);
#ifdef SS_TRACE
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_bar));
		if (threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x / 2))
			printf("refine cta %u round %u: own items %llu us, waited %llu us at the barrier\n", blockIdx.x, round, (t_work - t_prev) / 1000ull,
					 (t_bar - t_work) / 1000ull);
		t_prev = t_bar;
#endif
		const int32_t any = *reinterpret_cast<volatile int32_t *>(flag);
		/* the flag of round + 2 was last read behind the barrier of round - 1, which everybody has left */
		if (blockIdx.x == 0 && threadIdx.x == 0)
		{
			sb.changed[(round + 2u) % 3u] = 0;
			sb.ticket[(round + 2u) % 3u] = 0u; /* last used in round - 1, whose item loop everybody left before that round's barrier */
		}
		if (!any)
		{
			++round;
			break;
		}
	}
	if (blockIdx.x == 0 && threadIdx.x == 0)
	{
		sb.result[0] = parity;
		sb.result[1] = (int32_t)round;
	}
}

/* Phase 3: per scan, exclusive prefix sum of the per-subsequence block counts.  grid = self-sync scans. */
__global__ void __launch_bounds__(256)
k_ss_scan(const SsScan *__restrict__ ss, const uint32_t *__restrict__ nblk, uint32_t *__restrict__ first_blk)
{
	__shared__ uint32_t s_part[256];
	const SsScan s = ss[blockIdx.x];
	const int tid = threadIdx.x;
	const uint32_t chunk = (s.nsub + 255u) / 256u;
	const uint32_t lo = min(s.nsub, (uint32_t)tid * chunk), hi = min(s.nsub, lo + chunk);
	uint32_t sum = 0;
	for (uint32_t i = lo; i < hi; ++i)
		sum += nblk[s.base + i];
	s_part[tid] = sum;
	__syncthreads();
	if (tid == 0)
	{
		uint32_t run = 0;
		for (int i = 0; i < 256; ++i)
		{
			const uint32_t v = s_part[i];
			s_part[i] = run;
			run += v;
		}
	}
	__syncthreads();
	uint32_t run = s_part[tid];
	for (uint32_t i = lo; i < hi; ++i)
	{
		first_blk[s.base + i] = run;
		run += nblk[s.base + i];
	}
}

/*
 * Phase 3b: the write pass stores whole 128-byte blocks except where a block holds a subsequence boundary -- the thread
 * that began it and the one(s) that continue it each store single coefficients and rely on zeros elsewhere.  Only those
 * blocks are cleared here (one per subsequence that is entered in the middle of a block: its entry state has k != 0
 * and the block is number first_blk[i] of the scan); round 2 first cleared the whole coefficient slab with a memset
 * (3.2 GB per 256 x 1080p 4:4:4, 0.45 ms).  grid = decoding work items.
 */
__global__ void __launch_bounds__(SS_THREADS)
k_ss_zero_shared(const SsWork *__restrict__ work, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
					  int16_t *__restrict__ coef, SsBuffers sb)
{
	const SsWork w = work[blockIdx.x];
	if (threadIdx.x >= w.count)
		return;
	const uint32_t i = w.first + threadIdx.x, gi = w.scan_base + i;
	if (i == 0)
		return;
	const SsState *cur = sb.result[0] ? sb.exit_b : sb.exit_a;
	const SsState en = cur[gi - 1];
	if ((en.aux & 255u) == 0u)
		return; /* starts a block: that one leaves as a whole line */
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	int per = 0;
	for (int a = 0; a < sc->ncomp; ++a)
		per += sc->ncomp == 1 ? 1 : im->comp[sc->comp[a]].h * im->comp[sc->comp[a]].v;
	const uint32_t g = sb.first_blk[gi];
	if (g >= (uint32_t)sc->total_units * (uint32_t)per)
		return;
	const uint32_t m = g / (uint32_t)per;
	int r = (int)(g - m * (uint32_t)per);
	const int ux = (int)(m % (uint32_t)sc->units_per_row), uy = (int)(m / (uint32_t)sc->units_per_row);
	for (int a = 0; a < sc->ncomp; ++a)
	{
		const jcu_comp *cp = &im->comp[sc->comp[a]];
		const int hn = sc->ncomp == 1 ? 1 : cp->h, vn = sc->ncomp == 1 ? 1 : cp->v;
		if (r < hn * vn)
		{
			const int y = r / hn, x = r - y * hn;
			uint4 *blk = reinterpret_cast<uint4 *>(coef + cp->coef_off + ((size_t)(uy * vn + y) * cp->bw + (size_t)(ux * hn + x)) * 64);
#pragma unroll
			for (int q = 0; q < 8; ++q)
				blk[q] = make_uint4(0u, 0u, 0u, 0u);
			return;
		}
		r -= hn * vn;
	}
}

/*
 * Phase 5: DC prediction.  grid = (self-sync scans, 4 scan-component slots).  The blocks of one component in
 * decode order carry DC differences in coefficient 0; their running sum is the reference's dc_pred
 * (jpeg.c:323-325) and the stored value is (short)(dc * dequant[0]).
 */
#define SS_DC_THREADS 1024
#define SS_DC_ROUNDS 32 /* blocks per lane and tile: a tile = 32 warps x 32 rounds x 32 lanes = 32 768 blocks */
__global__ void __launch_bounds__(SS_DC_THREADS)
k_ss_dc(const SsScan *__restrict__ ss, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
		  int16_t *__restrict__ coef)
{
	__shared__ int s_warp[SS_DC_THREADS / 32];
	__shared__ int s_carry;
	const jcu_scan *sc = &scans[ss[blockIdx.x].scan];
	const int slot = blockIdx.y;
	if (slot >= sc->ncomp)
		return;
	const jcu_image *im = &images[sc->image];
	const jcu_comp *cp = &im->comp[sc->comp[slot]];
	const int hn = sc->ncomp == 1 ? 1 : cp->h, vn = sc->ncomp == 1 ? 1 : cp->v;
	const uint32_t per = (uint32_t)(hn * vn), upr = (uint32_t)sc->units_per_row;
	const int bw = cp->bw;
	const int q0 = (int)im->qt[sc->comp[slot]][0];
	const uint32_t n = (uint32_t)sc->total_units * per;
	int16_t *plane = coef + cp->coef_off;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	/* block j of the component in decode order -> its DC word.  A warp reads 32 CONSECUTIVE blocks per round (one
	   sector each, neighbours in the slab) and keeps the 32 rounds of a tile in registers: the slab is read once and
	   written once (the first version summed in one sweep and re-read everything in a second). */
	auto addr = [&](uint32_t j) -> int16_t * {
		const uint32_t m = j / per, r = j - m * per;
		const uint32_t uy = m / upr, ux = m - uy * upr;
		const uint32_t y = r / (uint32_t)hn, x = r - y * (uint32_t)hn;
		int16_t *blk = plane + ((size_t)(uy * (uint32_t)vn + y) * (size_t)bw + (ux * (uint32_t)hn + x)) * 64;
		return blk + coef_pos(0, coef_key(blk));
	};
	if (tid == 0)
		s_carry = 0;
	__syncthreads();
	const uint32_t tile_blocks = SS_DC_THREADS * SS_DC_ROUNDS;
	for (uint32_t t0 = 0; t0 < n; t0 += tile_blocks)
	{
		const uint32_t wlo = t0 + (uint32_t)wid * (32u * SS_DC_ROUNDS);
		int d[SS_DC_ROUNDS];
		int mine = 0;
#pragma unroll
		for (int q = 0; q < SS_DC_ROUNDS; ++q)
		{
			const uint32_t j = wlo + 32u * (uint32_t)q + (uint32_t)lane;
			d[q] = j < n ? (int)*addr(j) : 0;
		}
#pragma unroll
		for (int q = 0; q < SS_DC_ROUNDS; ++q)
			mine += d[q];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1)
			mine += __shfl_xor_sync(0xffffffffu, mine, o);
		if (lane == 0)
			s_warp[wid] = mine;
		__syncthreads();
		int pred = s_carry; /* dc_pred (jpeg.c:323) in front of this warp's first block */
		for (int w2 = 0; w2 < wid; ++w2)
			pred += s_warp[w2];
#pragma unroll
		for (int q = 0; q < SS_DC_ROUNDS; ++q)
		{
			int v = d[q];
#pragma unroll
			for (int o = 1; o < 32; o <<= 1)
			{
				const int u = __shfl_up_sync(0xffffffffu, v, o);
				if (lane >= o)
					v += u;
			}
			const uint32_t j = wlo + 32u * (uint32_t)q + (uint32_t)lane;
			if (j < n)
				*addr(j) = (int16_t)((pred + v) * q0);
			pred += __shfl_sync(0xffffffffu, v, 31);
		}
		__syncthreads();
		if (tid == SS_DC_THREADS - 1)
			s_carry = pred; /* the last warp's running value = everything up to the end of the tile */
		__syncthreads();
	}
}
