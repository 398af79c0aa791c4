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
 *   1 guess    thread i starts at byte i*SS_SUB_BYTES assuming "start of an MCU" and records its EXIT state
 *              (bit position, block-in-MCU, zig-zag index) at the first symbol start past its boundary.
 *   2 refine   every thread re-decodes its subsequence from its predecessor's exit state (Jacobi rounds,
 *              double-buffered); subsequence 0 starts from the true state, so correctness spreads at least
 *              one subsequence per round and, thanks to self-synchronisation, in practice converges in a few
 *              rounds; a thread whose entry state did not change skips the work.  Rounds stop when no exit
 *              state changed.
 *   3 scan     exclusive prefix sum of the number of blocks each subsequence completes.
 *   4 write    every thread decodes its subsequence once more from its (now true) entry state and writes
 *              (short)(value * dequant) coefficients (jpeg.c:345,365) into the pre-zeroed slab; DC symbols are
 *              written as raw differences.
 *   5 dc       per component: prefix sum of the DC differences in decode order = the reference's running
 *              dc_pred (jpeg.c:323-325), stored as (short)(dc * dequant[0]).
 *
 * Exit states must be comparable between decoders with different histories, so this bit reader is
 * canonical: it tops the 64-bit buffer up to more than 56 bits at EVERY symbol start, which makes
 * (read offset, buffered bits) a function of the logical bit position alone.  Byte stuffing (FF 00) is
 * undone inline (jpeg.c:167-187); a scan without restart markers contains no other FF xx.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"
#include "jpeg_entropy_baseline.cuh"

#define SS_THREADS 128
#define SS_SUB_BYTES 256

struct SsState
{
	uint32_t pos; /* read offset in the file (next byte to load) */
	uint32_t aux; /* buffered bits | zig-zag index k << 8 | block-in-MCU << 16 | zero bytes among the buffered bits << 28 */
	uint64_t buf; /* the buffered bits themselves (implied by pos/aux; kept so a successor can resume) */
};
__device__ __forceinline__ bool ss_same(const SsState &a, const SsState &b) { return a.pos == b.pos && a.aux == b.aux; }
/* exit states cross CTA boundaries INSIDE one launch of k_ss_refine_loop: read them from L2, never from an L1 line that
   an earlier round left behind */
static_assert(sizeof(SsState) == 16, "SsState is read as one 16-byte word");
__device__ __forceinline__ SsState ss_load_l2(const SsState *p)
{
	const uint4 v = __ldcg(reinterpret_cast<const uint4 *>(p));
	SsState s;
	s.pos = v.x;
	s.aux = v.y;
	s.buf = (uint64_t)v.z | ((uint64_t)v.w << 32);
	return s;
}

/* one CTA work item: up to SS_THREADS consecutive subsequences of one scan */
struct SsWork
{
	int32_t scan;
	uint32_t first;	  /* index within the scan of this CTA's first subsequence */
	uint32_t count;	  /* subsequences handled by this CTA */
	uint32_t scan_base; /* global index of the scan's subsequence 0 */
	uint32_t scan_nsub; /* subsequences in the scan */
	uint32_t pad;
};

struct SsBuffers
{
	SsState *exit_a, *exit_b; /* ping-pong exit states */
	SsState *entry_used;		  /* entry state the current exit state was computed from */
	uint32_t *nblk;			  /* blocks completed inside each subsequence */
	uint32_t *first_blk;		  /* exclusive prefix of nblk within the scan */
	int32_t *changed;			  /* [3] rotating flags: round t sets changed[t % 3] when an exit state changed */
	uint32_t *sync;			  /* [0] arrivals at the grid barrier of k_ss_refine_loop, [1] its generation */
	int32_t *result;			  /* [0] ping-pong side that holds the final exit states, [1] refine launches-worth of rounds run */
};

struct SBR
{
	uint32_t pos, end;
	uint64_t buf;
	int nbits;
	int pad; /* zero bits appended past the scan's end that are still buffered (bottom of the valid bits); a function of the position */
};

/* canonical top-up: > 56 valid bits; past the scan's end (or a marker) feeds zero bits */
__device__ __forceinline__ void sbr_topup(const uint8_t *__restrict__ base, SBR &br)
{
	while (br.nbits <= 56)
	{
		uint32_t b = 0;
		bool real = false;
		if (br.pos < br.end)
		{
			real = true;
			b = base[br.pos++];
			if (b == 0xff)
			{
				uint32_t q = br.pos;
				uint32_t c = q < br.end ? base[q] : 0u;
				while (c == 0xff)
				{
					++q;
					c = q < br.end ? base[q] : 0u;
				}
				if (c == 0)
					br.pos = q < br.end ? q + 1 : br.end;
				else
				{
					br.pos = br.end;
					b = 0;
					real = false;
				}
			}
		}
		if (!real)
			br.pad += 8;
		br.buf |= (uint64_t)b << (56 - br.nbits);
		br.nbits += 8;
	}
}

/* one symbol + magnitude from a topped-up buffer; false = invalid code */
template <int BITS, bool IS_DC>
__device__ __forceinline__ bool ss_symbol(SBR &br, const EntbLutT<BITS> *__restrict__ L, int &value, int &kadv)
{
	const uint32_t top = (uint32_t)(br.buf >> 32);
	uint32_t e = L->fast[top >> (32 - BITS)];
	if (e == 0)
	{
		const uint32_t t16 = top >> 16;
		int l = BITS + 1;
		while (t16 >= L->limit[l])
			++l;
		if (l > 16)
		{
			br.buf <<= 16;
			br.nbits -= 16;
			return false;
		}
		e = entb_entry(L->vals[(L->base[l] + (int)(t16 >> (16 - l))) & 255], l, IS_DC);
	}
	const int len = (int)ENTB_E_LEN(e), size = (int)ENTB_E_SIZE(e);
	kadv = (int)ENTB_E_KADV(e);
	const uint32_t after = top << len;
	const uint32_t s32 = 32u - (uint32_t)size;
	const uint32_t v = shr_clamp(after, s32), mask = shr_clamp(0xffffffffu, s32);
	const int t = (int)after >> 31;
	value = (int)v - (int)(mask & ~(uint32_t)t);
	const int total = len + size;
	br.buf <<= total;
	br.nbits -= total;
	return true;
}

enum
{
	SS_GUESS = 0,
	SS_REFINE = 1,
	SS_WRITE = 2
};

/* per self-sync scan */
struct SsScan
{
	int32_t scan;
	uint32_t base; /* global index of its subsequence 0 */
	uint32_t nsub;
	uint32_t pad;
};

/*
 * Phases 1, 2 and 4.  grid = SsWork items; thread t handles subsequence work.first + t of the CTA's scan.
 * parity selects the ping-pong side holding the CURRENT exit states (refine reads it and writes the other;
 * write reads it).
 */
template <int MODE>
__global__ void __launch_bounds__(SS_THREADS)
k_ss_decode(const SsWork *__restrict__ work, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
				const uint32_t *__restrict__ seg_start, const jcu_huff_spec *__restrict__ huff,
				const uint8_t *__restrict__ bitstream, int16_t *__restrict__ coef, int32_t *__restrict__ img_err, SsBuffers sb,
				int parity)
{
	__shared__ EntbShared sh;
	const int tid = threadIdx.x;
	const SsWork w = work[blockIdx.x];
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	entb_setup(sh, sc, im, huff, tid, SS_THREADS);
	if ((uint32_t)tid >= w.count)
		return;

	const uint32_t i = w.first + (uint32_t)tid, gi = w.scan_base + i;
	const uint8_t *__restrict__ base = bitstream + im->data_off;
	const uint32_t scan_start = seg_start[sc->seg_first], end = sc->data_end;
	const bool last = i + 1 == w.scan_nsub;
	const uint32_t limit = last ? end : scan_start + (i + 1) * SS_SUB_BYTES;
	if (parity < 0)
		parity = sb.result[0]; /* left there by k_ss_refine_loop */
	const SsState *cur = parity ? sb.exit_b : sb.exit_a;
	SsState *nxt = parity ? sb.exit_a : sb.exit_b;
	const int nblk_mcu = sh.nblk;

	SBR br;
	br.end = end;
	int b = 0, k = 0;
	if (MODE == SS_GUESS || i == 0)
	{
		if (MODE == SS_REFINE)
		{
			nxt[gi] = cur[gi]; /* subsequence 0 was decoded from the true start state in the guess phase */
			return;
		}
		uint32_t pos = scan_start + i * SS_SUB_BYTES;
		if (i > 0 && base[pos - 1] == 0xff && base[pos] == 0x00)
			++pos; /* do not start on the second byte of a stuffed pair */
		br.pos = pos;
		br.buf = 0;
		br.nbits = 0;
		br.pad = 0;
	}
	else
	{
		const SsState en = cur[gi - 1];
		if (MODE == SS_REFINE)
		{
			if (ss_same(en, sb.entry_used[gi]))
			{
				nxt[gi] = cur[gi];
				return;
			}
			sb.entry_used[gi] = en;
		}
		br.pos = en.pos;
		br.buf = en.buf;
		br.nbits = (int)(en.aux & 255u);
		k = (int)((en.aux >> 8) & 255u);
		b = (int)((en.aux >> 16) & 0xfffu);
		br.pad = (int)(en.aux >> 28) << 3;
	}

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
	bool failed = false;
	for (;;)
	{
		sbr_topup(base, br);
		if (MODE == SS_WRITE)
		{
			if (g >= total_blocks || (!last && br.pos >= limit))
				break;
		}
		else if (br.pos >= limit)
			break;
		const uint32_t ent = sh.blk[b];
		int val, kadv;
		if (k == 0)
		{
			const bool ok = ss_symbol<ENTB_DC_BITS, true>(br, &sh.dc[(ent >> 12) & 3u], val, kadv);
			if (MODE == SS_WRITE)
			{
				failed |= !ok;
				blk[coef_pos(0, key)] = (int16_t)val; /* DC difference; phase 5 turns it into (short)(dc * q) */
			}
			k = 1;
		}
		else
		{
			const bool ok = ss_symbol<ENTB_AC_BITS, false>(br, &sh.ac[(ent >> 16) & 3u], val, kadv);
			if (!ok)
				kadv = 1;
			const uint32_t e = sh.zq[ent & 15u][(k + kadv - 1) & 127];
			k += kadv;
			if (MODE == SS_WRITE)
			{
				failed |= !ok;
				if (ok && val)
					blk[coef_pos((int)ENTB_ZQ_NAT(e), key)] = (int16_t)(val * (int)(e >> 16));
			}
		}
		if (k >= 64)
		{
			k = 0;
			++nb;
			b = (b + 1 == nblk_mcu) ? 0 : b + 1;
			if (MODE == SS_WRITE)
			{
				++g;
				if (g < total_blocks)
					locate(g);
			}
		}
	}
	if (MODE == SS_WRITE)
	{
		if (failed)
			report_failure(img_err, sc, 0u, JCU_ERR_BAD_HUFFMAN);
		else if (g_entry < total_blocks && g >= total_blocks)
		{
			/* this thread finished the scan's last block: the left-over rule of jpeg_entropy.cuh (one segment, the last) */
			BitReader t;
			t.p = base + br.pos;
			t.end = base + end;
			t.buf = br.buf;
			t.nbits = br.nbits;
			t.stopped = br.pos >= end;
			t.pad = min(br.pad, br.nbits);
			/* a stuffed 0xFF more than 32 bits behind the end position is certainly unread; one inside the window needs
			   the exact fill level of the reference's bit buffer: replay its reader over the scan (rare, damaged files) */
			const int d_ff = br_first_ff_ahead(t);
			if (d_ff >= 32 || (d_ff >= 0 && entb_replay_trips(sh, base, im->data_len, scan_start, sc->total_units, false, end)))
				report_failure(img_err, sc, 0u, JCU_ERR_UNKNOWN_MARKER);
		}
		return;
	}
	SsState ex;
	ex.pos = br.pos;
	ex.aux = (uint32_t)br.nbits | ((uint32_t)k << 8) | ((uint32_t)b << 16) | ((uint32_t)(min(br.pad, br.nbits) >> 3) << 28);
	ex.buf = br.buf;
	sb.nblk[gi] = nb;
	if (MODE == SS_GUESS)
	{
		sb.exit_a[gi] = ex;
		SsState none;
		none.pos = 0xffffffffu;
		none.aux = 0xffffffffu;
		none.buf = 0;
		sb.entry_used[gi] = none;
	}
	else
	{
		if (!ss_same(ex, cur[gi]))
			atomicOr(sb.changed, 1);
		nxt[gi] = ex;
	}
}

/*
 * Phase 2, several Jacobi iterations per launch.  Same work split as k_ss_decode.  Iteration 0 takes every
 * entry state from the previous round (global memory); the following iterations pass exit states from thread
 * t-1 to thread t of the same CTA through shared memory, so a correction travels up to SS_INNER subsequences
 * per launch instead of one, and the many cheap tail rounds (a launch and a host round trip each) disappear.
 * A thread whose entry state did not change does not decode again.
 */
#define SS_INNER 8
__device__ __forceinline__ void ss_refine_item(EntbShared &sh, SsState *s_exit, const SsWork w, const jcu_scan *__restrict__ scans,
															  const jcu_image *__restrict__ images, const uint32_t *__restrict__ seg_start,
															  const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ bitstream, const SsBuffers &sb,
															  int parity, int32_t *changed_flag)
{
	const int tid = threadIdx.x;
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	entb_setup(sh, sc, im, huff, tid, SS_THREADS);
	const bool valid = (uint32_t)tid < w.count;
	const uint32_t i = w.first + (uint32_t)tid, gi = w.scan_base + i;
	const uint8_t *__restrict__ base = bitstream + im->data_off;
	const uint32_t scan_start = seg_start[sc->seg_first], end = sc->data_end;
	const bool last = i + 1 == w.scan_nsub;
	const uint32_t limit = last ? end : scan_start + (i + 1) * SS_SUB_BYTES;
	const SsState *cur = parity ? sb.exit_b : sb.exit_a;
	SsState *nxt = parity ? sb.exit_a : sb.exit_b;
	const int nblk_mcu = sh.nblk;

	SsState mine, used;
	mine.pos = mine.aux = 0;
	mine.buf = 0;
	used = mine;
	if (valid)
	{
		mine = ss_load_l2(&cur[gi]);
		used = ss_load_l2(&sb.entry_used[gi]);
	}
	const SsState first = mine;
	for (int it = 0; it < SS_INNER; ++it)
	{
		bool changed = false;
		if (valid && i > 0 && (it == 0 || tid > 0))
		{
			const SsState en = it == 0 ? ss_load_l2(&cur[gi - 1]) : s_exit[tid - 1];
			if (!ss_same(en, used))
			{
				used = en;
				SBR br;
				br.end = end;
				br.pos = en.pos;
				br.buf = en.buf;
				br.nbits = (int)(en.aux & 255u);
				br.pad = (int)(en.aux >> 28) << 3;
				int k = (int)((en.aux >> 8) & 255u), b = (int)((en.aux >> 16) & 0xfffu);
				uint32_t nb = 0;
				for (;;)
				{
					sbr_topup(base, br);
					if (br.pos >= limit)
						break;
					const uint32_t ent = sh.blk[b];
					int val, kadv;
					if (k == 0)
					{
						ss_symbol<ENTB_DC_BITS, true>(br, &sh.dc[(ent >> 12) & 3u], val, kadv);
						k = 1;
					}
					else
					{
						const bool ok = ss_symbol<ENTB_AC_BITS, false>(br, &sh.ac[(ent >> 16) & 3u], val, kadv);
						k += ok ? kadv : 1;
					}
					if (k >= 64)
					{
						k = 0;
						++nb;
						b = (b + 1 == nblk_mcu) ? 0 : b + 1;
					}
				}
				SsState ex;
				ex.pos = br.pos;
				ex.aux = (uint32_t)br.nbits | ((uint32_t)k << 8) | ((uint32_t)b << 16) | ((uint32_t)(min(br.pad, br.nbits) >> 3) << 28);
				ex.buf = br.buf;
				sb.nblk[gi] = nb;
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
	__syncthreads(); /* the next work item of this CTA rebuilds the tables and reuses s_exit */
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
 * out the work items (each item = SS_INNER in-CTA iterations as above), meet at a grid barrier and read the round's
 * "anything changed" flag; the side of the ping-pong that holds the final exit states goes to sb.result[0] for the
 * write pass.  The reference has no counterpart (its decoder is serial, jpeg.c:1198-1232); what this replaces is
 * one kernel launch + one device-to-host flag copy + one stream synchronisation per round.
 */
__global__ void __launch_bounds__(SS_THREADS)
k_ss_refine_loop(const SsWork *__restrict__ work, unsigned nwork, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
					  const uint32_t *__restrict__ seg_start, const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ bitstream,
					  SsBuffers sb, unsigned max_rounds)
{
	__shared__ EntbShared sh;
	__shared__ SsState s_exit[SS_THREADS];
	int parity = 0;
	unsigned round = 0;
	for (; round < max_rounds; ++round)
	{
		int32_t *flag = sb.changed + round % 3u;
		for (unsigned it = blockIdx.x; it < nwork; it += gridDim.x)
			ss_refine_item(sh, s_exit, work[it], scans, images, seg_start, huff, bitstream, sb, parity, flag);
		parity ^= 1;
		ss_grid_barrier(sb.sync, gridDim.x);
		const int32_t any = *reinterpret_cast<volatile int32_t *>(flag);
		/* the flag of round + 2 was last read behind the barrier of round - 1, which everybody has left */
		if (blockIdx.x == 0 && threadIdx.x == 0)
			sb.changed[(round + 2u) % 3u] = 0;
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
 * Phase 5: DC prediction.  grid = (self-sync scans, 4 scan-component slots).  The blocks of one component in
 * decode order carry DC differences in coefficient 0; their running sum is the reference's dc_pred
 * (jpeg.c:323-325) and the stored value is (short)(dc * dequant[0]).
 */
#define SS_DC_THREADS 1024
__global__ void __launch_bounds__(SS_DC_THREADS)
k_ss_dc(const SsScan *__restrict__ ss, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
		  int16_t *__restrict__ coef)
{
	__shared__ int s_part[SS_DC_THREADS];
	const jcu_scan *sc = &scans[ss[blockIdx.x].scan];
	const int slot = blockIdx.y;
	if (slot >= sc->ncomp)
		return;
	const jcu_image *im = &images[sc->image];
	const jcu_comp *cp = &im->comp[sc->comp[slot]];
	const int hn = sc->ncomp == 1 ? 1 : cp->h, vn = sc->ncomp == 1 ? 1 : cp->v, per = hn * vn;
	const int upr = sc->units_per_row, bw = cp->bw;
	const int q0 = (int)im->qt[sc->comp[slot]][0];
	const uint32_t n = (uint32_t)sc->total_units * (uint32_t)per;
	int16_t *plane = coef + cp->coef_off;
	const int tid = threadIdx.x;
	const uint32_t chunk = (n + SS_DC_THREADS - 1u) / SS_DC_THREADS;
	const uint32_t lo = min(n, (uint32_t)tid * chunk), hi = min(n, lo + chunk);
	auto dc_of = [&](uint32_t j) -> int16_t * {
		const uint32_t m = j / (uint32_t)per, r = j - m * (uint32_t)per;
		const int y = (int)r / hn, x = (int)r - y * hn;
		const int ux = (int)(m % (uint32_t)upr), uy = (int)(m / (uint32_t)upr);
		int16_t *blk = plane + ((size_t)(uy * vn + y) * bw + (ux * hn + x)) * 64;
		return blk + coef_pos(0, coef_key(blk));
	};
	int sum = 0;
	for (uint32_t j = lo; j < hi; ++j)
		sum += (int)*dc_of(j);
	s_part[tid] = sum;
	__syncthreads();
	if (tid == 0)
	{
		int run = 0;
		for (int i = 0; i < SS_DC_THREADS; ++i)
		{
			const int v = s_part[i];
			s_part[i] = run;
			run += v;
		}
	}
	__syncthreads();
	int pred = s_part[tid];
	for (uint32_t j = lo; j < hi; ++j)
	{
		int16_t *p = dc_of(j);
		pred += (int)*p;
		*p = (int16_t)(pred * q0);
	}
}
