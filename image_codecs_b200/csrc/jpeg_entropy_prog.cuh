/*
 * jpeg_entropy_prog.cuh -- warp-cooperative decode of progressive scans WITHOUT restart markers (sm_100a).
 *
 * Replaces stbi__jpeg_decode_block_prog_dc / _prog_ac (jpeg.c:372-558) driven by stbi__parse_entropy_coded_data
 * (jpeg.c:1198-1266) for scans that are one long restart segment.  Such a scan is inherently serial: where
 * block b+1 starts in the bitstream is known only after block b has been parsed, and in an AC refinement scan
 * the parse itself depends on which coefficients of the block are already non-zero (every one of them
 * consumes a correction bit, jpeg.c:489-497, 533-541), so not even a self-synchronising decoder can start in
 * the middle.  The design therefore shortens the serial dependency chain of ONE decoder to little more than
 * the table lookup, and moves everything else off that chain:
 *
 *   - ONE WARP PER SCAN, every lane executing the same parse on the same bits (warp-uniform control flow),
 *     so lane-parallel helpers cost nothing extra.
 *   - Stream: the lanes load the raw scan 128 bytes at a time (one word per lane, one chunk ahead), undo
 *     the byte stuffing of stbi__grow_buffer_unsafe (jpeg.c:167-187) together -- FF never emits, a 00 after
 *     FF emits FF, anything else after FF is the terminating marker; warp prefix sum for the compaction --
 *     and append to a 512-byte ring in shared memory.  The parser holds a 64-bit bit buffer and the next
 *     ring word in registers; past the end it is fed zero bits like the reference (jpeg.c:176-184).
 *   - AC refinement (jpeg.c:468-556): lane l owns the coefficients at zig-zag positions l and l+32 of the
 *     current block (the next block's are requested before the current one is parsed).  Two ballots give the
 *     block's history masks; each lane precomputes the zero-rank and non-zero-rank of its two positions.  A
 *     symbol (run r) then ends at the zero with rank j0 + r -- one compare per lane and a ballot -- and is
 *     followed by exactly (kz - k) - r correction bits; a lane picks the correction bit of its coefficient by
 *     its non-zero-rank.  No coefficient is touched on the serial chain.
 *   - AC first pass (jpeg.c:412-466): blocks inside an end-of-band run consume no bits, so a run advances the
 *     block counter by its whole length in one step.
 *   - DC refinement (jpeg.c:395-400) is one bit per block: 32 blocks per step, one per lane.
 *   - DC first pass (jpeg.c:382-392): one symbol per block on the same reader.
 *
 * Results are bit-identical to the reference, including its behaviour on damaged streams: positions past the
 * band map to coefficient 63 through the padded zig-zag table (jpeg.c:303), size != 1 in a refinement scan and
 * undecodable codes fail the image, and zero bits are fed past the end.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"
#include "jpeg_entropy.cuh"

#define PC_THREADS 32
#define PC_MAX_LUTS 2 /* 8 KB of shared memory per scan instead of 13: every scan of a 256-image batch is resident at once */
#define PC_RING_BYTES 512
#define PC_FULL 0xffffffffu

/* refinement-scan lookup (one entry per 9-bit prefix): bits consumed incl. the sign bit | run << 8 | class << 12 */
#define PC_RL_ZRL (1u << 12)
#define PC_RL_EOB (2u << 12)
#define PC_RL_SLOW (3u << 12) /* code longer than the prefix, or not allowed in a refinement scan */

struct PcShared
{
	HuffLut lut[PC_MAX_LUTS]; /* distinct tables of the scan (a DC scan with more goes to the per-segment decoder) */
	int lut_of[4];
	uint32_t ring[PC_RING_BYTES / 4];
	uint32_t tab[64];	 /* DC scans: per block of an MCU: slot | x << 4 | y << 8 | lut << 12 */
	uint16_t rlut[1 << ENT_LUT_BITS];
	uint8_t zpos[64];	 /* refinement: position of the band's t-th history-zero coefficient of the current block */
	unsigned long long dc_off[4]; /* DC scans, per scan component: first coefficient of the plane; blocks per row; MCU shape */
	int dc_bw[4], dc_hn[4], dc_vn[4];
	uint8_t zz[80];
};

struct PcStream
{
	/* consumer: top nb bits of buf are valid; nextw = ring word widx, already big-endian */
	uint64_t buf;
	int nb;
	uint32_t widx, nextw;
	/* producer */
	const uint8_t *base;	/* the file */
	uint32_t raw_pos;		/* file offset of the chunk held in pend (multiple of 4) */
	uint32_t start, end; /* entropy-coded bytes of the scan: [start, end) */
	uint32_t load_end;	/* loads stay below this offset */
	uint32_t out_bytes;	/* unstuffed bytes appended so far */
	uint32_t total;		/* unstuffed bytes of the whole scan, known once `ended` */
	uint32_t carry;		/* last raw byte of the previous chunk */
	uint32_t pend;			/* this lane's raw word of the chunk at raw_pos */
	bool ended;
};

__device__ __forceinline__ uint32_t pc_raw(const PcStream &s, uint32_t off)
{
	return off < s.load_end ? __ldg(reinterpret_cast<const uint32_t *>(s.base + off)) : 0u;
}

/* append the unstuffed bytes of one raw 128-byte chunk (or 128 zero bytes past the end) to the ring */
__device__ __forceinline__ void pc_produce(PcStream &s, uint8_t *ring_b, int lane)
{
	if (s.ended)
	{
#pragma unroll
		for (int i = 0; i < 4; ++i)
			ring_b[(s.out_bytes + 4u * (uint32_t)lane + (uint32_t)i) & (PC_RING_BYTES - 1)] = 0;
		s.out_bytes += 128u;
		__syncwarp();
		return;
	}
	const uint32_t w = s.pend;
	const uint32_t off = s.raw_pos + 4u * (uint32_t)lane;
	s.raw_pos += 128u;
	s.pend = pc_raw(s, s.raw_pos + 4u * (uint32_t)lane);
	uint32_t p = __shfl_up_sync(PC_FULL, w >> 24, 1);
	if (lane == 0)
		p = s.carry;
	s.carry = __shfl_sync(PC_FULL, w >> 24, 31);
	uint32_t packed = 0, cnt = 0;
	int endj = 4;
#pragma unroll
	for (int j = 0; j < 4; ++j)
	{
		const uint32_t o = off + (uint32_t)j;
		const bool before = o < s.start;
		const uint32_t b = before ? 0u : ((w >> (8 * j)) & 0xffu);
		const bool stop = o >= s.end || (p == 0xffu && b != 0xffu && b != 0u);
		if (stop && endj == 4)
			endj = j;
		if (!before && endj == 4 && b != 0xffu)
		{
			packed |= (p == 0xffu ? 0xffu : b) << (8u * cnt);
			++cnt;
		}
		p = b;
	}
	const uint32_t e = __ballot_sync(PC_FULL, endj < 4);
	if (e)
	{
		if (lane > __ffs((int)e) - 1)
			cnt = 0;
		s.ended = true;
	}
	uint32_t incl = cnt;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1)
	{
		const uint32_t t = __shfl_up_sync(PC_FULL, incl, d);
		if (lane >= d)
			incl += t;
	}
	const uint32_t pos = s.out_bytes + incl - cnt;
#pragma unroll
	for (int i = 0; i < 4; ++i)
		if ((uint32_t)i < cnt)
			ring_b[(pos + (uint32_t)i) & (PC_RING_BYTES - 1)] = (uint8_t)(packed >> (8 * i));
	s.out_bytes += __shfl_sync(PC_FULL, incl, 31);
	if (e)
		s.total = s.out_bytes;
	__syncwarp();
}

/* keep the ring at least 384 bytes ahead of the parser (a chunk adds at most 128) */
__device__ __forceinline__ void pc_fill(PcStream &s, uint32_t *ring, int lane)
{
	while (s.out_bytes - 4u * s.widx <= PC_RING_BYTES - 128u)
		pc_produce(s, reinterpret_cast<uint8_t *>(ring), lane);
}

__device__ __forceinline__ void pc_open(PcStream &s, uint32_t *ring, const uint8_t *file, uint32_t data_len, uint32_t start, uint32_t end,
													 int lane)
{
	s.base = file;
	s.start = start;
	s.end = end < start ? start : end;
	s.load_end = (data_len + 3u) & ~3u; /* the slab is padded past every file (jpeg_cuda.cu) */
	s.raw_pos = start & ~127u;
	s.out_bytes = 0;
	s.total = 0;
	s.carry = 0;
	s.ended = false;
	s.pend = pc_raw(s, s.raw_pos + 4u * (uint32_t)lane);
	s.widx = 0;
	pc_fill(s, ring, lane);
	s.nextw = __byte_perm(ring[0], 0, 0x0123);
	s.buf = 0;
	s.nb = 0;
}

/* at least 32 valid bits */
__device__ __forceinline__ void pc_need32(PcStream &s, uint32_t *ring, int lane)
{
	if (s.nb < 32)
	{
		s.buf |= (uint64_t)s.nextw << (32 - s.nb);
		s.nb += 32;
		++s.widx;
		s.nextw = __byte_perm(ring[s.widx & (PC_RING_BYTES / 4 - 1)], 0, 0x0123);
		pc_fill(s, ring, lane);
	}
}
__device__ __forceinline__ uint32_t pc_top(const PcStream &s) { return (uint32_t)(s.buf >> 32); }
__device__ __forceinline__ void pc_drop(PcStream &s, int n) /* 0 <= n <= 32 */
{
	s.buf <<= n;
	s.nb -= n;
}

/* one Huffman symbol from the next bits: code length, run, size; -1 = no such code.  When the table entry
   already holds the extended value (LUT_RESOLVED) the returned length covers the magnitude bits too. */
template <bool IS_DC>
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
		r = IS_DC ? 0 : sym >> 4;
		s = IS_DC ? sym : sym & 15;
		return l;
	}
	r = (int)((e >> 5) & 15u);
	if (e & LUT_RESOLVED)
	{
		resolved = true;
		value = (int)e >> 16;
		s = 1; /* refinement scans tell the exact size from `value` */
	}
	else
		s = (int)((e >> 9) & 15u);
	return (int)(e & 31u);
}
/* `s` magnitude bits that follow `len` code bits in top, extended (stbi__extend_receive, jpeg.c:250); 0 for s = 0 or s > 16 */
__device__ __forceinline__ int pc_extend(uint32_t top, int len, int s)
{
	if (s == 0 || s > 16)
		return 0;
	const int v = (int)((top << len) >> (32 - s));
	return v < (1 << (s - 1)) ? v + (int)(0xffffffffu << s) + 1 : v;
}

/*
 * Scan pipelining (PcPipe, round 2).  The scans of a progressive file form a dependency chain (first passes before
 * refinements of the same coefficients), and round 1 ran the chain level by level: one launch per level, each as long
 * as its slowest scan -- four to five scan durations per image.  But a scan only needs its predecessors to be AHEAD of
 * it, not finished: in pipelined mode every scan of an image runs at once, publishes how far down the image it is
 * final (in pixel rows, after every row of units) and waits, before each row of units, until its predecessors have
 * passed that row.  The chain then costs about its longest scan.  The work items are handed out by a ticket in
 * (level, scan) order, so a scan's predecessors always started before it and a waiting warp can never keep a
 * predecessor off the machine.  Coefficients written by another warp are read from L2 (ld.global.cg).
 */
#define PC_MAX_PREDS 4
struct PcPipe
{
	const int32_t *preds; /* per scan: PC_MAX_PREDS scan indices it must stay behind, -1 = none; NULL = not pipelined */
	int32_t *progress;	 /* per scan: pixel rows of the image below which the scan's output is final */
	uint32_t *ticket;		 /* next work item */
};
__device__ __forceinline__ int16_t pc_ld(const int16_t *p, bool l2) { return l2 ? __ldcg(p) : *p; }
/* before a row of units that covers pixel rows up to y_needed */
__device__ __forceinline__ void pc_wait(const PcPipe &pp, const int32_t *my_preds, int y_needed, int lane)
{
	if (!pp.preds)
		return;
	if (lane == 0)
		for (int i = 0; i < PC_MAX_PREDS; ++i)
		{
			const int p = my_preds[i];
			if (p < 0)
				continue;
			const volatile int32_t *f = pp.progress + p;
			unsigned spins = 0;
			while (*f < y_needed && ++spins < (1u << 28)) /* the cap only keeps a broken build from hanging the GPU */
				__nanosleep(200);
		}
	__syncwarp();
	__threadfence();
}
/* after a row of units: everything up to pixel row y_done is final */
__device__ __forceinline__ void pc_publish(const PcPipe &pp, int scan, int y_done, int lane)
{
	if (!pp.preds)
		return;
	__threadfence();
	__syncwarp();
	if (lane == 0)
		*reinterpret_cast<volatile int32_t *>(pp.progress + scan) = y_done;
}

/*
 * grid = CTA work items (one single-segment progressive scan each), 32 threads.
 */
__global__ void __launch_bounds__(PC_THREADS)
k_prog_coop(const EntWork *__restrict__ work, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
				const uint32_t *__restrict__ seg_start, const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ bitstream,
				int16_t *__restrict__ coef, int32_t *__restrict__ img_err, PcPipe pp)
{
	__shared__ PcShared sh;
	__shared__ uint32_t s_item;
	const int lane = threadIdx.x;
	const bool piped = pp.preds != nullptr;
	if (piped)
	{
		if (lane == 0)
			s_item = atomicAdd(pp.ticket, 1u);
		__syncwarp();
	}
	const EntWork w = work[piped ? s_item : blockIdx.x];
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	const int32_t *my_preds = piped ? pp.preds + (size_t)w.scan * PC_MAX_PREDS : nullptr;
	/* pixel rows one row of this scan's units covers */
	const int rows_px = 8 * (sc->ncomp == 1 ? im->vmax / max(1, im->comp[sc->comp[0]].v) : im->vmax);
	const int kind = sc->kind, ncs = sc->ncomp;
	const bool is_dc = kind == JCU_SCAN_PROG_DC_FIRST || kind == JCU_SCAN_PROG_DC_REFINE;
	for (int i = lane; i < 80; i += PC_THREADS)
		sh.zz[i] = c_natural[i];
	/* lookup tables: DC scans use the DC table of every scan component, AC scans one AC table */
	if (kind != JCU_SCAN_PROG_DC_REFINE)
		for (int i = 0; i < (is_dc ? ncs : 1); ++i)
		{
			const int spec = is_dc ? sc->dc_spec[i] : sc->ac_spec[i];
			if (spec < 0)
			{
				if (lane == 0)
					report_failure(img_err, sc, 0u, JCU_ERR_BAD_HUFFMAN);
				pc_publish(pp, w.scan, 0x7fffffff, lane);
				return;
			}
			int alias = -1, distinct = 0;
			for (int j = 0; j < i; ++j)
			{
				if ((is_dc ? sc->dc_spec[j] : sc->ac_spec[j]) == spec)
					alias = sh.lut_of[j];
				distinct = max(distinct, sh.lut_of[j] + 1);
			}
			if (alias < 0)
			{
				alias = min(distinct, PC_MAX_LUTS - 1); /* the host sends no scan with more distinct tables here */
				hufflut_build(&sh.lut[alias], &huff[spec], is_dc, lane, PC_THREADS);
			}
			if (lane == 0)
				sh.lut_of[i] = alias;
			__syncwarp();
		}
	/* blocks of one MCU (a single-component scan walks the component's own blocks: 1 x 1) */
	int nblk_mcu = 0;
	for (int i = 0; i < ncs; ++i)
	{
		const jcu_comp *cp = &im->comp[sc->comp[i]];
		const int hn = ncs == 1 ? 1 : cp->h, vn = ncs == 1 ? 1 : cp->v;
		const uint32_t li = kind == JCU_SCAN_PROG_DC_FIRST ? (uint32_t)(sh.lut_of[i] & 3) : 0u;
		for (int t = lane; t < hn * vn; t += PC_THREADS)
			if (nblk_mcu + t < 64)
				sh.tab[nblk_mcu + t] = (uint32_t)i | ((uint32_t)(t % hn) << 4) | ((uint32_t)(t / hn) << 8) | (li << 12);
		if (lane == 0)
		{
			sh.dc_off[i] = cp->coef_off;
			sh.dc_bw[i] = cp->bw;
			sh.dc_hn[i] = hn;
			sh.dc_vn[i] = vn;
		}
		nblk_mcu += hn * vn;
	}
	nblk_mcu = min(nblk_mcu, 64);
	if (kind == JCU_SCAN_PROG_AC_REFINE)
		for (int i = lane; i < (1 << ENT_LUT_BITS); i += PC_THREADS)
		{
			const uint32_t e = sh.lut[0].fast[i], len = e & 31u, r = (e >> 5) & 15u, s = (e >> 9) & 15u;
			uint32_t o = PC_RL_SLOW;
			if (len != 0u)
			{
				if (e & LUT_RESOLVED)
				{
					const int v = (int)e >> 16;
					if (v == 1 || v == -1)
						o = len | (r << 8); /* the table already counted the sign bit */
				}
				else if (s == 0u)
					o = len | (r << 8) | (r < 15u ? PC_RL_EOB : PC_RL_ZRL);
				else if (s == 1u)
					o = (len + 1u) | (r << 8);
			}
			sh.rlut[i] = (uint16_t)o;
		}
	__syncwarp();

	const int ss = sc->ss, se = sc->se, al = sc->al;
	const int total = sc->total_units, upr = sc->units_per_row;
	PcStream st;
	pc_open(st, sh.ring, bitstream + im->data_off, im->data_len, seg_start[sc->seg_first + w.seg_first], sc->data_end, lane);
	bool failed = false;

	if (kind == JCU_SCAN_PROG_DC_REFINE)
	{
		/* ---- one bit per block (jpeg.c:395-400): 32 blocks per step ---- */
		const uint32_t nblocks = (uint32_t)total * (uint32_t)nblk_mcu;
		const uint32_t per_row = (uint32_t)upr * (uint32_t)nblk_mcu; /* blocks of one row of units */
		uint32_t waited_to = 0; /* blocks covered by the rows waited for so far */
		for (uint32_t g0 = 0; g0 < nblocks; g0 += 32u)
		{
			if (piped && g0 + 32u > waited_to)
			{
				const uint32_t row_hi = (min(g0 + 31u, nblocks - 1u)) / per_row;
				pc_publish(pp, w.scan, (int)(g0 / per_row) * rows_px, lane); /* rows wholly below g0 are final */
				pc_wait(pp, my_preds, (int)(row_hi + 1u) * rows_px, lane);
				waited_to = (row_hi + 1u) * per_row;
			}
			pc_need32(st, sh.ring, lane);
			const uint32_t bits = pc_top(st);
			const uint32_t g = g0 + (uint32_t)lane;
			if (g < nblocks && ((bits >> (31 - lane)) & 1u))
			{
				const uint32_t m = g / (uint32_t)nblk_mcu, t = sh.tab[g - m * (uint32_t)nblk_mcu];
				const int slot = (int)(t & 15u);
				const int bx = (int)(m % (uint32_t)upr) * sh.dc_hn[slot] + (int)((t >> 4) & 15u);
				const int by = (int)(m / (uint32_t)upr) * sh.dc_vn[slot] + (int)((t >> 8) & 15u);
				int16_t *blk = coef + sh.dc_off[slot] + ((size_t)by * sh.dc_bw[slot] + bx) * 64;
				int16_t *p0 = &blk[coef_pos(0, coef_key(blk))];
				*p0 = (int16_t)(pc_ld(p0, piped) + (int16_t)(1 << al));
			}
			pc_drop(st, (int)min(32u, nblocks - g0));
		}
	}
	else if (kind == JCU_SCAN_PROG_DC_FIRST)
	{
		/* ---- one symbol per block (jpeg.c:382-392); the slab is zero, so only the DC value is written ---- */
		int dcp0 = 0, dcp1 = 0, dcp2 = 0, dcp3 = 0;
		int ux = 0, uy = 0;
		for (int m = 0; m < total && !failed; ++m)
		{
			if (piped && ux == 0)
			{
				pc_publish(pp, w.scan, uy * rows_px, lane);
				pc_wait(pp, my_preds, (uy + 1) * rows_px, lane);
			}
			for (int j = 0; j < nblk_mcu; ++j)
			{
				const uint32_t t = sh.tab[j];
				const int slot = (int)(t & 15u), li = (int)((t >> 12) & 3u);
				pc_need32(st, sh.ring, lane);
				const uint32_t top = pc_top(st);
				const uint32_t e = sh.lut[li].fast[top >> (32 - ENT_LUT_BITS)];
				int value, used;
				if ((e & 31u) == 0u)
				{
					int r, s;
					bool resolved;
					const int len = pc_symbol<true>(&sh.lut[li], top, r, s, value, resolved);
					if (len < 0)
					{
						failed = true;
						break;
					}
					value = pc_extend(top, len, s);
					used = len + ((s >= 1 && s <= 16) ? s : 0);
				}
				else if (e & LUT_RESOLVED)
				{
					value = (int)e >> 16;
					used = (int)(e & 31u);
				}
				else
				{
					const int len = (int)(e & 31u), s = (int)((e >> 9) & 15u);
					value = pc_extend(top, len, s);
					used = len + s;
				}
				pc_drop(st, used);
				int d;
				if (slot == 0)
					d = dcp0 += value;
				else if (slot == 1)
					d = dcp1 += value;
				else if (slot == 2)
					d = dcp2 += value;
				else
					d = dcp3 += value;
				if (lane == 0)
				{
					const int bx = ux * sh.dc_hn[slot] + (int)((t >> 4) & 15u), by = uy * sh.dc_vn[slot] + (int)((t >> 8) & 15u);
					int16_t *blk = coef + sh.dc_off[slot] + ((size_t)by * sh.dc_bw[slot] + bx) * 64;
					blk[coef_pos(0, coef_key(blk))] = (int16_t)(d * (1 << al));
				}
			}
			if (++ux == upr)
			{
				ux = 0;
				++uy;
			}
		}
	}
	else
	{
		int16_t *plane = coef + sh.dc_off[0];
		const int bw = sh.dc_bw[0];
		if (kind == JCU_SCAN_PROG_AC_FIRST)
		{
			/* ---- first pass of a band (jpeg.c:412-466): values go straight to the zeroed slab ---- */
			int u = 0;
			int row_end = 0; /* first unit behind the rows waited for */
			while (u < total && !failed)
			{
				if (piped && u >= row_end)
				{
					const int row = u / upr;
					pc_publish(pp, w.scan, row * rows_px, lane);
					pc_wait(pp, my_preds, (row + 1) * rows_px, lane);
					row_end = (row + 1) * upr;
				}
				int16_t *blk = plane + ((size_t)(u / upr) * bw + (u % upr)) * 64;
				const unsigned key = coef_key(blk);
				int k = ss, eob_run = 0;
				do
				{
					pc_need32(st, sh.ring, lane);
					const uint32_t top = pc_top(st);
					uint32_t e = sh.lut[0].fast[top >> (32 - ENT_LUT_BITS)];
					int r, s, value, len;
					bool resolved = false;
					if ((e & 31u) == 0u)
					{
						len = pc_symbol<false>(&sh.lut[0], top, r, s, value, resolved);
						if (len < 0)
						{
							failed = true;
							break;
						}
					}
					else
					{
						len = (int)(e & 31u);
						r = (int)((e >> 5) & 15u);
						s = (int)((e >> 9) & 15u);
						resolved = (e & LUT_RESOLVED) != 0u;
						value = (int)e >> 16;
					}
					if (s == 0 && !resolved)
					{
						if (r < 15)
						{
							eob_run = (1 << r) - 1;
							if (r)
								eob_run += (int)((top << len) >> (32 - r));
							pc_drop(st, len + r);
							break;
						}
						pc_drop(st, len);
						k += 16;
					}
					else
					{
						if (!resolved)
							value = pc_extend(top, len, s);
						pc_drop(st, len + s); /* a resolved entry carries s = 0: its length covers the magnitude */
						k += r;
						const int zig = sh.zz[k < 79 ? k : 79];
						++k;
						if (lane == 0)
							blk[coef_pos(zig, key)] = (int16_t)(value * (1 << al));
					}
				} while (k <= se);
				u += 1 + eob_run; /* the blocks of an end-of-band run consume no bits (jpeg.c:416-420) */
			}
		}
		else
		{
			/* ---- refinement of a band (jpeg.c:468-556) ---- */
			const int nat0 = sh.zz[lane], nat1 = sh.zz[lane + 32];
			const int bit = (int)(int16_t)(1 << al);
			const uint64_t band = (se >= 63 ? ~0ull : ((1ull << (se + 1)) - 1ull)) & ~((1ull << ss) - 1ull);
			const uint32_t band_lo = (uint32_t)band, band_hi = (uint32_t)(band >> 32);
			const uint32_t lt = (1u << lane) - 1u;
			const int z0 = lane, z1 = lane + 32;
			int eob_run = 0;
			int c0n = 0, c1n = 0; /* prefetched coefficients of the next block */
			int bx = 0;
			int16_t *row = plane, *blk = plane;
			int rowi = 0;
			if (total > 0)
			{
				pc_wait(pp, my_preds, rows_px, lane);
				const unsigned key = coef_key(blk);
				c0n = pc_ld(&blk[coef_pos(nat0, key)], piped);
				c1n = pc_ld(&blk[coef_pos(nat1, key)], piped);
			}
			for (int u = 0; u < total && !failed; ++u)
			{
				int16_t *cur = blk;
				const unsigned key = coef_key(cur);
				int c0 = c0n, c1 = c1n;
				const int o0 = c0, o1 = c1;
				if (++bx == upr)
				{
					bx = 0;
					row += (size_t)bw * 64;
					++rowi;
					if (piped && u + 1 < total)
					{
						/* the next block opens a new row of units: rows above the current one are final (the current block is
						   still being worked on), and the predecessors must have passed the new row before it is read */
						pc_publish(pp, w.scan, (rowi - 1) * rows_px, lane);
						pc_wait(pp, my_preds, (rowi + 1) * rows_px, lane);
					}
				}
				blk = row + (size_t)bx * 64;
				if (u + 1 < total)
				{
					const unsigned kn = coef_key(blk);
					c0n = pc_ld(&blk[coef_pos(nat0, kn)], piped);
					c1n = pc_ld(&blk[coef_pos(nat1, kn)], piped);
				}
				/* history masks inside the band, and this lane's ranks among the zeros / non-zeros below its positions */
				const uint32_t Mlo = __ballot_sync(PC_FULL, c0 != 0) & band_lo, Mhi = __ballot_sync(PC_FULL, c1 != 0) & band_hi;
				const uint32_t Zlo = ~Mlo & band_lo, Zhi = ~Mhi & band_hi;
				const int nnz = __popc(Mlo) + __popc(Mhi), nz = __popc(Zlo) + __popc(Zhi);
				const int nr0 = __popc(Mlo & lt), nr1 = __popc(Mlo) + __popc(Mhi & lt);
				const bool h0 = (Mlo >> lane) & 1u, h1 = (Mhi >> lane) & 1u; /* history non-zero */

				/* lanes take the correction bit of their coefficients: positions lo..hi, non-zero-rank counted from m0 */
				auto apply = [&](int lo, int hi, int m0, uint32_t s_hi, uint32_t s_lo) {
					if (h0 && z0 >= lo && z0 <= hi)
					{
						const int rank = nr0 - m0;
						const uint32_t b = rank < 32 ? (s_hi >> (31 - rank)) : (s_lo >> (63 - rank));
						if ((b & 1u) && (c0 & bit) == 0)
							c0 = c0 > 0 ? c0 + bit : c0 - bit;
					}
					if (h1 && z1 >= lo && z1 <= hi)
					{
						const int rank = nr1 - m0;
						const uint32_t b = rank < 32 ? (s_hi >> (31 - rank)) : (s_lo >> (63 - rank));
						if ((b & 1u) && (c1 & bit) == 0)
							c1 = c1 > 0 ? c1 + bit : c1 - bit;
					}
				};
				/* n <= 63 correction bits from a fresh window */
				auto correct = [&](int lo, int hi, int m0, int n) {
					if (n <= 0)
						return;
					pc_need32(st, sh.ring, lane);
					const uint32_t s_hi = pc_top(st);
					uint32_t s_lo = 0;
					if (n > 32)
					{
						pc_drop(st, 32);
						pc_need32(st, sh.ring, lane);
						s_lo = pc_top(st);
						pc_drop(st, n - 32);
					}
					else
						pc_drop(st, n);
					apply(lo, hi, m0, s_hi, s_lo);
				};

				if (eob_run)
				{
					--eob_run;
					correct(ss, se, 0, nnz);
				}
				else
				{
					/* position table of the band's zeros: a symbol with run r ends on the zero of rank j0 + r */
					__syncwarp();
					if ((Zlo >> lane) & 1u)
						sh.zpos[__popc(Zlo & lt)] = (uint8_t)z0;
					if ((Zhi >> lane) & 1u)
						sh.zpos[__popc(Zlo) + __popc(Zhi & lt)] = (uint8_t)z1;
					__syncwarp();
					int k = ss, j0 = 0, m0 = 0; /* position; zeros and non-zeros of the band below it */
					do
					{
						pc_need32(st, sh.ring, lane);
						const uint32_t top = pc_top(st);
						uint32_t o = sh.rlut[top >> (32 - ENT_LUT_BITS)];
						if (o >= PC_RL_SLOW)
						{
							int r_, s_, v_;
							bool res_;
							const int len = pc_symbol<false>(&sh.lut[0], top, r_, s_, v_, res_);
							if (len < 0 || res_ || s_ > 1)
							{
								failed = true; /* no such code, or a size other than 1 (jpeg.c:517) */
								break;
							}
							o = s_ == 0 ? ((uint32_t)len | ((uint32_t)r_ << 8) | (r_ < 15 ? PC_RL_EOB : PC_RL_ZRL))
											: ((uint32_t)(len + 1) | ((uint32_t)r_ << 8));
						}
						const int used = (int)(o & 63u), r = (int)((o >> 8) & 15u);
						const int target = j0 + r;
						if (o >= PC_RL_EOB)
						{
							eob_run = (1 << r) - 1;
							if (r)
								eob_run += (int)((top << used) >> (32 - r));
							pc_drop(st, used + r);
							correct(k, se, m0, nnz - m0);
							break;
						}
						if (target >= nz)
						{
							/* fewer than r + 1 zeros left: the run covers the rest of the band, nothing is placed */
							pc_drop(st, used);
							correct(k, se, m0, nnz - m0);
							break;
						}
						const int kz = (int)sh.zpos[target];
						const int n = kz - k - r;
						const int newval = (o & PC_RL_ZRL) ? 0 : (((top << (used - 1)) >> 31) ? bit : -bit);
						if (used + n <= 32)
						{
							const uint32_t s_hi = top << used;
							pc_drop(st, used + n);
							apply(k, kz, m0, s_hi, 0u);
						}
						else
						{
							pc_drop(st, used);
							correct(k, kz, m0, n);
						}
						if (kz == z0)
							c0 = newval; /* jpeg.c:545: the new coefficient (a zero for ZRL) lands on the zero the run stopped at */
						if (kz == z1)
							c1 = newval;
						k = kz + 1;
						j0 = target + 1;
						m0 += n;
					} while (k <= se);
				}
				if (c0 != o0)
					cur[coef_pos(nat0, key)] = (int16_t)c0;
				if (c1 != o1)
					cur[coef_pos(nat1, key)] = (int16_t)c1;
			}
		}
	}
	pc_publish(pp, w.scan, 0x7fffffff, lane); /* finished (or failed): nobody waits for this scan any more */
	if (failed)
	{
		if (lane == 0)
			report_failure(img_err, sc, 0u, JCU_ERR_BAD_HUFFMAN);
	}
	else
	{
		/* the left-over rule of jpeg_entropy.cuh (this scan is its own last segment): only when at least 25 real bits
		   lie behind the end position does one lane walk the scan's bytes */
		const long long pos = 32ll * (long long)st.widx - (long long)st.nb;
		if (lane == 0 && (!st.ended || (long long)st.total >= ((pos + 32) >> 3)))
		{
			const uint8_t *file = bitstream + im->data_off;
			if (leftover_trips_serial(file + seg_start[sc->seg_first + w.seg_first], file + im->data_len, pos, true,
													 sc->restart_interval > 0 && total % sc->restart_interval == 0))
				report_failure(img_err, sc, 0u, JCU_ERR_UNKNOWN_MARKER);
		}
	}
}
