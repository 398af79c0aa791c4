/*
 * jpeg_entropy_baseline.cuh -- baseline (SOF0/1) Huffman entropy decode, lock-step edition (sm_100a).
 *
 * Replaces stbi__jpeg_decode_block (jpeg.c:308-370) driven by stbi__parse_entropy_coded_data
 * (jpeg.c:1155-1234) for every baseline scan of a batch.  The progressive block decoders stay in
 * jpeg_entropy.cuh.  Two kernels per wave:
 *
 *   k_unstuff           one thread per restart segment copies the segment's entropy-coded bytes into a
 *                       second slab with the byte stuffing of stbi__grow_buffer_unsafe (jpeg.c:167-187)
 *                       already resolved: FF 00 -> FF, fill bytes dropped, everything from the first
 *                       marker on cut off; output as big-endian 32-bit words, zero padded to 16 bytes.
 *   k_entropy_baseline  one thread = one restart segment (stbi__jpeg_reset, jpeg.c:1142, clears every
 *                       carried state at a restart marker), ENTB_THREADS segments of ONE scan per CTA so
 *                       the scan's tables are built once in shared memory.  All lanes of a warp walk the
 *                       same (MCU, block) position of their own segment, so the per-block control flow is
 *                       warp-uniform; the only divergent loop is the AC symbol loop.
 *
 * With stuffing out of the way the bit reader is stateless: a lane keeps its bit position and the three
 * stream words around it in registers (next 32 bits = one funnel shift); crossing a word boundary slides the
 * window and re-reads the third word from a per-lane 64-byte ring in shared memory that 16-byte cp.async
 * copies top up at warp-uniform checkpoints (every ENTB_CHECK_EVERY symbols).  A checkpoint waits only for
 * the copies of three checkpoints ago, so the symbol loop never touches global memory, and the ring
 * provably never underflows (see ebr_checkpoint).  Past the end of a segment the ring is zero filled, which
 * is what the reference feeds after a marker or the end of the file (jpeg.c:176-184, common.c:39).
 *
 * One AC symbol is a single straight-line path:
 *
 *   entry  = lut[top 11 bits]           one 16-bit entry: k advance | magnitude bits | code length
 *   value  = extend(next `size` bits)   branch-free form of stbi__extend_receive (jpeg.c:250)
 *   slot[natural(k + run)] = (short)(value * q)      dequantised at decode time like jpeg.c:325,345,365
 *   k     += run + 1   (EOB: += 64, ZRL: += 16)
 *
 * EOB/ZRL need no branch: they carry size 0, so they store a zero at a position that has not been
 * written yet (coefficients are produced in increasing zig-zag order).  Codes longer than the table's
 * prefix take a canonical search (same result as jpeg.c:217-241).
 *
 * Output: int16 coefficients, natural order, 128-byte blocks, written once per block from a
 * zeroed shared-memory slot (the memset of jpeg.c:320).
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_entropy.cuh"
#include "jpeg_unstuff.h"

#ifndef ENTB_THREADS
#define ENTB_THREADS 256 /* restart segments per CTA: the tables are shared by more lanes, more warps fit an SM */
#endif
#define ENTB_SLOT_WORDS 32 /* 64 int16; byte o of lane l's block sits at o ^ entb_slot_key(l) */
#define ENTB_MAX_MCU_BLOCKS 64

#ifndef ENTB_DC_BITS
#define ENTB_DC_BITS 9
#endif
#ifndef ENTB_AC_BITS
#define ENTB_AC_BITS 11
#endif

/* 16-bit lookup entry: kadv (7) | size (4) << 7 | len (5) << 11 ; 0 = code longer than the table's prefix */
#define ENTB_E_KADV(e) ((e) & 127u)
#define ENTB_E_SIZE(e) (((e) >> 7) & 15u)
#define ENTB_E_LEN(e) ((e) >> 11)
template <int BITS>
struct EntbLutT
{
	uint16_t fast[1 << BITS];
	uint32_t limit[18]; /* limit[l] = (first code of length l + count) << (16 - l); [17] = ~0 */
	int32_t base[17];	  /* symbol index = base[l] + code */
	uint8_t vals[256];
};
typedef EntbLutT<ENTB_DC_BITS> EntbDc;
typedef EntbLutT<ENTB_AC_BITS> EntbAc;

template <int ND, int NA>
struct EntbSharedT
{
	EntbDc dc[ND];					 /* distinct DC tables of the scan, compact ids */
	EntbAc ac[NA];
	uint32_t zq[4][128];			 /* per slot: 2 * natural index | quantiser << 16, indexed by zig-zag position;
											 positions 64..127 alias natural 63 like the reference's padded table (jpeg.c:303) */
	uint32_t blk[ENTB_MAX_MCU_BLOCKS]; /* per block of an MCU: slot | dx << 4 | dy << 8 | dc lut << 12 | ac lut << 16 */
	int nblk;
	unsigned long long coff[4]; /* per slot: first coefficient of the component in the slab */
	int bw[4], hn[4], vn[4];	 /* per slot: blocks per plane row; blocks per MCU across / down (1 x 1 in single-component scans) */
};
typedef EntbSharedT<4, 4> EntbShared;
typedef EntbSharedT<2, 2> EntbShared2; /* luma + chroma pair: the usual case, half the table space */
#define ENTB_ZQ_NAT(e) (((e) & 0xffffu) >> 1)

/* entry for symbol `sym` of code length l (l <= 9) */
__device__ __forceinline__ uint32_t entb_entry(int sym, int l, bool is_dc)
{
	int size, kadv;
	if (is_dc)
	{
		size = sym > 15 ? 0 : sym; /* DC categories > 15 decode to 0 (jpeg.c:259) */
		kadv = 0;
	}
	else
	{
		size = sym & 15;
		const int run = sym >> 4;
		kadv = size ? run + 1 : (sym == 0xF0 ? 16 : 64); /* s == 0: ZRL or end of block (jpeg.c:352-355) */
	}
	return (uint32_t)kadv | ((uint32_t)size << 7) | ((uint32_t)l << 11);
}

template <int BITS>
__device__ __forceinline__ void entb_lut_build(EntbLutT<BITS> *L, const jcu_huff_spec *spec, bool is_dc, int tid, int nthreads)
{
	__shared__ uint32_t s_first[17];
	__shared__ uint32_t s_count[17];
	__shared__ int32_t s_valptr[17];
	__syncthreads();
	if (tid == 0)
	{
		uint32_t code = 0;
		int k = 0;
		for (int l = 1; l <= 16; ++l)
		{
			uint32_t n = spec->bits[l - 1];
			s_first[l] = code;
			s_count[l] = n;
			s_valptr[l] = k;
			L->base[l] = k - (int)code;
			code += n;
			k += (int)n;
			L->limit[l] = code << (16 - l);
			code <<= 1;
		}
		L->limit[17] = 0xffffffffu;
	}
	for (int i = tid; i < 256; i += nthreads)
		L->vals[i] = spec->vals[i];
	__syncthreads();
	for (int e = tid; e < (1 << BITS); e += nthreads)
	{
		uint32_t entry = 0;
		for (int l = 1; l <= BITS; ++l)
		{
			uint32_t code = (uint32_t)e >> (BITS - l);
			uint32_t rel = code - s_first[l];
			if (code >= s_first[l] && rel < s_count[l])
			{
				entry = entb_entry(spec->vals[(s_valptr[l] + (int)rel) & 255], l, is_dc);
				break;
			}
		}
		L->fast[e] = (uint16_t)entry;
	}
	__syncthreads();
}


/* per-scan tables, built once per CTA by all its threads (slots sharing a DHT table share one lookup) */
template <int ND, int NA>
__device__ __forceinline__ void entb_setup(EntbSharedT<ND, NA> &sh, const jcu_scan *sc, const jcu_image *im,
														 const jcu_huff_spec *__restrict__ huff, int tid, int nthreads)
{
	const int ncs = sc->ncomp;
	int lut_dc[4], lut_ac[4]; /* compact table ids per slot; resolved after unrolling: no dynamic indexing */
	int nd = 0, na = 0;
#pragma unroll
	for (int i = 0; i < 4; ++i)
	{
		lut_dc[i] = 0;
		lut_ac[i] = 0;
		if (i < ncs)
		{
			int ad = -1, aa = -1;
#pragma unroll
			for (int j = 0; j < 4; ++j)
				if (j < i)
				{
					if (sc->dc_spec[j] == sc->dc_spec[i] && ad < 0)
						ad = lut_dc[j];
					if (sc->ac_spec[j] == sc->ac_spec[i] && aa < 0)
						aa = lut_ac[j];
				}
			if (ad < 0)
			{
				ad = min(nd, ND - 1); /* the host picks ND/NA >= distinct tables of every scan in the launch */
				++nd;
				if (sc->dc_spec[i] >= 0)
					entb_lut_build(&sh.dc[ad], &huff[sc->dc_spec[i]], true, tid, nthreads);
			}
			if (aa < 0)
			{
				aa = min(na, NA - 1);
				++na;
				if (sc->ac_spec[i] >= 0)
					entb_lut_build(&sh.ac[aa], &huff[sc->ac_spec[i]], false, tid, nthreads);
			}
			lut_dc[i] = ad;
			lut_ac[i] = aa;
		}
	}
	for (int i = tid; i < ncs * 128; i += nthreads)
	{
		const int s = i >> 7, z = i & 127;
		const int nat = c_natural[z < 79 ? z : 79];
		sh.zq[s][z] = (uint32_t)(2 * nat) | ((uint32_t)im->qt[sc->comp[s]][nat] << 16);
	}
	if (tid == 0)
	{
		int n = 0;
#pragma unroll
		for (int i = 0; i < 4; ++i)
			if (i < ncs)
			{
				const jcu_comp *cp = &im->comp[sc->comp[i]];
				const int hn = ncs == 1 ? 1 : cp->h, vn = ncs == 1 ? 1 : cp->v;
				sh.coff[i] = cp->coef_off;
				sh.bw[i] = cp->bw;
				sh.hn[i] = hn;
				sh.vn[i] = vn;
				for (int y = 0; y < vn; ++y)
					for (int x = 0; x < hn; ++x)
						if (n < ENTB_MAX_MCU_BLOCKS)
							sh.blk[n++] = (uint32_t)i | ((uint32_t)x << 4) | ((uint32_t)y << 8) | ((uint32_t)lut_dc[i] << 12) |
											  ((uint32_t)lut_ac[i] << 16);
			}
		sh.nblk = n;
	}
	__syncthreads();
}

/* ------------------------------------------------------------------------------------------------------
 * Unstuffed slab.  Segment g (global index into the segment table) of a file at slab offset data_off whose
 * raw bytes start at file offset s is written at  data_off + align16(s) + 32 * g , zero padded to 16 bytes and
 * followed by one all-zero 16-byte unit (the source of every fetch past the end).  Raw segments are at least
 * 2 bytes apart (the RSTn marker) and never grow, so in 16-byte units the area of segment g holds
 * ceil(s'/16) - ceil(s/16) + 2 >= ceil(raw/16) + 1 units: the areas cannot overlap; the slab needs
 * bitstream_bytes + 32 * (n_segs + 2) bytes.  useg_len[g] = bytes produced, rounded up to 16, + the padding bytes (0..15) in the low 4 bits.
 * ------------------------------------------------------------------------------------------------------ */
__device__ __forceinline__ size_t entb_useg_off(uint64_t data_off, uint32_t seg_start, uint32_t g)
{
	return (size_t)data_off + (((size_t)seg_start + 15u) & ~(size_t)15u) + 32u * (size_t)g;
}

/*
 * One thread per restart segment; all lanes of a warp step through aligned raw words of their own segment
 * together.  A word without 0xFF is four output bytes; a word whose 0xFF bytes are all followed by 0x00 is
 * compacted branch-free (zero-byte mask & shifted FF mask -> 4-bit drop mask -> PRMT selector); only an
 * 0xFF followed by something else (a marker: the end of the segment, or fill bytes) takes the byte loop
 * with the exact rules of the reference.  Raw bytes are read as aligned 16-byte units, the next unit
 * requested before the current one is processed.
 */
__global__ void __launch_bounds__(ENTB_THREADS)
k_unstuff(const EntWork *__restrict__ work, const jcu_scan *__restrict__ scans, const jcu_image *__restrict__ images,
			 const uint32_t *__restrict__ seg_start, const uint8_t *__restrict__ bitstream, uint8_t *__restrict__ uslab,
			 uint32_t *__restrict__ useg_len)
{
	const EntWork w = work[blockIdx.x];
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	const bool active = threadIdx.x < w.nseg;
	const uint32_t g = (uint32_t)sc->seg_first + w.seg_first + (active ? threadIdx.x : 0u);
	const uint8_t *__restrict__ base = bitstream + im->data_off; /* 16-byte aligned */
	const uint32_t len = im->data_len;
	uint32_t pos = seg_start[g];
	/* The host walks the scan with the same byte rules (jpeg_parse.c walk_entropy_segments), so the first
	   marker after `pos` is the one that ends this segment: the loop below stops at least 2 raw bytes before
	   the next segment's start and the output area (sized by the raw distance) cannot overflow. */
	__shared__ uint32_t s_stage[4 * ENTB_THREADS];
	UnstuffOut o;
	o.wp = reinterpret_cast<uint32_t *>(uslab + entb_useg_off(im->data_off, pos, g));
	o.pend = 0u;
	o.nacc = 0u;
	o.words = 0u;
	o.stage = (uint32_t)__cvta_generic_to_shared(s_stage) + threadIdx.x * 4u;
	o.stride = ENTB_THREADS * 4u;
	bool more = active && pos < len;
	/* Raw bytes are read as aligned 32-byte sectors (two 16-byte loads), the next sector requested before the
	   current one is processed: the L1 is too small to keep a half-used sector per lane until the lane comes
	   back for the other half.  The slab is padded (jpeg_cuda.cu: + 64 bytes), so a whole sector may be read
	   whenever it starts inside the file. */
	uint32_t ub = pos & ~31u;
	uint4 na = make_uint4(0u, 0u, 0u, 0u), nb4 = na;
	if (more)
	{
		na = __ldg(reinterpret_cast<const uint4 *>(base + ub));
		nb4 = __ldg(reinterpret_cast<const uint4 *>(base + ub + 16u));
	}
	while (__any_sync(0xffffffffu, more))
	{
		const uint4 u0 = na, u1 = nb4;
		const uint32_t nb = ub + 32u;
		if (more && nb < len)
		{
			na = __ldg(reinterpret_cast<const uint4 *>(base + nb));
			nb4 = __ldg(reinterpret_cast<const uint4 *>(base + nb + 16u));
		}
		const uint32_t wv[9] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w, na.x};
#pragma unroll
		for (int k = 0; k < 8; ++k)
		{
			const uint32_t wpos = ub + 4u * (uint32_t)k;
			const uint32_t d = pos - wpos; /* leading bytes of this word that are not ours (or already consumed) */
			const bool mine = more && d < 4u;
			const uint32_t wd = wv[k];
			const uint32_t ffq = ((wd & 0x7f7f7f7fu) + 0x01010101u) & wd & 0x80808080u;
			const bool fast = !mine || (d == 0u && wpos + 4u <= len && ffq == 0u);
			if (__all_sync(0xffffffffu, fast))
			{
				if (mine)
				{
					unstuff_append(o, __byte_perm(wd, 0, 0x0123), 4u);
					pos = wpos + 4u;
				}
				continue;
			}
			if (mine)
				more = unstuff_word(base, len, wpos, wd, wv[k + 1], pos, o);
		}
		if (more)
		{
			const uint32_t want = pos & ~31u; /* a fill run may have jumped over whole sectors */
			if (want != nb)
			{
				na = want < len ? __ldg(reinterpret_cast<const uint4 *>(base + want)) : make_uint4(0u, 0u, 0u, 0u);
				nb4 = want < len ? __ldg(reinterpret_cast<const uint4 *>(base + want + 16u)) : make_uint4(0u, 0u, 0u, 0u);
			}
			ub = want;
		}
	}
	if (!active)
		return;
	const uint32_t exact = o.words * 4u + o.nacc; /* bytes of the segment once the stuffing is gone */
	if (o.nacc)
		unstuff_put(o, o.pend);
	while (o.words & 3u)
		unstuff_put(o, 0u);
	*reinterpret_cast<uint4 *>(o.wp) = make_uint4(0u, 0u, 0u, 0u);
	useg_len[g] = o.words * 4u + (o.words * 4u - exact); /* padded length (multiple of 16) + padding bytes in the low 4 bits */
}

/* ------------------------------------------------------------------------------------------------------
 * Bit reader over the unstuffed slab.  A lane keeps its bit position `bp` and the three stream words at
 * bp/32, +1, +2 in registers (cur, nxt, nx2); the next 32 stream bits are funnelshift(cur:nxt, bp % 32).
 * When a symbol crosses a word boundary the window slides and nx2 is re-read from the lane's ring of
 * ENTB_RING_WORDS words in shared memory (row u of the ring = 16-byte unit (index % 4 == u) of every lane,
 * filled by one 16-byte cp.async: the L1 pipe pays per distinct line touched, so 4-byte copies cost 4x).
 * That read is consumed at the next crossing at the earliest, so the only shared-memory latency on the
 * symbol-to-symbol dependency chain is the table lookup.
 * ------------------------------------------------------------------------------------------------------ */
#define ENTB_RING_WORDS 16
#define ENTB_RING_ROW (ENTB_THREADS * 16) /* one 16-byte unit of every lane */
#define ENTB_RING_BYTES (ENTB_RING_WORDS / 4 * ENTB_RING_ROW)
/* bits 7..8 of a bit position select the unit: shift them to the unit stride (ENTB_THREADS is a power of two) */
#define ENTB_RING_UNIT_SHIFT (ENTB_THREADS == 128 ? 4 : ENTB_THREADS == 256 ? 5 : 6)
static_assert(ENTB_THREADS == 128 || ENTB_THREADS == 256 || ENTB_THREADS == 512, "ENTB_THREADS: 128, 256 or 512");
#define ENTB_CHECK_EVERY 2 /* symbols between ring checkpoints */
#define ENTB_DYN_SMEM (ENTB_THREADS * ENTB_SLOT_WORDS * 4 + ENTB_RING_BYTES + ENTB_THREADS * 8)

struct EBR
{
	uint32_t bp;			 /* bit position in the segment's unstuffed stream + 64, so that bp / 32 is the index of nx2 */
	uint32_t cur, nxt, nx2; /* words bp/32, bp/32 + 1, bp/32 + 2 */
	uint32_t fw;			 /* next word to fetch (multiple of 4) */
	uint32_t nwords;		 /* words the segment holds (multiple of 4); the 4 words behind them are zero */
	const uint32_t *src; /* the segment in the unstuffed slab (16-byte aligned) */
};

__device__ __forceinline__ uint32_t lds_u32(uint32_t a)
{
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
	return v;
}
__device__ __forceinline__ uint32_t lds_u32_m4(uint32_t a)
{
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1+-4];" : "=r"(v) : "r"(a));
	return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t a)
{
	uint32_t v;
	asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a));
	return v;
}
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v)
{
	asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ebr_ring_addr(uint32_t ring_lane, uint32_t word)
{
	return ring_lane + ((word >> 2) & (ENTB_RING_WORDS / 4 - 1u)) * ENTB_RING_ROW + ((word & 3u) << 2);
}

/*
 * Warp-uniform checkpoint.  Each checkpoint commits one cp.async group and (WAIT) waits for all but the two
 * newest: a copy has three intervals to land.  After a checkpoint that found room, fw >= (unit of bp) + 16
 * words.  Between checkpoints j+2 and j+3 the reader touches at most word
 * (bp_j + 4 intervals * ENTB_CHECK_EVERY symbols * 31 bits) / 32 + 2, i.e. less than 44 bytes past bp_j,
 * which is below (unit of bp_j) + 64 bytes: issued at or before checkpoint j, landed.  One unit per
 * checkpoint keeps up because an interval consumes < 128 bits.
 */
template <int MAX_UNITS, bool WAIT>
__device__ __forceinline__ void ebr_checkpoint(uint32_t ring_lane, EBR &br)
{
	if (WAIT)
		asm volatile("cp.async.wait_group 2;" ::: "memory");
#pragma unroll
	for (int u = 0; u < MAX_UNITS; ++u)
	{
		/* the slot still holds words fw-16..fw-13: they must lie below the unit bp is in */
		if (br.fw + 4u <= (((br.bp >> 5) - 2u) & ~3u) + ENTB_RING_WORDS)
		{
			const uint32_t *g = br.src + min(br.fw, br.nwords); /* past the end: the segment's zero unit */
			const uint32_t d = ebr_ring_addr(ring_lane, br.fw);
			asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(g) : "memory");
			br.fw += 4u;
		}
	}
	asm volatile("cp.async.commit_group;" ::: "memory");
}

/* Additions of the symbol loop issued on the FMA pipe: the loop is bound by the ALU pipe (ncu: 73 % busy, math-pipe
   throttle 1.3 stalls per issue) while the FMA pipe idles at 15 %.  The multiplier comes from constant memory so
   that ptxas cannot fold the multiply-add back into an IADD3 / LEA. */
__constant__ uint32_t c_entb_one = 1u;
__device__ __forceinline__ uint32_t fma_add(uint32_t a, uint32_t b) { return a * c_entb_one + b; }
__device__ __forceinline__ uint32_t fma_sub(uint32_t a, uint32_t b) { return b * (0u - c_entb_one) + a; } /* a - b */

__device__ __forceinline__ uint32_t shr_clamp(uint32_t x, uint32_t n) /* n >= 32 gives 0 (PTX shr semantics) */
{
	uint32_t d;
	asm("shr.u32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(n));
	return d;
}

/*
 * One Huffman symbol plus its magnitude bits.  Returns false for "bad huffman code" (jpeg.c:317,351).
 * value = extended magnitude (0 when the symbol carries no bits); e = the symbol's table entry.
 * `between` runs after the table lookup has been issued and before its result is needed.
 */
template <int BITS, bool IS_DC, typename F>
__device__ __forceinline__ bool entb_symbol(uint32_t ring_lane, EBR &br, uint32_t fast_addr, const EntbLutT<BITS> *__restrict__ L,
														  int &value, uint32_t &e, F between)
{
	const uint32_t top = __funnelshift_l(br.nxt, br.cur, br.bp);
	e = lds_u16(fma_add((top >> (32 - BITS)) << 1, fast_addr));
	between();
	if (e == 0u)
	{
		const uint32_t t16 = top >> 16;
		int l = BITS + 1;
		while (t16 >= L->limit[l])
			++l;
		if (l > 16)
			return false;
		e = entb_entry(L->vals[(L->base[l] + (int)(t16 >> (16 - l))) & 255], l, IS_DC);
	}
	const uint32_t len = ENTB_E_LEN(e), size = ENTB_E_SIZE(e);
	const uint32_t after = top << len;
	const uint32_t s32 = fma_sub(32u, size);
	const uint32_t v = shr_clamp(after, s32), mask = shr_clamp(0xffffffffu, s32);
	const int t = (int)after >> 31;
	value = (int)fma_sub(v, mask & ~(uint32_t)t);
	const uint32_t nbp = fma_add(len, br.bp) + size;
	/* a symbol is < 32 bits: at most one word boundary.  Predicated in place: no temporary whose copy would
	   wait for the read that was just issued. */
	/* ring address of word nbp / 32 (the new nx2): unit (w / 4) % 4 at ENTB_RING_ROW strides, word w % 4 inside it;
	   both fields come from bits 5..8 of nbp */
	const uint32_t m = nbp & 0x1e0u;
	const uint32_t ra = fma_add(((m >> 3) | (m << ENTB_RING_UNIT_SHIFT)) & (0xcu | (3u * ENTB_RING_ROW)), ring_lane);
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p mov.u32 %0, %1;\n\t@p mov.u32 %1, %2;\n\t"
					 "@p ld.shared.u32 %2, [%4];\n\t}"
					 : "+r"(br.cur), "+r"(br.nxt), "+r"(br.nx2)
					 : "r"((nbp ^ br.bp) & 32u), "r"(ra));
	br.bp = nbp;
	return true;
}

/* ------------------------------------------------------------------------------------------------------
 * Rare path of the left-over rule (jpeg_entropy.cuh).  When a scan simply runs out of units, how many bytes the
 * reference's reader holds beyond the end position depends on where its last refill happened.  If a stuffed 0xFF
 * lies inside that window (within 32 bits of the end position -- damaged files only), one thread replays the
 * reference's reader literally over the raw bytes of the segment: stbi__grow_buffer_unsafe (jpeg.c:167-187), the
 * code_bits bookkeeping of stbi__jpeg_huff_decode (:205-243) and stbi__extend_receive (:249-266), the refill
 * checks of stbi__jpeg_decode_block (:308-370) and the restart check (:1180-1189).  No coefficient is produced.
 * Returns true when the reference leaves the scan without having seen the marker behind it and its marker loop
 * (jpeg.c:1722-1735) then runs into an 0xFF before that marker.
 * ------------------------------------------------------------------------------------------------------ */
struct ReplayReader
{
	const uint8_t *base;
	uint32_t len, pos, buf;
	int cb;
	bool nomore;
};
__device__ __forceinline__ uint32_t replay_get8(ReplayReader &r)
{
	const uint32_t b = r.pos < r.len ? r.base[r.pos] : 0u; /* stbi__get8 at the end of the data: 0 */
	++r.pos;
	return b;
}
__device__ __noinline__ void replay_grow(ReplayReader &r)
{
	do
	{
		const uint32_t b = r.nomore ? 0u : replay_get8(r);
		if (b == 0xffu)
		{
			uint32_t c = replay_get8(r);
			while (c == 0xffu)
				c = replay_get8(r);
			if (c != 0u)
			{
				r.nomore = true;
				return;
			}
		}
		if (b)
			r.buf |= b << (24 - r.cb);
		r.cb += 8;
	} while (r.cb <= 24);
}
enum
{
	REPLAY_FINE = 0,
	REPLAY_TRIPS = 1, /* the reference leaves the scan without having seen the marker behind it and its marker loop runs into an 0xFF */
	REPLAY_BAD = 2		/* "bad huffman code": no such code */
};
/* One symbol as the reference's reader consumes it.  The refill that MEETS the marker adds nothing (replay_grow), but
   a baseline block refills below 16 bits itself (jpeg.c:313, 332) and stbi__jpeg_huff_decode does so again (:199):
   behind a freshly met marker the second call pads with zeros, so a code that straddles the end of the data decodes
   from zero bits ("bad huffman code" at jpeg.c:218, 237 cannot happen here; it can in a progressive AC refinement,
   which has only the one refill).  AC symbols the reference's fast-AC table holds -- code + magnitude in 9
   bits, 1..7 magnitude bits (8 would leave -128..127), jpeg.c:138-165 -- are consumed at :335-346 without a
   further refill. */
template <int BITS>
__device__ __forceinline__ int replay_symbol(ReplayReader &r, const EntbLutT<BITS> *__restrict__ L, bool is_dc, uint32_t &e)
{
	if (r.cb < 16)
		replay_grow(r);
	const bool fast_candidate = !is_dc;
	uint32_t top = r.buf;
	if (fast_candidate)
	{
		const uint32_t e9 = L->fast[top >> (32 - BITS)];
		const int len9 = (int)ENTB_E_LEN(e9), size9 = (int)ENTB_E_SIZE(e9);
		if (e9 != 0u && size9 >= 1 && size9 <= 7 && len9 + size9 <= 9)
		{
			e = e9;
			r.buf <<= len9 + size9;
			r.cb -= len9 + size9;
			return REPLAY_FINE;
		}
	}
	if (r.cb < 16)
		replay_grow(r); /* stbi__jpeg_huff_decode's own refill */
	top = r.buf;
	e = L->fast[top >> (32 - BITS)];
	if (e == 0u)
	{
		const uint32_t t16 = top >> 16;
		int l = BITS + 1;
		while (t16 >= L->limit[l])
			++l;
		if (l > 16)
			return REPLAY_BAD; /* no such code, jpeg.c:226-230 */
		e = entb_entry(L->vals[(L->base[l] + (int)(t16 >> (16 - l))) & 255], l, is_dc);
	}
	const int len = (int)ENTB_E_LEN(e), size = (int)ENTB_E_SIZE(e);
	if (len > r.cb)
		return REPLAY_BAD; /* jpeg.c:218, 237: unreachable behind the two refills, kept for the letter of it */
	r.buf <<= len;
	r.cb -= len;
	if (size)
	{
		if (r.cb < size)
			replay_grow(r);
		r.buf <<= size;
		r.cb -= size;
	}
	return REPLAY_FINE;
}
template <typename SH>
__device__ __noinline__ int entb_replay(const SH &sh, const uint8_t *base, uint32_t len, uint32_t start, int n_units, bool at_restart,
													 uint32_t data_end)
{
	ReplayReader r;
	r.base = base;
	r.len = len;
	r.pos = start;
	r.buf = 0u;
	r.cb = 0;
	r.nomore = false;
	const int nb = sh.nblk;
	for (int u = 0; u < n_units; ++u)
		for (int b = 0; b < nb; ++b)
		{
			const uint32_t ent = sh.blk[b];
			uint32_t e;
			if (replay_symbol<ENTB_DC_BITS>(r, &sh.dc[(ent >> 12) & 3u], true, e) != REPLAY_FINE)
				return REPLAY_BAD;
			int k = 1;
			do
			{
				if (replay_symbol<ENTB_AC_BITS>(r, &sh.ac[(ent >> 16) & 3u], false, e) != REPLAY_FINE)
					return REPLAY_BAD;
				k += (int)ENTB_E_KADV(e);
			} while (k < 64);
		}
	if (at_restart && r.cb < 24)
		replay_grow(r);
	if (r.nomore)
		return REPLAY_FINE;
	for (uint32_t q = r.pos; q < len && q < data_end; ++q)
		if (base[q] == 0xffu)
			return REPLAY_TRIPS;
	return REPLAY_FINE;
}

/*
 * grid = CTA work items (EntWork) of one wave restricted to baseline scans; thread t decodes restart
 * segment (work.seg_first + t) of the CTA's scan from the unstuffed slab.  <ND, NA> = distinct DC / AC
 * tables the CTA can hold: the host launches <2, 2> when no scan of the wave needs more (the usual luma +
 * chroma pair), which fits one more CTA per SM.  Dynamic shared memory (ENTB_DYN_SMEM): slots, then rings.
 *
 * The warp issues in order, so every shared-memory round trip that is consumed right away stalls all of
 * it.  The symbol loop therefore has exactly one exposed round trip (the table lookup): the dequantiser
 * entry of symbol i is requested as soon as its position is known and the coefficient is stored during
 * the lookup of symbol i+1.
 */
template <int ND, int NA>
__global__ void __launch_bounds__(ENTB_THREADS)
k_entropy_baseline(const EntWork *__restrict__ work, const jcu_scan *__restrict__ scans,
						 const jcu_image *__restrict__ images, const uint32_t *__restrict__ seg_start,
						 const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ uslab,
						 const uint32_t *__restrict__ useg_len, int16_t *__restrict__ coef, int32_t *__restrict__ img_err,
						 const uint8_t *__restrict__ bitstream)
{
	__shared__ EntbSharedT<ND, NA> sh;
	extern __shared__ __align__(16) uint32_t s_dyn[];

	const int tid = threadIdx.x;
	const EntWork w = work[blockIdx.x];
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	entb_setup(sh, sc, im, huff, tid, ENTB_THREADS);

	/* ---- this thread's segment ---- */
	const bool active = (uint32_t)tid < w.nseg;
	const uint32_t seg = w.seg_first + (uint32_t)tid;
	const int ri = sc->restart_interval, total = sc->total_units, upr = sc->units_per_row;
	const int n_units = ri ? ri : total;
	const int u0 = ri ? (int)seg * ri : 0;
	const int u1 = ri ? min(u0 + ri, total) : total;
	const int nblk = sh.nblk;
	const uint32_t ring_lane = (uint32_t)__cvta_generic_to_shared(s_dyn + ENTB_THREADS * ENTB_SLOT_WORDS) + (uint32_t)tid * 16u;

	EBR br;
	br.bp = 64u;
	br.fw = 0u;
	br.nwords = 0u;
	br.src = reinterpret_cast<const uint32_t *>(uslab);
	if (active)
	{
		const uint32_t g = (uint32_t)sc->seg_first + seg;
		br.nwords = (useg_len[g] & ~15u) >> 2;
		br.src = reinterpret_cast<const uint32_t *>(uslab + entb_useg_off(im->data_off, seg_start[g], g));
	}
	ebr_checkpoint<ENTB_RING_WORDS / 4, false>(ring_lane, br);
	asm volatile("cp.async.wait_all;" ::: "memory");
	br.cur = lds_u32(ebr_ring_addr(ring_lane, 0u));
	br.nxt = lds_u32(ebr_ring_addr(ring_lane, 1u));
	br.nx2 = lds_u32(ebr_ring_addr(ring_lane, 2u));
	int dc0 = 0, dc1 = 0, dc2 = 0, dc3 = 0;
	bool failed = false;

	uint32_t *slot_w = &s_dyn[tid * ENTB_SLOT_WORDS];
	const uint32_t slot_addr = (uint32_t)__cvta_generic_to_shared(slot_w);
	const uint32_t lane = (uint32_t)tid & 31u;
	/* byte offset o of the lane's block sits at o ^ slot_key: chunk index ^ (lane & 7), word in chunk ^ (lane >> 3),
	   so the 32 lanes of a warp storing the same coefficient position hit 32 different banks */
	const uint32_t slot_key = ((lane & 7u) << 4) | ((lane >> 3) << 2);
	unsigned long long *s_ptr = reinterpret_cast<unsigned long long *>(s_dyn + ENTB_THREADS * ENTB_SLOT_WORDS + ENTB_RING_BYTES / 4);
	const int wbase = tid & ~31;
#pragma unroll
	for (int i = 0; i < 32; ++i)
		slot_w[i] = 0u;

	int ux = u0 % upr, uy = u0 / upr;
	for (int uu = 0; uu < n_units; ++uu)
	{
		const bool live_u = active && (u0 + uu) < u1;
#pragma unroll 1
		for (int b = 0; b < nblk; ++b)
		{
			const uint32_t ent = sh.blk[b];
			const int s = (int)(ent & 15u);
			const EntbDc *Ldc = &sh.dc[(ent >> 12) & (ND - 1)];
			const EntbAc *Lac = &sh.ac[(ent >> 16) & (NA - 1)];
			const uint32_t dc_fast = (uint32_t)__cvta_generic_to_shared(Ldc->fast);
			const uint32_t ac_fast = (uint32_t)__cvta_generic_to_shared(Lac->fast);
			const uint32_t zq_addr = (uint32_t)__cvta_generic_to_shared(sh.zq[s]);
			const bool live = live_u && !failed;
			bool ok = false;
			if (live)
			{
				int diff;
				uint32_t e;
				ebr_checkpoint<1, true>(ring_lane, br);
				uint32_t zqp = lds_u32(zq_addr); /* pending store: dequantiser entry and value of the previous symbol */
				ok = entb_symbol<ENTB_DC_BITS, true>(ring_lane, br, dc_fast, Ldc, diff, e, [] {});
				if (ok)
				{
					int dcp = s == 0 ? dc0 : s == 1 ? dc1 : s == 2 ? dc2 : dc3;
					dcp += diff;
					if (s == 0) dc0 = dcp;
					else if (s == 1) dc1 = dcp;
					else if (s == 2) dc2 = dcp;
					else dc3 = dcp;
					int valp = dcp;
					/* kp walks the zig-zag table: kp = &zq[k]; the entry of the coefficient is the one before it */
					uint32_t kp = zq_addr + 4u;
					const uint32_t kend = zq_addr + 256u;
					/* two symbols per ring checkpoint (ENTB_CHECK_EVERY), written out so that no counter is needed */
					static_assert(ENTB_CHECK_EVERY == 2, "the loop below is unrolled for two symbols per checkpoint");
					for (;;)
					{
						int val;
						ebr_checkpoint<1, true>(ring_lane, br);
						if (!entb_symbol<ENTB_AC_BITS, false>(ring_lane, br, ac_fast, Lac, val, e, [&] {
								 sts_u16(slot_addr + ((zqp & 0xffffu) ^ slot_key), (uint32_t)(valp * (int)(zqp >> 16)));
							 }))
						{
							ok = false;
							break;
						}
						kp += ENTB_E_KADV(e) << 2;
						zqp = lds_u32_m4(kp);
						valp = val;
						if (kp >= kend)
							break;
						if (!entb_symbol<ENTB_AC_BITS, false>(ring_lane, br, ac_fast, Lac, val, e, [&] {
								 sts_u16(slot_addr + ((zqp & 0xffffu) ^ slot_key), (uint32_t)(valp * (int)(zqp >> 16)));
							 }))
						{
							ok = false;
							break;
						}
						kp += ENTB_E_KADV(e) << 2;
						zqp = lds_u32_m4(kp);
						valp = val;
						if (kp >= kend)
							break;
					}
					if (ok)
					sts_u16(slot_addr + ((zqp & 0xffffu) ^ slot_key), (uint32_t)(valp * (int)(zqp >> 16)));
				}
				failed = !ok;
			}
			/* ---- flush: the warp writes its 32 blocks as whole 128-byte lines (8 lanes per block), which costs the
			   L1 pipe one line per 8 lanes instead of one per lane; the owner publishes its destination first ---- */
			{
				const int bx = ux * sh.hn[s] + (int)((ent >> 4) & 15u), by = uy * sh.vn[s] + (int)((ent >> 8) & 15u);
				const unsigned long long dst =
					(live && ok) ? (unsigned long long)(coef + sh.coff[s] + ((size_t)by * sh.bw[s] + bx) * 64) : 0ull;
				s_ptr[tid] = dst;
			}
			__syncwarp();
#pragma unroll
			for (int i = 0; i < 8; ++i)
			{
				const int o = 4 * i + (int)(lane >> 3); /* owner lane of the block this lane helps to write */
				const uint32_t c = lane & 7u;			  /* natural chunk of that block */
				uint4 *d4 = reinterpret_cast<uint4 *>(s_ptr[wbase + o]);
				uint4 *src = reinterpret_cast<uint4 *>(&s_dyn[(wbase + o) * ENTB_SLOT_WORDS]) + (c ^ (uint32_t)(o & 7));
				const uint4 v = *src;
				*src = make_uint4(0u, 0u, 0u, 0u);
				/* words of the chunk are permuted by (owner lane >> 3) == (i >> 1): a compile-time renaming */
				const uint32_t wv[4] = {v.x, v.y, v.z, v.w};
				const int r = (i >> 1) & 3;
				if (d4)
					d4[c ^ coef_key(d4)] = make_uint4(wv[0 ^ r], wv[1 ^ r], wv[2 ^ r], wv[3 ^ r]); /* slab chunk permutation */
			}
			__syncwarp();
		}
		if (++ux == upr)
		{
			ux = 0;
			++uy;
		}
	}
	asm volatile("cp.async.wait_all;" ::: "memory"); /* no copy may land after the CTA's shared memory is reassigned */
	if (failed)
		report_failure(img_err, sc, seg, JCU_ERR_BAD_HUFFMAN);
	else if (active && u0 < total)
	{
		/* the left-over rule of jpeg_entropy.cuh on the unstuffed copy: R = real bits behind the end position */
		const uint32_t g = (uint32_t)sc->seg_first + seg;
		const uint32_t ul = useg_len[g];
		const int exact = (int)(ul & ~15u) - (int)(ul & 15u), pos = (int)(br.bp - 64u);
		const int R = 8 * exact - pos;
		const bool at_restart = ri > 0 && total % ri == 0;
		const bool inner = seg + 1 < sc->nseg || sc->end_rst; /* an RSTn follows this segment */
		bool trips = false;
		if (inner)
			/* the marker loop runs into the RSTn itself when the reader has not met it: 25 real bits and more behind the
			   end position always; exactly 24 only if all 24 sat in the buffer already (no refill at jpeg.c:1185) */
			trips = R >= 25 || (R == 24 && entb_replay(sh, bitstream + im->data_off, im->data_len, seg_start[g], u1 - u0, true, sc->data_end) == REPLAY_TRIPS);
		else if (R >= 9)
		{
			/* last segment: the first 0xFF data byte (a stuffed FF 00 in the file) wholly behind the end position */
			const uint8_t *ub = reinterpret_cast<const uint8_t *>(br.src);
			int f = -1;
			for (int i = (pos + 7) >> 3; i < exact && f < 0; ++i)
				if (ub[i ^ 3] == 0xffu) /* the slab holds big-endian words */
					f = i;
			if (f >= 0)
			{
				if (8 * f >= pos + 32)
					trips = true; /* beyond anything the reference can have buffered */
				else if (at_restart)
					trips = R >= 25 && 8 * f >= pos + leftover_ahead_bits(R, true);
				else
					trips = entb_replay(sh, bitstream + im->data_off, im->data_len, seg_start[g], u1 - u0, false, sc->data_end) == REPLAY_TRIPS;
			}
		}
		if (trips)
			report_failure(img_err, sc, seg, JCU_ERR_UNKNOWN_MARKER);
	}
}
