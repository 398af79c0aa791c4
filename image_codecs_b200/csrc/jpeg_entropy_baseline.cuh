/*
 * jpeg_entropy_baseline.cuh -- baseline (SOF0/1) Huffman entropy decode, lock-step edition (sm_100a).
 *
 * Replaces stbi__jpeg_decode_block (jpeg.c:308-370) driven by stbi__parse_entropy_coded_data
 * (jpeg.c:1155-1234) for every baseline scan of a batch in one launch.  The progressive block
 * decoders stay in jpeg_entropy.cuh.
 *
 * Decomposition: one thread = one restart segment (stbi__jpeg_reset, jpeg.c:1142, clears every
 * carried state at a restart marker), ENTB_THREADS segments of ONE scan per CTA so the scan's
 * tables are built once in shared memory.  All lanes of a warp walk the same (MCU, block) position
 * of their own segment, so the per-block control flow is warp-uniform; the only divergent loop is
 * the AC symbol loop, whose body is a single straight-line path:
 *
 *   entry  = lut[top 9 bits]            one 32-bit entry: code length | magnitude bits | k advance
 *   value  = extend(next `size` bits)   branch-free form of stbi__extend_receive (jpeg.c:250)
 *   slot[natural(k + run)] = (short)(value * q)      dequantised at decode time like jpeg.c:325,345,365
 *   k     += run + 1   (EOB: += 64, ZRL: += 16)
 *
 * EOB/ZRL need no branch: they carry size 0, so they store a zero at a position that has not been
 * written yet (coefficients are produced in increasing zig-zag order).  Codes longer than 9 bits
 * take a canonical search (same result as jpeg.c:217-241).  The bit buffer holds >= 32 valid bits
 * at every symbol start (a code is <= 16 bits and its magnitude <= 15), so one symbol works on
 * the top 32-bit word only.  Refill is word-at-a-time with the next word already in a register;
 * a word containing 0xFF (stuffing, fill bytes, markers: jpeg.c:167-187) drops to a byte loop.
 *
 * Output: int16 coefficients, natural order, 128-byte blocks, written once per block from a
 * zeroed shared-memory slot (the memset of jpeg.c:320).
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_entropy.cuh"

#define ENTB_THREADS 128
#define ENTB_SLOT_WORDS 36 /* 64 int16 + 4 pad words (bank spread of the 16-byte flush reads) */
#define ENTB_MAX_MCU_BLOCKS 64

#define ENTB_DC_BITS 9
#define ENTB_AC_BITS 11

/* 16-bit lookup entry: len (5) | size (4) << 5 | kadv (7) << 9 ; 0 = code longer than the table's prefix */
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

struct EntbShared
{
	EntbDc dc[4];					 /* per scan-component slot */
	EntbAc ac[4];
	uint32_t zq[4][128];			 /* per slot: natural index | quantiser << 16, indexed by zig-zag position;
											 positions 64..127 alias natural 63 like the reference's padded table (jpeg.c:303) */
	uint32_t blk[ENTB_MAX_MCU_BLOCKS]; /* per block of an MCU: slot | dx << 4 | dy << 8 | dc lut << 12 | ac lut << 16 */
	int nblk;
	unsigned long long coff[4]; /* per slot: first coefficient of the component in the slab */
	int bw[4], hn[4], vn[4];	 /* per slot: blocks per plane row; blocks per MCU across / down (1 x 1 in single-component scans) */
};

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
	return (uint32_t)l | ((uint32_t)size << 5) | ((uint32_t)kadv << 9);
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


/* ---- bit reader: 64-bit MSB-aligned buffer over one file (base pointer is CTA-uniform),
 *      next aligned word already in a register ---- */
struct EBR
{
	uint32_t pos, len; /* byte offset in the file; file length */
	uint64_t buf;
	int nbits;
	uint32_t w; /* word at pos if pos is 4-aligned and pos + 4 <= len, else 0xffffffff (forces the byte path) */
};

__device__ __forceinline__ uint32_t ebr_prefetch(const uint8_t *__restrict__ base, uint32_t pos, uint32_t len)
{
	return (((pos & 3u) == 0u) && pos + 4u <= len) ? __ldg(reinterpret_cast<const uint32_t *>(base + pos)) : 0xffffffffu;
}

/* byte path with the stuffing rules of stbi__grow_buffer_unsafe (jpeg.c:167-187): FF 00 -> FF, FF FF.. fill,
   FF xx -> marker: stop consuming (pos = len) and feed zero bits; past the end of the file reads 0 (common.c:39).
   Loops until at least 32 bits are valid and the reader is back on an aligned word (or out of data). */
__device__ __forceinline__ void ebr_slow(const uint8_t *__restrict__ base, EBR &br)
{
	do
	{
		uint32_t b = 0;
		if (br.pos < br.len)
		{
			b = base[br.pos++];
			if (b == 0xff)
			{
				uint32_t q = br.pos;
				uint32_t c = q < br.len ? base[q] : 0u;
				while (c == 0xff)
				{
					++q;
					c = q < br.len ? base[q] : 0u;
				}
				if (c == 0)
					br.pos = q < br.len ? q + 1 : br.len;
				else
				{
					br.pos = br.len;
					b = 0;
				}
			}
		}
		br.buf |= (uint64_t)b << (56 - br.nbits);
		br.nbits += 8;
	} while (br.nbits <= 56 && (br.nbits < 32 || br.pos >= br.len || (br.pos & 3u) != 0u));
	br.w = ebr_prefetch(base, br.pos, br.len);
}

__device__ __forceinline__ uint32_t shr_clamp(uint32_t x, uint32_t n) /* n >= 32 gives 0 (PTX shr semantics) */
{
	uint32_t d;
	asm("shr.u32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(n));
	return d;
}

/*
 * One Huffman symbol plus its magnitude bits.  Returns false for "bad huffman code" (jpeg.c:317,351).
 * value = extended magnitude (0 when the symbol carries no bits); kadv = k advance (AC tables).
 * The refill is predicated, not branched: nearly every lock-step iteration has SOME lane short of
 * bits, so a branch would serialise a 3-lane refill against the rest of the warp every time.
 */
template <int BITS, bool IS_DC>
__device__ __forceinline__ bool entb_symbol(const uint8_t *__restrict__ base, EBR &br, const EntbLutT<BITS> *__restrict__ L,
														  int &value, int &kadv)
{
	{
		const bool need = br.nbits < 32;
		const uint32_t w = br.w;
		const uint32_t ff = ((w & 0x7f7f7f7fu) + 0x01010101u) & w & 0x80808080u;
		if (need && ff)
			ebr_slow(base, br); /* leaves >= 32 bits */
		else
		{
			const uint64_t ins = (uint64_t)__byte_perm(w, 0, 0x0123) << ((32 - br.nbits) & 63);
			br.buf |= need ? ins : 0ull;
			br.nbits += need ? 32 : 0;
			br.pos += need ? 4u : 0u;
			if (need)
				br.w = ebr_prefetch(base, br.pos, br.len);
		}
	}
	const uint32_t top = (uint32_t)(br.buf >> 32);
	uint32_t e = L->fast[top >> (32 - BITS)];
	if (e == 0)
	{
		const uint32_t t16 = top >> 16;
		int l = BITS + 1;
		while (t16 >= L->limit[l])
			++l;
		if (l > 16)
			return false;
		e = entb_entry(L->vals[(L->base[l] + (int)(t16 >> (16 - l))) & 255], l, IS_DC);
	}
	const int len = (int)(e & 31u), size = (int)((e >> 5) & 15u);
	kadv = (int)(e >> 9);
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

/* per-scan tables, built once per CTA by all its threads (slots sharing a DHT table share one lookup) */
__device__ __forceinline__ void entb_setup(EntbShared &sh, const jcu_scan *sc, const jcu_image *im,
														 const jcu_huff_spec *__restrict__ huff, int tid, int nthreads)
{
	const int ncs = sc->ncomp;
	int lut_dc[4], lut_ac[4]; /* indices resolved after unrolling: no dynamic indexing */
#pragma unroll
	for (int i = 0; i < 4; ++i)
	{
		lut_dc[i] = i;
		lut_ac[i] = i;
		if (i < ncs)
		{
#pragma unroll
			for (int j = 0; j < 4; ++j)
				if (j < i)
				{
					if (sc->dc_spec[j] == sc->dc_spec[i] && lut_dc[i] == i)
						lut_dc[i] = lut_dc[j];
					if (sc->ac_spec[j] == sc->ac_spec[i] && lut_ac[i] == i)
						lut_ac[i] = lut_ac[j];
				}
			if (lut_dc[i] == i && sc->dc_spec[i] >= 0)
				entb_lut_build(&sh.dc[i], &huff[sc->dc_spec[i]], true, tid, nthreads);
			if (lut_ac[i] == i && sc->ac_spec[i] >= 0)
				entb_lut_build(&sh.ac[i], &huff[sc->ac_spec[i]], false, tid, nthreads);
		}
	}
	for (int i = tid; i < ncs * 128; i += nthreads)
	{
		const int s = i >> 7, z = i & 127;
		const int nat = c_natural[z < 79 ? z : 79];
		sh.zq[s][z] = (uint32_t)nat | ((uint32_t)im->qt[sc->comp[s]][nat] << 16);
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

/*
 * grid = CTA work items (EntWork) of one wave restricted to baseline scans; thread t decodes restart
 * segment (work.seg_first + t) of the CTA's scan.
 */
__global__ void __launch_bounds__(ENTB_THREADS)
k_entropy_baseline(const EntWork *__restrict__ work, const jcu_scan *__restrict__ scans,
						 const jcu_image *__restrict__ images, const uint32_t *__restrict__ seg_start,
						 const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ bitstream,
						 int16_t *__restrict__ coef, int32_t *__restrict__ img_err)
{
	__shared__ EntbShared sh;
	__shared__ __align__(16) uint32_t s_slots[ENTB_THREADS * ENTB_SLOT_WORDS];

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
	const uint8_t *__restrict__ base = bitstream + im->data_off;

	EBR br;
	br.len = im->data_len;
	br.pos = active ? seg_start[sc->seg_first + seg] : br.len;
	br.buf = 0;
	br.nbits = 0;
	br.w = ebr_prefetch(base, br.pos, br.len);
	int dc0 = 0, dc1 = 0, dc2 = 0, dc3 = 0;
	bool failed = false;

	uint32_t *slot_w = &s_slots[tid * ENTB_SLOT_WORDS];
	int16_t *slot = reinterpret_cast<int16_t *>(slot_w);
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
			const EntbDc *Ldc = &sh.dc[(ent >> 12) & 3u];
			const EntbAc *Lac = &sh.ac[(ent >> 16) & 3u];
			const uint32_t *zq = sh.zq[s];
			const bool live = live_u && !failed;
			bool ok = false;
			if (live)
			{
				int diff, kadv;
				ok = entb_symbol<ENTB_DC_BITS, true>(base, br, Ldc, diff, kadv);
				if (ok)
				{
					int dcp = s == 0 ? dc0 : s == 1 ? dc1 : s == 2 ? dc2 : dc3;
					dcp += diff;
					if (s == 0) dc0 = dcp;
					else if (s == 1) dc1 = dcp;
					else if (s == 2) dc2 = dcp;
					else dc3 = dcp;
					slot[0] = (int16_t)(dcp * (int)(zq[0] >> 16));
					int k = 1;
					do
					{
						int val;
						if (!entb_symbol<ENTB_AC_BITS, false>(base, br, Lac, val, kadv))
						{
							ok = false;
							break;
						}
						const uint32_t e = zq[k + kadv - 1];
						k += kadv;
						slot[e & 0xffffu] = (int16_t)(val * (int)(e >> 16));
					} while (k < 64);
				}
				failed = !ok;
			}
			__syncwarp();
			if (live)
			{
				const int bx = ux * sh.hn[s] + (int)((ent >> 4) & 15u), by = uy * sh.vn[s] + (int)((ent >> 8) & 15u);
				uint4 *d4 = reinterpret_cast<uint4 *>(coef + sh.coff[s] + ((size_t)by * sh.bw[s] + bx) * 64);
				uint4 *s4 = reinterpret_cast<uint4 *>(slot_w);
				const unsigned key = coef_key(d4);
#pragma unroll
				for (int q = 0; q < 8; ++q)
				{
					const uint4 v = s4[q];
					if (ok)
						d4[q ^ key] = v; /* slab chunk permutation (jpeg_coef_layout.cuh) */
					s4[q] = make_uint4(0u, 0u, 0u, 0u);
				}
			}
		}
		if (++ux == upr)
		{
			ux = 0;
			++uy;
		}
	}
	if (failed)
		atomicOr(&img_err[sc->image], 1);
}
