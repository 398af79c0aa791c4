/*
 * jpeg_entropy.cuh -- Huffman entropy decode on the device (sm_100a).
 *
 * Replaces, for every scan of every image in a batch at once, the reference's serial
 * stbi__parse_entropy_coded_data (jpeg.c:1155) and its block decoders:
 *   baseline              stbi__jpeg_decode_block          jpeg.c:308
 *   progressive DC        stbi__jpeg_decode_block_prog_dc  jpeg.c:372
 *   progressive AC        stbi__jpeg_decode_block_prog_ac  jpeg.c:406
 * built on a device bit reader with the byte-stuffing rules of stbi__grow_buffer_unsafe
 * (jpeg.c:167) and canonical Huffman decode equivalent to stbi__jpeg_huff_decode (jpeg.c:193).
 *
 * Parallel decomposition (not in the reference, which is strictly serial):
 *   - one CTA works on up to ENT_THREADS restart segments of ONE scan of ONE image, so the
 *     scan's lookup tables are built once per CTA in shared memory from the raw DHT contents;
 *   - one thread decodes one restart segment (stbi__jpeg_reset, jpeg.c:1142, shows the state
 *     that resets at a restart: bit buffer, DC predictors, eob_run -- so segments are
 *     independent).  Streams without restart markers have one segment per scan here; the
 *     self-synchronising decoder (jpeg_selfsync.cuh) splits those further.
 *
 * Output: int16 coefficients, natural (row-major) order, one 128-byte block at a time, DC
 * de-predicted.  Baseline scans store (short)(value * dequant) exactly as the reference does at
 * decode time (jpeg.c:325,345,365); progressive scans store raw values and the output stage
 * applies the 16-bit-wrapping dequantisation of stbi__jpeg_dequantize (jpeg.c:1319).
 *
 * Lookup (this repo's own format, one 32-bit entry per 9-bit prefix):
 *   bits 0-4  bits to consume (0 = code longer than 9 bits -> canonical search)
 *   bits 5-8  run            bits 9-12  magnitude bits still to read
 *   bit 13    value resolved: code + magnitude fitted in 9 bits, bits 16-31 = extended value
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"

#define ENT_THREADS 128
#define ENT_SMALL_THREADS 32 /* CTA size for scans of few segments (no restart markers: one) */
#define ENT_LUT_BITS 9
#define ENT_SLOT_WORDS 36 /* 64 int16 = 32 words, +4 pad words: 16B chunks of the 8 lanes of a phase hit distinct banks */

/* zig-zag index -> natural index, with 16 trailing 63s like the reference's padded table
   (jpeg.c:303-305) so corrupt runs cannot index outside the block */
__constant__ uint8_t c_natural[80] = {0,	1,	 8,  16, 9,	 2,  3,	10, 17, 24, 32, 25, 18, 11, 4,  5,	12, 19, 26, 33,
												  40, 48, 41, 34, 27, 20, 13, 6,	 7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36,
												  29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54,
												  47, 55, 62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

struct HuffLut
{
	uint32_t fast[1 << ENT_LUT_BITS];
	uint32_t limit[18]; /* limit[l] = (first code of length l + count) << (16 - l); [17] = ~0 */
	int32_t base[17];	  /* symbol index = base[l] + code */
	uint8_t vals[256];
};

#define LUT_RESOLVED 0x2000u

/* Build one lookup table from raw DHT contents.  Called by all threads of the CTA. */
__device__ __forceinline__ void hufflut_build(HuffLut *L, const jcu_huff_spec *spec, bool is_dc, int tid, int nthreads)
{
	__shared__ uint32_t s_first[17]; /* first code of each length */
	__shared__ uint32_t s_count[17];
	__shared__ int32_t s_valptr[17];
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
	for (int e = tid; e < (1 << ENT_LUT_BITS); e += nthreads)
	{
		uint32_t entry = 0;
		for (int l = 1; l <= ENT_LUT_BITS; ++l)
		{
			uint32_t code = (uint32_t)e >> (ENT_LUT_BITS - l);
			uint32_t rel = code - s_first[l];
			if (code >= s_first[l] && rel < s_count[l])
			{
				int sym = spec->vals[s_valptr[l] + (int)rel];
				int r = is_dc ? 0 : (sym >> 4);
				int s = is_dc ? (sym > 15 ? 0 : sym) : (sym & 15); /* DC sizes > 15 decode to 0 (jpeg.c:259) */
				if (s && l + s <= ENT_LUT_BITS)
				{
					int mag = (e >> (ENT_LUT_BITS - l - s)) & ((1 << s) - 1);
					if (mag < (1 << (s - 1)))
						mag += (int)(0xffffffffu << s) + 1;
					entry = (uint32_t)(l + s) | ((uint32_t)r << 5) | LUT_RESOLVED | ((uint32_t)mag << 16);
				}
				else
					entry = (uint32_t)l | ((uint32_t)r << 5) | ((uint32_t)s << 9);
				break;
			}
		}
		L->fast[e] = entry;
	}
	__syncthreads();
}

/* MSB-first bit reader over one restart segment */
/* ---- per-image failure word.  The reference stops at the first failing event in file order, so the word keeps
 * the smallest key (scan ordinal, restart segment, code); it is stored complemented so that 0 = no failure and
 * atomicMax keeps the earliest event (JCU_ERR_CODE in jpeg_cuda_abi.h reads the code back). ---- */
__device__ __forceinline__ void report_failure(int32_t *img_err, const jcu_scan *sc, uint32_t seg, uint32_t code)
{
	const uint32_t key = (((min((uint32_t)sc->wave, 1023u) << 20) | min(seg, 0xfffffu)) << 2) | code;
	atomicMax(reinterpret_cast<unsigned int *>(&img_err[sc->image]), ~key);
}

/*
 * "Left-over bytes" rule.  The reference leaves a restart interval through
 *     if (code_bits < 24) grow_buffer(); if (!RESTART(marker)) return 1;        (jpeg.c:1180-1189, 1223-1230, ...)
 * and after a scan looks for the next marker by skipping to the next 0xFF byte and taking the byte behind it
 * (jpeg.c:1722-1735).  A segment of a damaged stream that is decoded before its bytes are used up therefore
 * (a) ends the scan if it is not the last one -- the RSTn marker, or a stuffed FF 00 before it, is then taken
 * for the next marker segment: "unknown marker" -- and (b) if it is the last one, fails the same way when the
 * unread bytes contain a stuffed 0xFF.  grow_buffer fills to 25..32 bits, so with R real bits left behind the
 * end position the marker has been seen iff R <= 24 (R == 24 with exactly 24 bits buffered is the one case
 * that depends on the refill history; it is taken as seen), and the reader has consumed 24 + (R % 8 ? R % 8 : 8)
 * bits beyond the end position.  That is exact wherever the interval ends with the refill above (`at_restart`:
 * every segment but the last, and the last one too when the scan's units are a multiple of the interval).  A
 * scan that simply runs out of units ends without it: the bits the reference then holds (R % 8 plus 0, 8, 16
 * or 24) depend on where its last refill happened, i.e. on the whole history of the segment, which a parallel
 * decoder does not have; the middle value is assumed, so a stuffed 0xFF within two bytes of that point can be
 * judged differently.
 */
__device__ __forceinline__ int leftover_ahead_bits(int R, bool at_restart)
{
	return (at_restart ? 24 : 16) + ((R & 7) ? (R & 7) : 8);
}

/* next raw byte of a segment with the stuffing resolved; false at the marker that ends it (or the end of the file) */
__device__ __forceinline__ bool raw_next_unstuffed(const uint8_t *&q, const uint8_t *end, uint32_t &b)
{
	if (q >= end)
		return false;
	b = *q++;
	if (b == 0xff)
	{
		const uint8_t *r = q;
		uint32_t c = r < end ? *r : 0u;
		while (c == 0xff)
		{
			++r;
			c = r < end ? *r : 0u;
		}
		if (c != 0)
			return false;
		q = r < end ? r + 1 : end;
	}
	return true;
}

/* serial form of the rule for a decoder that only knows where it stopped: pos = unstuffed bits consumed since
   seg_begin.  With i0 = ceil((pos + 25) / 8) -- the first byte the reference has not consumed -- and L unstuffed
   bytes in the segment: nothing happens when L < i0; otherwise a segment that is not the last trips, and the last
   one trips iff a byte in [i0, L) is 0xFF (i0 one byte earlier when the scan does not end at a restart boundary,
   see above).  O(segment) for one thread: callers run it only when L >= i0. */
__device__ __noinline__ bool leftover_trips_serial(const uint8_t *seg_begin, const uint8_t *file_end, long long pos, bool last_segment,
																	bool at_restart)
{
	const long long i0 = ((pos + 25 + 7) >> 3) - ((last_segment && !at_restart) ? 1 : 0);
	const uint8_t *q = seg_begin;
	uint32_t b;
	for (long long i = 0; i < i0; ++i)
		if (!raw_next_unstuffed(q, file_end, b))
			return false;
	if (!last_segment)
		return true;
	while (raw_next_unstuffed(q, file_end, b))
		if (b == 0xffu)
			return true;
	return false;
}

struct BitReader
{
	const uint8_t *p;
	const uint8_t *end;
	uint64_t buf; /* valid bits are the top nbits */
	int nbits;
	int stopped; /* a marker was reached: feed zero bits from now on (jpeg.c:171-183) */
	int pad;		 /* zero bits appended behind the stream so far (they sit at the bottom of the valid bits) */
};

__device__ __forceinline__ void br_init(BitReader &br, const uint8_t *p, const uint8_t *end)
{
	br.p = p;
	br.end = end;
	br.buf = 0;
	br.nbits = 0;
	br.stopped = 0;
	br.pad = 0;
}

/* one byte with the reference's stuffing rules: FF 00 -> FF, FF FF.. fill, FF xx -> stop */
__device__ __noinline__ void br_refill_slow(BitReader &br)
{
	do
	{
		uint32_t b = 0;
		bool real = false;
		if (!br.stopped && br.p < br.end)
		{
			real = true;
			b = *br.p++;
			if (b == 0xff)
			{
				const uint8_t *q = br.p;
				uint32_t c = q < br.end ? *q : 0u;
				while (c == 0xff)
				{
					++q;
					c = q < br.end ? *q : 0u;
				}
				if (c == 0)
					br.p = q < br.end ? q + 1 : br.end;
				else
				{
					br.stopped = 1;
					b = 0;
					real = false;
				}
			}
		}
		if (!real)
			br.pad += 8;
		br.buf |= (uint64_t)b << (56 - br.nbits);
		br.nbits += 8;
	} while (br.nbits <= 56 && (br.stopped || br.p >= br.end || ((uintptr_t)br.p & 3) != 0));
}

/* top up to more than 32 valid bits: aligned 32-bit loads while no byte of the word is 0xFF */
__device__ __forceinline__ void br_refill(BitReader &br)
{
	while (br.nbits <= 32)
	{
		if (!br.stopped && ((uintptr_t)br.p & 3) == 0 && br.p + 4 <= br.end)
		{
			const uint32_t w = *reinterpret_cast<const uint32_t *>(br.p);
			const uint32_t ff = ((w & 0x7f7f7f7fu) + 0x01010101u) & w & 0x80808080u;
			if (ff == 0)
			{
				br.buf |= (uint64_t)__byte_perm(w, 0, 0x0123) << (32 - br.nbits);
				br.nbits += 32;
				br.p += 4;
				continue;
			}
		}
		br_refill_slow(br);
	}
}

/* the left-over rule (above) for a segment that this reader has just finished; true = the reference fails here */
__device__ __noinline__ bool br_reference_trips(const BitReader &br, bool last_segment, bool at_restart)
{
	const int real = br.nbits - br.pad; /* stream bits still buffered */
	const uint8_t *q = br.p;
	uint32_t b;
	int R = real;
	bool more = !br.stopped;
	while (more && R < 25)
	{
		more = raw_next_unstuffed(q, br.end, b);
		if (more)
			R += 8;
	}
	if (R < 25)
		return false;
	if (!last_segment)
		return true;
	/* last segment: is there a (stuffed) 0xFF in what the reference has not consumed? */
	const int a = leftover_ahead_bits(real, at_restart);
	for (int o = a; o + 8 <= real; o += 8)
		if (((br.buf >> (56 - o)) & 0xffu) == 0xffu)
			return true;
	if (br.stopped)
		return false;
	q = br.p;
	for (int skip = a - real; skip > 0; skip -= 8)
		if (!raw_next_unstuffed(q, br.end, b))
			return false;
	while (raw_next_unstuffed(q, br.end, b))
		if (b == 0xffu)
			return true;
	return false;
}

/* distance in bits from the reader's position to the first 0xFF data byte that lies wholly behind it (a stuffed
   FF 00 in the file), -1 if the segment has none */
__device__ __noinline__ int br_first_ff_ahead(const BitReader &br)
{
	const int real = br.nbits - br.pad;
	if (real < 0)
		return -1;
	for (int o = real & 7; o + 8 <= real; o += 8)
		if (((br.buf >> (56 - o)) & 0xffu) == 0xffu)
			return o;
	if (br.stopped)
		return -1;
	const uint8_t *q = br.p;
	uint32_t b;
	for (int o = real; raw_next_unstuffed(q, br.end, b); o += 8)
	{
		if (b == 0xffu)
			return o;
		if (o > (1 << 30))
			break;
	}
	return -1;
}

__device__ __forceinline__ uint32_t br_peek(const BitReader &br, int n) /* 1 <= n <= 32 */
{
	return (uint32_t)(br.buf >> (64 - n));
}
__device__ __forceinline__ void br_skip(BitReader &br, int n)
{
	br.buf <<= n;
	br.nbits -= n;
}
__device__ __forceinline__ uint32_t br_bits(BitReader &br, int n) /* 0 <= n <= 16 */
{
	if (n == 0)
		return 0;
	if (br.nbits < n)
		br_refill(br);
	uint32_t v = br_peek(br, n);
	br_skip(br, n);
	return v;
}

/* canonical search for codes longer than ENT_LUT_BITS; returns the symbol or -1 ("bad huffman code") */
__device__ __noinline__ int huff_decode_long(BitReader &br, const HuffLut *L)
{
	uint32_t top = br_peek(br, 16);
	int l = ENT_LUT_BITS + 1;
	while (top >= L->limit[l])
		++l;
	if (l > 16)
	{
		br_skip(br, 16);
		return -1;
	}
	int idx = L->base[l] + (int)(top >> (16 - l));
	br_skip(br, l);
	return (int)L->vals[idx & 255];
}

/* plain symbol decode (progressive paths) */
__device__ __forceinline__ int huff_decode(BitReader &br, const HuffLut *L)
{
	if (br.nbits <= 32)
		br_refill(br);
	uint32_t e = L->fast[br_peek(br, ENT_LUT_BITS)];
	if ((e & 31u) == 0)
		return huff_decode_long(br, L);
	if (e & LUT_RESOLVED)
	{
		/* the table folded the magnitude bits in: undo that, callers want the bare symbol */
		int s = 0;
		/* code length = total - s, where s is recoverable from the value's magnitude */
		int v = (int)e >> 16;
		int a = v < 0 ? -v : v;
		while (a)
		{
			++s;
			a >>= 1;
		}
		br_skip(br, (int)(e & 31u) - s);
		return (int)(((e >> 5) & 15u) << 4) | s;
	}
	br_skip(br, (int)(e & 31u));
	return (int)((((e >> 5) & 15u) << 4) | ((e >> 9) & 15u));
}

/* receive n bits and sign-extend JPEG-style (stbi__extend_receive, jpeg.c:250) */
__device__ __forceinline__ int receive_extend(BitReader &br, int n)
{
	if (n == 0 || n > 16)
		return 0;
	int v = (int)br_bits(br, n);
	if (v < (1 << (n - 1)))
		v += (int)(0xffffffffu << n) + 1;
	return v;
}

struct ScanShared
{
	HuffLut lut[4];	/* per scan-component slot: a progressive scan uses DC tables or AC tables, never both */
	int lut_of_dc[4]; /* slot -> lut index (slots sharing a spec share a table) */
	int lut_of_ac[4];
	uint8_t zz[80];
};

/* one CTA work item */
struct EntWork
{
	int32_t scan;
	uint32_t seg_first; /* index (within the scan) of this CTA's first segment */
	uint32_t nseg;
};

/* ---- progressive block decoders operating directly on the coefficient slab ---- */

__device__ __forceinline__ int prog_ac_first(BitReader &br, const HuffLut *L, const uint8_t *zz, int16_t *blk, int ss,
															int se, int al, int &eob_run)
{
	if (eob_run)
	{
		--eob_run;
		return 1;
	}
	int k = ss;
	do
	{
		int rs = huff_decode(br, L);
		if (rs < 0)
			return 0;
		int s = rs & 15, r = rs >> 4;
		if (s == 0)
		{
			if (r < 15)
			{
				eob_run = 1 << r;
				if (r)
					eob_run += (int)br_bits(br, r);
				--eob_run;
				break;
			}
			k += 16;
		}
		else
		{
			k += r;
			int zig = zz[k < 79 ? k : 79];
			++k;
			blk[coef_pos(zig, coef_key(blk))] = (int16_t)(receive_extend(br, s) * (1 << al));
		}
	} while (k <= se);
	return 1;
}

__device__ __forceinline__ void refine_one(int16_t *p, int bit)
{
	int v = *p;
	if ((v & bit) == 0)
		*p = (int16_t)(v > 0 ? v + bit : v - bit);
}

__device__ __forceinline__ int prog_ac_refine(BitReader &br, const HuffLut *L, const uint8_t *zz, int16_t *blk, int ss,
															 int se, int al, int &eob_run)
{
	const int bit = (int)(int16_t)(1 << al);
	if (eob_run)
	{
		--eob_run;
		for (int k = ss; k <= se; ++k)
		{
			int16_t *p = &blk[coef_pos(zz[k], coef_key(blk))];
			if (*p != 0)
				if (br_bits(br, 1))
					refine_one(p, bit);
		}
		return 1;
	}
	int k = ss;
	do
	{
		int rs = huff_decode(br, L);
		if (rs < 0)
			return 0;
		int s = rs & 15, r = rs >> 4;
		if (s == 0)
		{
			if (r < 15)
			{
				eob_run = (1 << r) - 1;
				if (r)
					eob_run += (int)br_bits(br, r);
				r = 64; /* force end of block */
			}
		}
		else
		{
			if (s != 1)
				return 0;
			s = br_bits(br, 1) ? bit : -bit;
		}
		while (k <= se)
		{
			int16_t *p = &blk[coef_pos(zz[k++], coef_key(blk))];
			if (*p != 0)
			{
				if (br_bits(br, 1))
					refine_one(p, bit);
			}
			else
			{
				if (r == 0)
				{
					*p = (int16_t)s;
					break;
				}
				--r;
			}
		}
	} while (k <= se);
	return 1;
}

/*
 * Progressive entropy decode (the three decoders of jpeg.c:372-558), in place on the coefficient slab:
 * grid = CTA work items of one wave (scans of one dependency level across the batch).  Thread t of a CTA
 * decodes restart segment (work.seg_first + t) of the CTA's scan.  T = segments per CTA: scans without
 * restart markers are one segment each, so they get 32-thread CTAs (a decoder is a serial, latency-bound
 * thread: what counts is how many scans are resident, and a CTA slot per scan is the limit).
 * Baseline scans never come here (jpeg_entropy_baseline.cuh, jpeg_selfsync.cuh).
 */
template <int T>
__global__ void __launch_bounds__(T)
k_entropy_segments(const EntWork *__restrict__ work, const jcu_scan *__restrict__ scans,
						 const jcu_image *__restrict__ images, const uint32_t *__restrict__ seg_start,
						 const jcu_huff_spec *__restrict__ huff, const uint8_t *__restrict__ bitstream,
						 int16_t *__restrict__ coef, int32_t *__restrict__ img_err)
{
	__shared__ ScanShared sh;

	const int tid = threadIdx.x;
	const EntWork w = work[blockIdx.x];
	const jcu_scan *sc = &scans[w.scan];
	const jcu_image *im = &images[sc->image];
	const int kind = sc->kind, ncs = sc->ncomp;

	/* ---- lookup tables, once per CTA ---- */
	for (int i = tid; i < 80; i += T)
		sh.zz[i] = c_natural[i];
	for (int i = 0; i < ncs; ++i)
	{
		if (sc->dc_spec[i] >= 0)
		{
			int alias = -1;
			for (int j = 0; j < i; ++j)
				if (sc->dc_spec[j] == sc->dc_spec[i])
					alias = sh.lut_of_dc[j];
			if (alias < 0)
			{
				hufflut_build(&sh.lut[i], &huff[sc->dc_spec[i]], true, tid, T);
				alias = i;
			}
			if (tid == 0)
				sh.lut_of_dc[i] = alias;
		}
		if (sc->ac_spec[i] >= 0)
		{
			int alias = -1;
			for (int j = 0; j < i; ++j)
				if (sc->ac_spec[j] == sc->ac_spec[i])
					alias = sh.lut_of_ac[j];
			if (alias < 0)
			{
				hufflut_build(&sh.lut[i], &huff[sc->ac_spec[i]], false, tid, T);
				alias = i;
			}
			if (tid == 0)
				sh.lut_of_ac[i] = alias;
		}
		__syncthreads();
	}

	/* ---- this thread's segment ---- */
	const bool active = (uint32_t)tid < w.nseg;
	const uint32_t seg = w.seg_first + (uint32_t)tid;
	const int ri = sc->restart_interval, total = sc->total_units, upr = sc->units_per_row;
	const int n_units = ri ? ri : total;
	const int u0 = ri ? (int)seg * ri : 0;
	const int u1 = ri ? min(u0 + ri, total) : total;

	BitReader br;
	{
		const uint8_t *file = bitstream + im->data_off;
		uint32_t off = active ? seg_start[sc->seg_first + seg] : 0u;
		br_init(br, file + off, file + im->data_len);
	}
	int dc_pred[4] = {0, 0, 0, 0};
	int eob_run = 0;
	bool failed = false;

	const int ss = sc->ss, se = sc->se, al = sc->al;

	for (int uu = 0; uu < n_units; ++uu)
	{
		const int u = u0 + uu;
		const bool live = active && u < u1 && !failed;
		const int ux = u % upr, uy = u / upr;
#pragma unroll 1
		for (int i = 0; i < ncs; ++i)
		{
			const jcu_comp *cp = &im->comp[sc->comp[i]];
			const int bh_n = ncs == 1 ? 1 : cp->v, bw_n = ncs == 1 ? 1 : cp->h;
			const HuffLut *Ldc = &sh.lut[sh.lut_of_dc[i] & 3];
			const HuffLut *Lac = &sh.lut[sh.lut_of_ac[i] & 3];
			int dcp = dc_pred[i];
			for (int y = 0; y < bh_n; ++y)
				for (int x = 0; x < bw_n; ++x)
				{
					const int bx = ux * bw_n + x, by = uy * bh_n + y;
					int16_t *blk = coef + cp->coef_off + ((size_t)by * cp->bw + bx) * 64;
					if (live && !failed)
					{
						int ok = 1;
						if (kind == JCU_SCAN_PROG_DC_FIRST)
						{
							uint4 *d4 = reinterpret_cast<uint4 *>(blk);
#pragma unroll
							for (int q = 0; q < 8; ++q)
								d4[q] = make_uint4(0u, 0u, 0u, 0u);
							int t = huff_decode(br, Ldc);
							if (t < 0)
								ok = 0;
							else
							{
								dcp += receive_extend(br, t);
								blk[coef_pos(0, coef_key(blk))] = (int16_t)(dcp * (1 << al));
							}
						}
						else if (kind == JCU_SCAN_PROG_DC_REFINE)
						{
							if (br_bits(br, 1))
							{
								int16_t *p0 = &blk[coef_pos(0, coef_key(blk))];
								*p0 = (int16_t)(*p0 + (int16_t)(1 << al));
							}
						}
						else if (kind == JCU_SCAN_PROG_AC_FIRST)
							ok = prog_ac_first(br, Lac, sh.zz, blk, ss, se, al, eob_run);
						else
							ok = prog_ac_refine(br, Lac, sh.zz, blk, ss, se, al, eob_run);
						if (!ok)
							failed = true;
					}
				}
			dc_pred[i] = dcp;
		}
	}
	if (failed)
		report_failure(img_err, sc, seg, JCU_ERR_BAD_HUFFMAN);
	else if (active && u0 < total && br_reference_trips(br, seg + 1 >= sc->nseg && !sc->end_rst, ri > 0 && total % ri == 0))
		report_failure(img_err, sc, seg, JCU_ERR_UNKNOWN_MARKER);
}
