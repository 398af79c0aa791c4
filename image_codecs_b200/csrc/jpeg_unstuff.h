/*
 * jpeg_unstuff.h -- byte-stuffing removal for one restart segment, word at a time.
 *
 * The rules are those of stbi__grow_buffer_unsafe (jpeg.c:167-187): FF 00 -> FF; FF FF.. are fill bytes
 * before a marker or a stuffed 00; FF xx ends the segment (marker) and nothing after it is data; reading
 * past the end of the file yields 0 (common.c:39).  Compiles for the device (k_unstuff in
 * jpeg_entropy_baseline.cuh) and for the host (tests/test_unstuff_core.py checks unstuff_word against the
 * plain byte loop on random streams).
 */
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define UNSTUFF_HD __host__ __device__ __forceinline__
#else
#define UNSTUFF_HD static inline
#endif

/* compaction table: index = drop mask of the 4 stream bytes of a word (bit i = byte i is dropped);
   low 16 bits = byte selector that left-aligns the survivors in stream order (big-endian), count << 16 */
#define UNSTUFF_LUT_INIT                                                                                                     \
	{                                                                                                                         \
		0x40123u, 0x31234u, 0x30234u, 0x22344u, 0x30134u, 0x21344u, 0x20344u, 0x13444u, 0x30124u, 0x21244u, 0x20244u,       \
			0x12444u, 0x20144u, 0x11444u, 0x10444u, 0x04444u                                                                   \
	}
#if defined(__CUDACC__)
__constant__ uint32_t c_unstuff_lut[16] = UNSTUFF_LUT_INIT;
#endif
static const uint32_t h_unstuff_lut[16] = UNSTUFF_LUT_INIT;

/* PRMT with a zero second operand: nibble j of sel picks the source byte of result byte j (4 = zero) */
UNSTUFF_HD uint32_t unstuff_perm(uint32_t w, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
	return __byte_perm(w, 0u, sel);
#else
	uint32_t r = 0;
	for (int j = 0; j < 4; ++j)
	{
		const uint32_t src = (sel >> (4 * j)) & 7u;
		const uint32_t byte = src < 4u ? (w >> (8u * src)) & 0xffu : 0u;
		r |= byte << (8 * j);
	}
	return r;
#endif
}
UNSTUFF_HD uint32_t unstuff_shl(uint32_t x, uint32_t n) { return n >= 32u ? 0u : x << n; }

struct UnstuffOut
{
	uint32_t *wp;	 /* next output word (device: next 16-byte unit) */
	uint32_t pend;	 /* pending bytes, left aligned */
	uint32_t nacc;	 /* 0..3 */
	uint32_t words; /* words produced */
#if defined(__CUDACC__)
	uint32_t stage;  /* device: shared-memory address of this lane's staging column, rows `stride` bytes apart */
	uint32_t stride;
#endif
};
/* one output word.  On the device four words are staged in shared memory and leave as one 16-byte store:
   scattered 4-byte stores cost the L2 one transaction each, four times as many. */
UNSTUFF_HD void unstuff_put(UnstuffOut &o, uint32_t w)
{
#if defined(__CUDA_ARCH__)
	asm volatile("st.shared.u32 [%0], %1;" ::"r"(o.stage + (o.words & 3u) * o.stride), "r"(w) : "memory");
	++o.words;
	if ((o.words & 3u) == 0u)
	{
		uint4 v;
		asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v.x) : "r"(o.stage));
		asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v.y) : "r"(o.stage + o.stride));
		asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v.z) : "r"(o.stage + 2u * o.stride));
		asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v.w) : "r"(o.stage + 3u * o.stride));
		*reinterpret_cast<uint4 *>(o.wp) = v;
		o.wp += 4;
	}
#else
	*o.wp++ = w;
	++o.words;
#endif
}
UNSTUFF_HD void unstuff_emit(UnstuffOut &o, uint32_t b)
{
	o.pend |= b << (24u - 8u * o.nacc);
	if (++o.nacc == 4u)
	{
		unstuff_put(o, o.pend);
		o.pend = 0u;
		o.nacc = 0u;
	}
}
/* append cnt (0..4) bytes held left-aligned in be */
UNSTUFF_HD void unstuff_append(UnstuffOut &o, uint32_t be, uint32_t cnt)
{
	const uint32_t sh = 8u * o.nacc;
	const uint32_t merged = o.pend | (be >> sh);
	const uint32_t rest = unstuff_shl(be, 32u - sh);
	const uint32_t total = o.nacc + cnt;
	if (total >= 4u)
	{
		unstuff_put(o, merged);
		o.pend = rest;
		o.nacc = total - 4u;
	}
	else
	{
		o.pend = merged;
		o.nacc = total;
	}
}
/* one byte with the rules above; false = marker or end of file */
UNSTUFF_HD bool unstuff_byte(const uint8_t *base, uint32_t &pos, uint32_t len, UnstuffOut &o)
{
	if (pos >= len)
		return false;
	const uint32_t b = base[pos++];
	if (b == 0xffu)
	{
		uint32_t q = pos;
		uint32_t c = q < len ? base[q] : 0u;
		while (c == 0xffu)
		{
			++q;
			c = q < len ? base[q] : 0u;
		}
		if (c != 0u)
		{
			pos = len;
			return false; /* marker: nothing more from this segment */
		}
		pos = q < len ? q + 1u : len;
	}
	unstuff_emit(o, b);
	return true;
}

/*
 * The aligned raw word wd at file offset wpos (wpos < len), of which the bytes below `pos` are not ours
 * (before the segment, or a stuffed 00 already accounted for); n0w = the word after it (only its first
 * byte matters, and only when wpos + 4 < len).  A word whose 0xFF bytes are all followed by 0x00 is
 * compacted without a branch: zero-byte mask & shifted FF mask -> 4-bit drop mask -> byte selector.  An
 * 0xFF followed by anything else (a marker, i.e. the end of the segment, or fill bytes) hands over to the
 * byte loop at that 0xFF.  Returns false when the segment is finished; else pos is word aligned again, or
 * wpos + 5 when the next word starts with the 00 of a stuffed pair.
 */
UNSTUFF_HD bool unstuff_word(const uint8_t *base, uint32_t len, uint32_t wpos, uint32_t wd, uint32_t n0w, uint32_t &pos,
									  UnstuffOut &o)
{
	const uint32_t d = pos - wpos; /* 0..3 */
	const uint32_t e = len - wpos; /* >= 1 */
	if (e < 4u)
		wd &= (1u << (8u * e)) - 1u; /* past the end of the file reads 0 */
	const uint32_t igm = 0x80808080u & ((1u << (8u * d)) - 1u);
	const uint32_t tm = e < 4u ? 0x80808080u & ~((1u << (8u * e)) - 1u) : 0u;
	uint32_t ffm = ((wd & 0x7f7f7f7fu) + 0x01010101u) & wd & 0x80808080u & ~igm;
	const uint32_t zm = ~(((wd & 0x7f7f7f7fu) + 0x7f7f7f7fu) | wd) & 0x80808080u;
	const bool n0z = wpos + 4u >= len || (n0w & 0xffu) == 0u;
	const uint32_t znext = (zm >> 8) | (n0z ? 0x80000000u : 0u);
	const uint32_t bad = ffm & ~znext; /* 0xFF followed by something other than 0x00 */
	uint32_t cut = 0u, i0 = 4u;
	if (bad)
	{
		uint32_t t = bad;
		i0 = 0u;
		while (!(t & 0xffu))
		{
			t >>= 8;
			++i0;
		}
		cut = 0x80808080u & ~((1u << (8u * i0)) - 1u);
		ffm &= ~cut;
	}
	const uint32_t drop = igm | tm | cut | (zm & (ffm << 8));
	const uint32_t idx = (((drop >> 7) * 0x01020408u) >> 24) & 15u;
#if defined(__CUDA_ARCH__)
	const uint32_t ent = c_unstuff_lut[idx];
#else
	const uint32_t ent = h_unstuff_lut[idx];
#endif
	unstuff_append(o, unstuff_perm(wd, ent & 0xffffu), ent >> 16);
	if (bad)
	{
		bool more;
		/* the byte after that 0xFF is at hand (it lies inside the file, else it would have read as 0x00): anything
		   but another 0xFF makes it a marker and the segment ends here, which is how every segment of a valid
		   file ends -- no byte loop, no dependent loads */
		const uint32_t after = i0 < 3u ? (wd >> (8u * (i0 + 1u))) & 0xffu : n0w & 0xffu;
		if (after != 0xffu)
		{
			pos = len;
			return false;
		}
		pos = wpos + i0;
		do
			more = unstuff_byte(base, pos, len, o);
		while (more && (pos & 3u) != 0u);
		return more;
	}
	const bool carry = (ffm >> 31) != 0u && wpos + 4u < len; /* then the next byte is the stuffed 0x00: skip it */
	pos = wpos + 4u + (carry ? 1u : 0u);
	return pos < len;
}

/* Scalar statement of what one lane of k_unstuff does for a segment starting at `start` (test harness only;
   the kernel runs the same word steps with its lanes in lock step).  Returns the number of output words
   before padding; *out must hold (len - start) / 4 + 2 words. */
UNSTUFF_HD uint32_t unstuff_segment_scalar(const uint8_t *base, uint32_t len, uint32_t start, uint32_t *out)
{
	UnstuffOut o;
	uint32_t pos = start;
	bool more = pos < len;
	o.wp = out;
	o.pend = 0u;
	o.nacc = 0u;
	o.words = 0u;
	while (more)
	{
		const uint32_t wpos = pos & ~3u;
		uint32_t wd = 0u, n0w = 0u;
		for (uint32_t i = 0; i < 4u; ++i) /* the kernel reads whole padded units; bytes past len are masked either way */
		{
			wd |= (uint32_t)base[wpos + i] << (8u * i);
			n0w |= (uint32_t)base[wpos + 4u + i] << (8u * i);
		}
		const uint32_t ffq = ((wd & 0x7f7f7f7fu) + 0x01010101u) & wd & 0x80808080u;
		if (pos == wpos && wpos + 4u <= len && ffq == 0u)
		{
			unstuff_append(o, unstuff_perm(wd, 0x0123u), 4u);
			pos = wpos + 4u;
			more = pos < len; /* the kernel finds this out at the next word (e = 0 never happens: it checks pos < len too) */
		}
		else
			more = unstuff_word(base, len, wpos, wd, n0w, pos, o);
	}
	if (o.nacc)
		unstuff_put(o, o.pend);
	return o.words;
}
