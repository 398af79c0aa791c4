/*
 * tools/jpeg_testenc.c -- test/bench tooling: a from-scratch JPEG *encoder* used only to
 * generate input streams the reference's codec/jpeg_write.c cannot emit (it has no DRI/RSTn,
 * no 1-component output, no 4:4:4 at q<=90, no progressive -- jpeg_write.c:220-221,247-250;
 * SURVEY.md F5).  Not part of the accelerated path, not an oracle.
 *
 * Follows ITU-T T.81: Annex A (DCT, zig-zag), Annex B (markers), Annex F (sequential Huffman),
 * Annex G (progressive: spectral selection + successive approximation), Annex K.2 (optimal
 * table generation), Annex K.3 (typical tables).
 *
 *   features: baseline SOF0 / progressive SOF2; 1, 3 or 4 components; any H,V in {1,2,4};
 *             restart intervals; typical (Annex K) or optimised Huffman tables (two-pass);
 *             JFIF YCbCr, Adobe RGB / CMYK / YCCK colour signalling; 8- or 16-bit DQT;
 *             libjpeg-style or custom progressive scan scripts; multi-scan (non-interleaved)
 *             baseline.
 *
 * C interface (see tools/testenc.py for the ctypes binding):
 *   int  tje_encode(const tje_params *p, const unsigned char *pixels, unsigned char **out, size_t *out_len);
 *   void tje_free(void *p);
 *   int  tje_encode_many(...)   -- thread pool over a list of images (bench input generation)
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define TJE_API __attribute__((visibility("default")))

enum
{
	TJE_CS_YCBCR = 0, /* 3 comps, JFIF APP0, ids 1,2,3; RGB input converted */
	TJE_CS_GRAY = 1,	/* 1 comp; input 1 channel (or 3 -> luma) */
	TJE_CS_RGB = 2,	/* 3 comps, ids 'R','G','B', no transform, Adobe APP14 transform=0 */
	TJE_CS_CMYK = 3,	/* 4 comps, Adobe APP14 transform=0; input 4 channels stored as is */
	TJE_CS_YCCK = 4,	/* 4 comps, Adobe APP14 transform=2; input CMYK -> YCC on CMY, K kept */
	TJE_CS_YCBCR_NOJFIF = 5 /* 3 comps ids 1,2,3, no APP0/APP14 */
};

typedef struct
{
	int width, height;
	int in_channels; /* channels per input pixel */
	int colorspace;  /* TJE_CS_* */
	int quality;	  /* 1..100, IJG scaling */
	int h[4], v[4];  /* sampling factors per component (0 -> 1) */
	int restart_interval; /* in MCUs, 0 = none */
	int progressive; /* 0 baseline, 1 = libjpeg default script, 2 = spectral selection only, 3 = custom script */
	int optimize;	  /* 1 = two-pass optimal Huffman tables (forced for progressive != 0 when eob runs are wanted) */
	int dqt16;		  /* 1 = write 16-bit quant tables */
	int noninterleaved; /* baseline only: one scan per component */
	int n_script;	  /* custom script entries (progressive == 3) */
	int script[32][6]; /* {comp_mask, Ss, Se, Ah, Al, unused} */
	int dri_after_tables; /* 1 = emit DRI after SOF (default before) -- marker-order variety */
	int pad_fill_bytes; /* extra 0xFF fill bytes before each RSTn / EOI (legal per B.1.1.2) */
} tje_params;

/* ------------------------------------------------------------------------------------ */
static const uint8_t k_zigzag[64] = {0,	 1,  8,  16, 9,  2,	3,	 10, 17, 24, 32, 25, 18, 11, 4,  5,	12, 19, 26, 33, 40, 48,
												 41, 34, 27, 20, 13, 6,	7,	 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
												 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

/* Annex K.1 quantisation tables (natural order) */
static const uint8_t k_qlum[64] = {16, 11, 10, 16, 24,  40,	51,  61,	 12, 12, 14, 19, 26,	 58,	60,  55,
											  14, 13, 16, 24, 40,  57,	69,  56,	 14, 17, 22, 29, 51,	 87,	80,  62,
											  18, 22, 37, 56, 68,  109, 103, 77,	 24, 35, 55, 64, 81,	 104, 113, 92,
											  49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const uint8_t k_qchr[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
											  99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
											  99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};

/* Annex K.3 typical Huffman tables: BITS[16] + HUFFVAL */
static const uint8_t k_dc_lum_bits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const uint8_t k_dc_chr_bits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const uint8_t k_dc_vals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t k_ac_lum_bits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const uint8_t k_ac_chr_bits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const uint8_t k_ac_lum_vals[162] = {
	 0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71,
	 0x14, 0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72,
	 0x82, 0x09, 0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37,
	 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59,
	 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83,
	 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3,
	 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3,
	 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2,
	 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
static const uint8_t k_ac_chr_vals[162] = {
	 0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22,
	 0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1,
	 0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36,
	 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58,
	 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a,
	 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a,
	 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba,
	 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda,
	 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};

/* ------------------------------------------------------------------------------------ */
typedef struct
{
	uint8_t *data;
	size_t size, cap;
	uint32_t acc; /* bit accumulator, MSB-first */
	int nacc;
} bytesink;

static void bs_reserve(bytesink *b, size_t extra)
{
	if (b->size + extra > b->cap)
	{
		size_t nc = b->cap ? b->cap * 2 : (1u << 16);
		while (nc < b->size + extra)
			nc *= 2;
		b->data = (uint8_t *)realloc(b->data, nc);
		b->cap = nc;
	}
}
static void bs_byte(bytesink *b, int v)
{
	bs_reserve(b, 1);
	b->data[b->size++] = (uint8_t)v;
}
static void bs_u16(bytesink *b, int v)
{
	bs_byte(b, v >> 8);
	bs_byte(b, v & 255);
}
static void bs_bytes(bytesink *b, const void *p, size_t n)
{
	bs_reserve(b, n);
	memcpy(b->data + b->size, p, n);
	b->size += n;
}
static void bs_bits(bytesink *b, unsigned code, int n)
{
	if (n == 0)
		return;
	b->acc = (b->acc << n) | (code & ((1u << n) - 1));
	b->nacc += n;
	while (b->nacc >= 8)
	{
		int byte = (b->acc >> (b->nacc - 8)) & 255;
		bs_byte(b, byte);
		if (byte == 255)
			bs_byte(b, 0);
		b->nacc -= 8;
	}
}
static void bs_flush_ones(bytesink *b)
{
	if (b->nacc)
		bs_bits(b, (1u << (8 - b->nacc)) - 1, 8 - b->nacc);
	b->acc = 0;
	b->nacc = 0;
}

/* ------------------------------------------------------------------------------------ */
typedef struct
{
	uint8_t bits[16];
	uint8_t vals[256];
	int nvals;
	uint16_t code[256];
	uint8_t len[256];
	long freq[257];
} hufftab;

static void huff_derive(hufftab *t)
{
	int k = 0, l, i;
	unsigned code = 0;
	memset(t->len, 0, sizeof(t->len));
	for (l = 1; l <= 16; ++l)
	{
		for (i = 0; i < t->bits[l - 1]; ++i, ++k)
		{
			t->code[t->vals[k]] = (uint16_t)code++;
			t->len[t->vals[k]] = (uint8_t)l;
		}
		code <<= 1;
	}
	t->nvals = k;
}
static void huff_set(hufftab *t, const uint8_t *bits, const uint8_t *vals, int n)
{
	memcpy(t->bits, bits, 16);
	memcpy(t->vals, vals, (size_t)n);
	huff_derive(t);
}

/* Annex K.2: code sizes from frequencies, limited to 16 bits; symbol 256 reserved so no all-ones code */
static void huff_optimal(hufftab *t)
{
	long freq[257];
	int codesize[257], others[257], bits[33];
	int i, j, c1, c2, p;
	memcpy(freq, t->freq, sizeof(freq));
	freq[256] = 1;
	memset(codesize, 0, sizeof(codesize));
	memset(bits, 0, sizeof(bits));
	for (i = 0; i < 257; ++i)
		others[i] = -1;
	for (;;)
	{
		long v;
		c1 = -1;
		v = 1000000000L;
		for (i = 0; i <= 256; ++i)
			if (freq[i] && freq[i] <= v)
			{
				v = freq[i];
				c1 = i;
			}
		c2 = -1;
		v = 1000000000L;
		for (i = 0; i <= 256; ++i)
			if (freq[i] && freq[i] <= v && i != c1)
			{
				v = freq[i];
				c2 = i;
			}
		if (c2 < 0)
			break;
		freq[c1] += freq[c2];
		freq[c2] = 0;
		codesize[c1]++;
		while (others[c1] >= 0)
		{
			c1 = others[c1];
			codesize[c1]++;
		}
		others[c1] = c2;
		codesize[c2]++;
		while (others[c2] >= 0)
		{
			c2 = others[c2];
			codesize[c2]++;
		}
	}
	for (i = 0; i <= 256; ++i)
		if (codesize[i])
			bits[codesize[i] > 32 ? 32 : codesize[i]]++;
	for (i = 32; i > 16; --i)
		while (bits[i] > 0)
		{
			j = i - 2;
			while (bits[j] == 0)
				j--;
			bits[i] -= 2;
			bits[i - 1]++;
			bits[j + 1] += 2;
			bits[j]--;
		}
	while (bits[i] == 0)
		i--;
	bits[i]--; /* remove the reserved symbol */
	for (i = 0; i < 16; ++i)
		t->bits[i] = (uint8_t)bits[i + 1];
	p = 0;
	for (i = 1; i <= 32; ++i)
		for (j = 0; j <= 255; ++j)
			if (codesize[j] == i)
				t->vals[p++] = (uint8_t)j;
	huff_derive(t);
}

/* ------------------------------------------------------------------------------------ */
typedef struct
{
	int h, v, tq, id;
	int bw, bh;		/* blocks incl. MCU padding */
	int cbw, cbh;	/* blocks covering the real component size (non-interleaved scans) */
	int16_t *coef; /* bw*bh*64, zig-zag order, quantised */
	int td, ta;
} tcomp;

typedef struct
{
	const tje_params *p;
	int ncomp, hmax, vmax, mcux, mcuy;
	tcomp comp[4];
	uint16_t qt[2][64]; /* natural order */
	hufftab dc[2], ac[2];
	bytesink out;
	/* scan state */
	int gather; /* 1 = statistics pass */
	int Ss, Se, Ah, Al;
	int last_dc[4];
	unsigned eobrun;
	int ac_tbl;
	uint8_t *corr; /* buffered correction bits */
	int ncorr;
	int restarts_left, next_rst;
} encoder;

static int nbits_of(int v)
{
	int n = 0;
	if (v < 0)
		v = -v;
	while (v)
	{
		n++;
		v >>= 1;
	}
	return n;
}

static void emit_sym(encoder *e, hufftab *t, int sym)
{
	if (e->gather)
		t->freq[sym]++;
	else
		bs_bits(&e->out, t->code[sym], t->len[sym]);
}
static void emit_raw(encoder *e, unsigned v, int n)
{
	if (!e->gather)
		bs_bits(&e->out, v, n);
}
static void emit_corr(encoder *e, const uint8_t *bits, int n)
{
	int i;
	if (e->gather)
		return;
	for (i = 0; i < n; ++i)
		bs_bits(&e->out, bits[i], 1);
}
static void flush_eobrun(encoder *e)
{
	if (e->eobrun)
	{
		int nb = nbits_of((int)e->eobrun) - 1;
		emit_sym(e, &e->ac[e->ac_tbl], nb << 4);
		if (nb)
			emit_raw(e, e->eobrun & ((1u << nb) - 1), nb);
		e->eobrun = 0;
		emit_corr(e, e->corr, e->ncorr);
		e->ncorr = 0;
	}
}

/* ---- block coders ---- */
static void enc_block_baseline(encoder *e, const int16_t *zz, int ci)
{
	tcomp *c = &e->comp[ci];
	int diff = zz[0] - e->last_dc[ci], nb, k, r = 0, v;
	e->last_dc[ci] = zz[0];
	v = diff;
	if (v < 0)
		v--;
	nb = nbits_of(diff);
	emit_sym(e, &e->dc[c->td], nb);
	emit_raw(e, (unsigned)v, nb);
	for (k = 1; k < 64; ++k)
	{
		int a = zz[k];
		if (a == 0)
		{
			r++;
			continue;
		}
		while (r > 15)
		{
			emit_sym(e, &e->ac[c->ta], 0xF0);
			r -= 16;
		}
		v = a;
		if (v < 0)
			v--;
		nb = nbits_of(a);
		emit_sym(e, &e->ac[c->ta], (r << 4) | nb);
		emit_raw(e, (unsigned)v, nb);
		r = 0;
	}
	if (r)
		emit_sym(e, &e->ac[c->ta], 0);
}

static void enc_block_dc_first(encoder *e, const int16_t *zz, int ci)
{
	tcomp *c = &e->comp[ci];
	int val = zz[0] >> e->Al; /* arithmetic shift = point transform for DC (G.1.2.1) */
	int diff = val - e->last_dc[ci], nb = nbits_of(diff), v = diff;
	e->last_dc[ci] = val;
	if (v < 0)
		v--;
	emit_sym(e, &e->dc[c->td], nb);
	emit_raw(e, (unsigned)v, nb);
}
static void enc_block_dc_refine(encoder *e, const int16_t *zz)
{
	emit_raw(e, (unsigned)(zz[0] >> e->Al) & 1u, 1);
}
static void enc_block_ac_first(encoder *e, const int16_t *zz)
{
	int k, r = 0;
	hufftab *t = &e->ac[e->ac_tbl];
	for (k = e->Ss; k <= e->Se; ++k)
	{
		int a = zz[k], neg = a < 0, nb, v;
		if (neg)
			a = -a;
		a >>= e->Al; /* AC point transform divides magnitude (G.1.2.2) */
		if (a == 0)
		{
			r++;
			continue;
		}
		flush_eobrun(e);
		while (r > 15)
		{
			emit_sym(e, t, 0xF0);
			r -= 16;
		}
		nb = nbits_of(a);
		v = neg ? ~a : a;
		emit_sym(e, t, (r << 4) | nb);
		emit_raw(e, (unsigned)v, nb);
		r = 0;
	}
	if (r > 0)
	{
		e->eobrun++;
		if (e->eobrun == 0x7FFF)
			flush_eobrun(e);
	}
}
static void enc_block_ac_refine(encoder *e, const int16_t *zz)
{
	int k, r = 0, eob = 0, nbr = 0;
	int absv[64];
	uint8_t *br = e->corr + e->ncorr; /* this block's correction bits follow any carried ones */
	hufftab *t = &e->ac[e->ac_tbl];
	for (k = e->Ss; k <= e->Se; ++k)
	{
		int a = zz[k];
		if (a < 0)
			a = -a;
		a >>= e->Al;
		absv[k] = a;
		if (a == 1)
			eob = k;
	}
	for (k = e->Ss; k <= e->Se; ++k)
	{
		int a = absv[k];
		if (a == 0)
		{
			r++;
			continue;
		}
		while (r > 15 && k <= eob)
		{
			flush_eobrun(e);
			emit_sym(e, t, 0xF0);
			r -= 16;
			emit_corr(e, br, nbr);
			br = e->corr; /* eobrun flush reset ncorr to 0 */
			nbr = 0;
		}
		if (a > 1)
		{
			br[nbr++] = (uint8_t)(a & 1);
			continue;
		}
		flush_eobrun(e);
		emit_sym(e, t, (r << 4) | 1);
		emit_raw(e, zz[k] < 0 ? 0u : 1u, 1);
		emit_corr(e, br, nbr);
		br = e->corr;
		nbr = 0;
		r = 0;
	}
	if (r > 0 || nbr > 0)
	{
		e->eobrun++;
		/* keep this block's bits buffered behind the pending EOB run */
		if (br != e->corr + e->ncorr)
			memmove(e->corr + e->ncorr, br, (size_t)nbr);
		e->ncorr += nbr;
		if (e->eobrun == 0x7FFF || e->ncorr > 1000 - 64 + 1)
			flush_eobrun(e);
	}
}

static void scan_restart(encoder *e)
{
	int i;
	flush_eobrun(e);
	if (!e->gather)
	{
		bs_flush_ones(&e->out);
		for (i = 0; i < e->p->pad_fill_bytes; ++i)
			bs_byte(&e->out, 0xFF);
		bs_byte(&e->out, 0xFF);
		bs_byte(&e->out, 0xD0 + e->next_rst);
	}
	e->next_rst = (e->next_rst + 1) & 7;
	memset(e->last_dc, 0, sizeof(e->last_dc));
}

static void enc_one_block(encoder *e, int ci, int bx, int by)
{
	tcomp *c = &e->comp[ci];
	const int16_t *zz = c->coef + 64 * ((size_t)by * c->bw + bx);
	if (!e->p->progressive)
		enc_block_baseline(e, zz, ci);
	else if (e->Ss == 0)
	{
		if (e->Ah == 0)
			enc_block_dc_first(e, zz, ci);
		else
			enc_block_dc_refine(e, zz);
	}
	else
	{
		if (e->Ah == 0)
			enc_block_ac_first(e, zz);
		else
			enc_block_ac_refine(e, zz);
	}
}

/* entropy-code one scan over the listed components */
static void encode_scan_data(encoder *e, const int *cis, int ncs)
{
	int mx, my, k, x, y, ri = e->p->restart_interval, todo = ri;
	memset(e->last_dc, 0, sizeof(e->last_dc));
	e->eobrun = 0;
	e->ncorr = 0;
	e->next_rst = 0;
	if (ncs == 1)
	{
		tcomp *c = &e->comp[cis[0]];
		long nblk = (long)c->cbw * c->cbh, b = 0;
		e->ac_tbl = c->ta;
		for (my = 0; my < c->cbh; ++my)
			for (mx = 0; mx < c->cbw; ++mx)
			{
				enc_one_block(e, cis[0], mx, my);
				++b;
				if (ri && --todo == 0 && b < nblk)
				{
					scan_restart(e);
					todo = ri;
				}
			}
	}
	else
	{
		long nm = (long)e->mcux * e->mcuy, m = 0;
		for (my = 0; my < e->mcuy; ++my)
			for (mx = 0; mx < e->mcux; ++mx)
			{
				for (k = 0; k < ncs; ++k)
				{
					tcomp *c = &e->comp[cis[k]];
					for (y = 0; y < c->v; ++y)
						for (x = 0; x < c->h; ++x)
							enc_one_block(e, cis[k], mx * c->h + x, my * c->v + y);
				}
				++m;
				if (ri && --todo == 0 && m < nm)
				{
					scan_restart(e);
					todo = ri;
				}
			}
	}
	flush_eobrun(e);
	if (!e->gather)
	{
		bs_flush_ones(&e->out);
	}
}

/* ---- markers ---- */
static void write_dqt(encoder *e, int t)
{
	int i, p16 = e->p->dqt16;
	bs_u16(&e->out, 0xFFDB);
	bs_u16(&e->out, 2 + 1 + (p16 ? 128 : 64));
	bs_byte(&e->out, (p16 << 4) | t);
	for (i = 0; i < 64; ++i)
	{
		int q = e->qt[t][k_zigzag[i]];
		if (p16)
			bs_u16(&e->out, q);
		else
			bs_byte(&e->out, q);
	}
}
static void write_dht(encoder *e, int cls, int id)
{
	hufftab *t = cls ? &e->ac[id] : &e->dc[id];
	bs_u16(&e->out, 0xFFC4);
	bs_u16(&e->out, 2 + 1 + 16 + t->nvals);
	bs_byte(&e->out, (cls << 4) | id);
	bs_bytes(&e->out, t->bits, 16);
	bs_bytes(&e->out, t->vals, (size_t)t->nvals);
}
static void write_sos(encoder *e, const int *cis, int ncs)
{
	int i;
	bs_u16(&e->out, 0xFFDA);
	bs_u16(&e->out, 6 + 2 * ncs);
	bs_byte(&e->out, ncs);
	for (i = 0; i < ncs; ++i)
	{
		bs_byte(&e->out, e->comp[cis[i]].id);
		bs_byte(&e->out, (e->comp[cis[i]].td << 4) | e->comp[cis[i]].ta);
	}
	bs_byte(&e->out, e->Ss);
	bs_byte(&e->out, e->Se);
	bs_byte(&e->out, (e->Ah << 4) | e->Al);
}

/* one scan: optional stats pass + table emission, then SOS + data */
static void do_scan(encoder *e, const int *cis, int ncs, int Ss, int Se, int Ah, int Al)
{
	int i, need_dc, need_ac, used_dc[2] = {0, 0}, used_ac[2] = {0, 0};
	e->Ss = Ss;
	e->Se = Se;
	e->Ah = Ah;
	e->Al = Al;
	need_dc = (!e->p->progressive) || (Ss == 0 && Ah == 0);
	need_ac = (!e->p->progressive) || (Ss > 0);
	for (i = 0; i < ncs; ++i)
	{
		used_dc[e->comp[cis[i]].td] = 1;
		used_ac[e->comp[cis[i]].ta] = 1;
	}
	if (e->p->optimize)
	{
		for (i = 0; i < 2; ++i)
		{
			memset(e->dc[i].freq, 0, sizeof(e->dc[i].freq));
			memset(e->ac[i].freq, 0, sizeof(e->ac[i].freq));
		}
		e->gather = 1;
		encode_scan_data(e, cis, ncs);
		e->gather = 0;
		for (i = 0; i < 2; ++i)
		{
			if (need_dc && used_dc[i])
			{
				huff_optimal(&e->dc[i]);
				write_dht(e, 0, i);
			}
			if (need_ac && used_ac[i])
			{
				huff_optimal(&e->ac[i]);
				write_dht(e, 1, i);
			}
		}
	}
	write_sos(e, cis, ncs);
	encode_scan_data(e, cis, ncs);
}

/* ---- forward DCT (separable, double precision basis; speed is irrelevant to correctness) ---- */
static float g_dct[8][8];
static pthread_once_t g_dct_once = PTHREAD_ONCE_INIT;
static void dct_init(void)
{
	int u, x;
	for (u = 0; u < 8; ++u)
		for (x = 0; x < 8; ++x)
			g_dct[u][x] = (float)((u == 0 ? sqrt(0.125) : 0.5) * cos((2 * x + 1) * u * M_PI / 16.0));
}
static void fdct_quant(const float *px, int stride, const uint16_t *q, int16_t *zz_out)
{
	float tmp[64], res[64];
	int u, v, x, y;
	for (y = 0; y < 8; ++y)
		for (u = 0; u < 8; ++u)
		{
			float s = 0;
			for (x = 0; x < 8; ++x)
				s += px[y * stride + x] * g_dct[u][x];
			tmp[y * 8 + u] = s;
		}
	for (v = 0; v < 8; ++v)
		for (u = 0; u < 8; ++u)
		{
			float s = 0;
			for (y = 0; y < 8; ++y)
				s += tmp[y * 8 + u] * g_dct[v][y];
			res[v * 8 + u] = s;
		}
	for (x = 0; x < 64; ++x)
	{
		int nat = k_zigzag[x];
		float f = res[nat] / (float)q[nat];
		int iv = (int)(f < 0 ? f - 0.5f : f + 0.5f);
		if (iv > 1023 * 16) iv = 1023 * 16;
		if (iv < -1023 * 16) iv = -1023 * 16;
		zz_out[x] = (int16_t)iv;
	}
}

static int clampi(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }

/* build full-resolution component planes (float, level shifted) from the input pixels */
static void make_planes(encoder *e, const unsigned char *px, float **full)
{
	const tje_params *p = e->p;
	int w = p->width, h = p->height, nc = p->in_channels, x, y, c;
	for (y = 0; y < h; ++y)
		for (x = 0; x < w; ++x)
		{
			const unsigned char *s = px + ((size_t)y * w + x) * nc;
			size_t o = (size_t)y * w + x;
			float r = s[0], g = nc > 1 ? s[1] : s[0], b = nc > 2 ? s[2] : s[0];
			switch (p->colorspace)
			{
			case TJE_CS_GRAY:
				full[0][o] = (nc >= 3 ? 0.299f * r + 0.587f * g + 0.114f * b : r) - 128.f;
				break;
			case TJE_CS_RGB:
				full[0][o] = r - 128.f;
				full[1][o] = g - 128.f;
				full[2][o] = b - 128.f;
				break;
			case TJE_CS_CMYK:
				for (c = 0; c < 4; ++c)
					full[c][o] = (float)s[c < nc ? c : nc - 1] - 128.f;
				break;
			case TJE_CS_YCCK:
				full[0][o] = 0.299f * r + 0.587f * g + 0.114f * b - 128.f;
				full[1][o] = -0.168736f * r - 0.331264f * g + 0.5f * b;
				full[2][o] = 0.5f * r - 0.418688f * g - 0.081312f * b;
				full[3][o] = (float)s[nc > 3 ? 3 : nc - 1] - 128.f;
				break;
			default:
				full[0][o] = 0.299f * r + 0.587f * g + 0.114f * b - 128.f;
				full[1][o] = -0.168736f * r - 0.331264f * g + 0.5f * b;
				full[2][o] = 0.5f * r - 0.418688f * g - 0.081312f * b;
				break;
			}
		}
}

/* downsample + pad + DCT + quantise one component */
static void make_coefs(encoder *e, int ci, const float *full)
{
	tcomp *c = &e->comp[ci];
	int w = e->p->width, h = e->p->height;
	int hs = e->hmax / c->h, vs = e->vmax / c->v;
	int pw = c->bw * 8, ph = c->bh * 8, x, y, dx, dy, bx, by;
	float *pl = (float *)malloc(sizeof(float) * (size_t)pw * ph);
	float inv = 1.f / (float)(hs * vs);
	for (y = 0; y < ph; ++y)
		for (x = 0; x < pw; ++x)
		{
			float s = 0;
			for (dy = 0; dy < vs; ++dy)
				for (dx = 0; dx < hs; ++dx)
				{
					int sx = clampi(x * hs + dx, 0, w - 1), sy = clampi(y * vs + dy, 0, h - 1);
					s += full[(size_t)sy * w + sx];
				}
			pl[(size_t)y * pw + x] = s * inv;
		}
	c->coef = (int16_t *)malloc(sizeof(int16_t) * 64 * (size_t)c->bw * c->bh);
	for (by = 0; by < c->bh; ++by)
		for (bx = 0; bx < c->bw; ++bx)
			fdct_quant(pl + (size_t)by * 8 * pw + bx * 8, pw, e->qt[c->tq], c->coef + 64 * ((size_t)by * c->bw + bx));
	free(pl);
}

TJE_API void tje_free(void *p) { free(p); }

TJE_API int tje_encode(const tje_params *p_in, const unsigned char *pixels, unsigned char **out, size_t *out_len)
{
	encoder E, *e = &E;
	tje_params P = *p_in;
	const tje_params *p = &P;
	int i, ci, q, all[4] = {0, 1, 2, 3};
	float *full[4] = {0, 0, 0, 0};
	memset(e, 0, sizeof(*e));
	if (P.progressive)
		P.optimize = 1; /* the Annex K tables have no EOBn (n>0) symbols */
	e->p = p;
	pthread_once(&g_dct_once, dct_init);
	if (p->width < 1 || p->height < 1 || p->width > 65535 || p->height > 65535)
		return 0;
	switch (p->colorspace)
	{
	case TJE_CS_GRAY: e->ncomp = 1; break;
	case TJE_CS_CMYK:
	case TJE_CS_YCCK: e->ncomp = 4; break;
	default: e->ncomp = 3; break;
	}
	e->hmax = e->vmax = 1;
	for (i = 0; i < e->ncomp; ++i)
	{
		tcomp *c = &e->comp[i];
		c->h = p->h[i] ? p->h[i] : 1;
		c->v = p->v[i] ? p->v[i] : 1;
		if (c->h > e->hmax) e->hmax = c->h;
		if (c->v > e->vmax) e->vmax = c->v;
		c->id = (p->colorspace == TJE_CS_RGB) ? "RGB"[i] : i + 1;
		{
			int chroma = (p->colorspace == TJE_CS_YCBCR || p->colorspace == TJE_CS_YCBCR_NOJFIF || p->colorspace == TJE_CS_YCCK) && (i == 1 || i == 2);
			c->tq = chroma ? 1 : 0;
			c->td = c->ta = chroma ? 1 : 0;
		}
	}
	e->mcux = (p->width + 8 * e->hmax - 1) / (8 * e->hmax);
	e->mcuy = (p->height + 8 * e->vmax - 1) / (8 * e->vmax);
	for (i = 0; i < e->ncomp; ++i)
	{
		tcomp *c = &e->comp[i];
		int cx = (p->width * c->h + e->hmax - 1) / e->hmax, cy = (p->height * c->v + e->vmax - 1) / e->vmax;
		c->bw = e->mcux * c->h;
		c->bh = e->mcuy * c->v;
		c->cbw = (cx + 7) / 8;
		c->cbh = (cy + 7) / 8;
	}
	/* quant tables, IJG quality scaling */
	q = clampi(p->quality ? p->quality : 90, 1, 100);
	q = q < 50 ? 5000 / q : 200 - 2 * q;
	for (i = 0; i < 64; ++i)
	{
		int hi = p->dqt16 ? 32767 : 255;
		e->qt[0][i] = (uint16_t)clampi((k_qlum[i] * q + 50) / 100, 1, hi);
		e->qt[1][i] = (uint16_t)clampi((k_qchr[i] * q + 50) / 100, 1, hi);
	}
	huff_set(&e->dc[0], k_dc_lum_bits, k_dc_vals, 12);
	huff_set(&e->dc[1], k_dc_chr_bits, k_dc_vals, 12);
	huff_set(&e->ac[0], k_ac_lum_bits, k_ac_lum_vals, 162);
	huff_set(&e->ac[1], k_ac_chr_bits, k_ac_chr_vals, 162);
	e->corr = (uint8_t *)malloc(65536);

	for (i = 0; i < e->ncomp; ++i)
		full[i] = (float *)malloc(sizeof(float) * (size_t)p->width * p->height);
	make_planes(e, pixels, full);
	for (i = 0; i < e->ncomp; ++i)
	{
		make_coefs(e, i, full[i]);
		free(full[i]);
	}

	/* ---- headers ---- */
	bs_u16(&e->out, 0xFFD8);
	if (p->colorspace == TJE_CS_YCBCR || p->colorspace == TJE_CS_GRAY)
	{
		static const uint8_t jfif[] = {0xFF, 0xE0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
		bs_bytes(&e->out, jfif, sizeof(jfif));
	}
	if (p->colorspace == TJE_CS_RGB || p->colorspace == TJE_CS_CMYK || p->colorspace == TJE_CS_YCCK)
	{
		uint8_t adobe[] = {0xFF, 0xEE, 0, 14, 'A', 'd', 'o', 'b', 'e', 0, 100, 0, 0, 0, 0, 0};
		adobe[15] = (uint8_t)(p->colorspace == TJE_CS_YCCK ? 2 : 0);
		bs_bytes(&e->out, adobe, sizeof(adobe));
	}
	{
		static const char com[] = "image-codecs_b200 test stream";
		bs_u16(&e->out, 0xFFFE);
		bs_u16(&e->out, 2 + (int)sizeof(com) - 1);
		bs_bytes(&e->out, com, sizeof(com) - 1);
	}
	write_dqt(e, 0);
	if (e->ncomp >= 3 && p->colorspace != TJE_CS_RGB && p->colorspace != TJE_CS_CMYK)
		write_dqt(e, 1);
	if (p->restart_interval && !p->dri_after_tables)
	{
		bs_u16(&e->out, 0xFFDD);
		bs_u16(&e->out, 4);
		bs_u16(&e->out, p->restart_interval);
	}
	bs_u16(&e->out, p->progressive ? 0xFFC2 : 0xFFC0);
	bs_u16(&e->out, 8 + 3 * e->ncomp);
	bs_byte(&e->out, 8);
	bs_u16(&e->out, p->height);
	bs_u16(&e->out, p->width);
	bs_byte(&e->out, e->ncomp);
	for (i = 0; i < e->ncomp; ++i)
	{
		bs_byte(&e->out, e->comp[i].id);
		bs_byte(&e->out, (e->comp[i].h << 4) | e->comp[i].v);
		bs_byte(&e->out, e->comp[i].tq);
	}
	if (!p->optimize)
	{
		int two = 0;
		for (i = 0; i < e->ncomp; ++i)
			if (e->comp[i].td)
				two = 1;
		write_dht(e, 0, 0);
		write_dht(e, 1, 0);
		if (two)
		{
			write_dht(e, 0, 1);
			write_dht(e, 1, 1);
		}
	}
	if (p->restart_interval && p->dri_after_tables)
	{
		bs_u16(&e->out, 0xFFDD);
		bs_u16(&e->out, 4);
		bs_u16(&e->out, p->restart_interval);
	}

	/* ---- scans ---- */
	if (!p->progressive)
	{
		if (p->noninterleaved && e->ncomp > 1)
			for (ci = 0; ci < e->ncomp; ++ci)
				do_scan(e, &ci, 1, 0, 63, 0, 0);
		else
			do_scan(e, all, e->ncomp, 0, 63, 0, 0);
	}
	else if (p->progressive == 3)
	{
		for (i = 0; i < p->n_script; ++i)
		{
			int cis[4], n = 0;
			for (ci = 0; ci < e->ncomp; ++ci)
				if (p->script[i][0] & (1 << ci))
					cis[n++] = ci;
			do_scan(e, cis, n, p->script[i][1], p->script[i][2], p->script[i][3], p->script[i][4]);
		}
	}
	else if (p->progressive == 2)
	{
		/* spectral selection only */
		do_scan(e, all, e->ncomp, 0, 0, 0, 0);
		for (ci = 0; ci < e->ncomp; ++ci)
		{
			do_scan(e, &ci, 1, 1, 5, 0, 0);
			do_scan(e, &ci, 1, 6, 63, 0, 0);
		}
	}
	else
	{
		/* libjpeg's default script shape: DC Al=1; low/high AC bands; successive approximation */
		int y0 = 0;
		do_scan(e, all, e->ncomp, 0, 0, 0, 1);
		if (e->ncomp >= 3)
		{
			int c1 = 1, c2 = 2;
			do_scan(e, &y0, 1, 1, 5, 0, 2);
			do_scan(e, &c2, 1, 1, 63, 0, 1);
			do_scan(e, &c1, 1, 1, 63, 0, 1);
			do_scan(e, &y0, 1, 6, 63, 0, 2);
			do_scan(e, &y0, 1, 1, 63, 2, 1);
			do_scan(e, all, e->ncomp, 0, 0, 1, 0);
			do_scan(e, &c2, 1, 1, 63, 1, 0);
			do_scan(e, &c1, 1, 1, 63, 1, 0);
			do_scan(e, &y0, 1, 1, 63, 1, 0);
			if (e->ncomp == 4)
			{
				int c3 = 3;
				do_scan(e, &c3, 1, 1, 63, 0, 1);
				do_scan(e, &c3, 1, 1, 63, 1, 0);
			}
		}
		else
		{
			do_scan(e, &y0, 1, 1, 5, 0, 2);
			do_scan(e, &y0, 1, 6, 63, 0, 2);
			do_scan(e, &y0, 1, 1, 63, 2, 1);
			do_scan(e, all, 1, 0, 0, 1, 0);
			do_scan(e, &y0, 1, 1, 63, 1, 0);
		}
	}
	for (i = 0; i < p->pad_fill_bytes; ++i)
		bs_byte(&e->out, 0xFF);
	bs_u16(&e->out, 0xFFD9);

	for (i = 0; i < e->ncomp; ++i)
		free(e->comp[i].coef);
	free(e->corr);
	*out = e->out.data;
	*out_len = e->out.size;
	return 1;
}

/* ---- procedural test images ---- */
/*
 * kind 0: SURVEY.md 8(c) probe image -- gradient XOR low-amplitude LCG noise
 *         (s = s*1664525 + 1013904223 per pixel, n = (s>>24)&31,
 *          R=(x*255/w)^n, G=(y*255/h)+n, B=((x+y)&255)^(n<<1)), seed 12345 reproduces the probe.
 * kind 1: smooth sin/cos field plus mild noise (lower bpp).
 * kind 2: hard-edged blocks and text-like stripes (long AC runs, large coefficients).
 */
TJE_API void tje_synth_rgb(unsigned char *rgb, int w, int h, int kind, unsigned seed)
{
	uint32_t s = seed;
	int x, y;
	for (y = 0; y < h; ++y)
		for (x = 0; x < w; ++x)
		{
			unsigned char *o = rgb + ((size_t)y * w + x) * 3;
			unsigned n;
			s = s * 1664525u + 1013904223u;
			n = (s >> 24) & 31u;
			if (kind == 0)
			{
				o[0] = (unsigned char)(((unsigned)(x * 255 / w)) ^ n);
				o[1] = (unsigned char)((unsigned)(y * 255 / h) + n);
				o[2] = (unsigned char)((((unsigned)(x + y)) & 255u) ^ (n << 1));
			}
			else if (kind == 1)
			{
				float fx = (float)x / (float)w, fy = (float)y / (float)h, ph = (float)(seed % 97u) * 0.13f;
				float r = 128.f + 100.f * sinf(6.283f * (fx * 2.3f + ph)) * cosf(6.283f * (fy * 1.7f));
				float g = 128.f + 90.f * sinf(6.283f * (fx * fy * 3.1f + ph * 0.5f));
				float b = 128.f + 80.f * cosf(6.283f * (fx * 1.1f - fy * 2.9f + ph));
				o[0] = (unsigned char)clampi((int)r + (int)(n & 7u) - 3, 0, 255);
				o[1] = (unsigned char)clampi((int)g + (int)((n >> 1) & 7u) - 3, 0, 255);
				o[2] = (unsigned char)clampi((int)b + (int)((n >> 2) & 7u) - 3, 0, 255);
			}
			else
			{
				unsigned cell = ((unsigned)(x / 24) * 2654435761u) ^ ((unsigned)(y / 18) * 40503u) ^ seed;
				int stripe = ((x + (int)(seed & 7u)) % 7) < 2;
				o[0] = (unsigned char)((cell >> 3) & 255u);
				o[1] = (unsigned char)(stripe ? 255 : (cell >> 11) & 255u);
				o[2] = (unsigned char)(((cell >> 19) & 255u) ^ (n & 3u));
			}
		}
}

/* ---- thread pool: synthesise + encode many images (bench input generation) ---- */
typedef struct
{
	const tje_params *p;
	int count, kind;
	unsigned base_seed;
	unsigned char **outs;
	size_t *lens;
	int tid, nthreads, ok;
} many_job;

static void *many_main(void *arg)
{
	many_job *j = (many_job *)arg;
	int i;
	unsigned char *rgb = (unsigned char *)malloc((size_t)j->p->width * j->p->height * 3);
	j->ok = 1;
	for (i = j->tid; i < j->count; i += j->nthreads)
	{
		tje_params p = *j->p;
		p.in_channels = 3;
		tje_synth_rgb(rgb, p.width, p.height, j->kind < 0 ? (i & 1) : j->kind, j->base_seed + (unsigned)i);
		if (!tje_encode(&p, rgb, &j->outs[i], &j->lens[i]))
			j->ok = 0;
	}
	free(rgb);
	return NULL;
}

/* kind < 0 alternates kinds 0 and 1 so bpp spans ~2..3.2 at q90 */
TJE_API int tje_encode_many(const tje_params *p, int count, int kind, unsigned base_seed, int nthreads,
									 unsigned char **outs, size_t *lens)
{
	pthread_t th[256];
	many_job jobs[256];
	int t, ok = 1;
	if (nthreads < 1) nthreads = 1;
	if (nthreads > 256) nthreads = 256;
	if (nthreads > count) nthreads = count > 0 ? count : 1;
	for (t = 0; t < nthreads; ++t)
	{
		jobs[t].p = p;
		jobs[t].count = count;
		jobs[t].kind = kind;
		jobs[t].base_seed = base_seed;
		jobs[t].outs = outs;
		jobs[t].lens = lens;
		jobs[t].tid = t;
		jobs[t].nthreads = nthreads;
		pthread_create(&th[t], NULL, many_main, &jobs[t]);
	}
	for (t = 0; t < nthreads; ++t)
	{
		pthread_join(th[t], NULL);
		ok &= jobs[t].ok;
	}
	return ok;
}
