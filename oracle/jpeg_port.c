/*
 * oracle/jpeg_port.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * A plain-C, serial CPU restatement of the reference's JPEG decode path, written for this
 * repository (no reference source is copied).  It is the checker of last resort: the preferred
 * oracle is the reference's own code compiled from /root/reference into oracle/_ref/
 * (oracle/ref_unity.c).  This port is PINNED against that build: tests/test_oracle.py requires
 * byte-identical output on the whole stream matrix whenever oracle/_ref is present, and
 * tests/golden/ holds hashes produced by the reference build.
 *
 * Structure (deliberately unlike the reference's decode-and-IDCT-as-you-go, jpeg.c:1155-1234):
 *   1. marker pass     -> frame, tables, scan list                 (jpeg.c:1349-1756)
 *   2. entropy pass    -> all scans decoded into coefficient planes (jpeg.c:308-558, 1155-1317)
 *   3. dequant + IDCT  -> sample planes                            (jpeg.c:615-679, 1319-1347)
 *   4. resample+colour -> packed pixels                            (jpeg.c:1765-2018, 2224-2440)
 * Each routine cites the reference lines whose arithmetic it restates.
 *
 * Same C interface as oracle/ref_unity.c (see oracle/oracle_lib.py).
 */
#include <limits.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define PORT_API __attribute__((visibility("default")))

typedef unsigned char u8;

static __thread const char *port_reason;
PORT_API const char *stbi_failure_reason(void) { return port_reason; }
PORT_API void stbi_image_free(void *p) { free(p); }
PORT_API void oracle_free(void *p) { free(p); }
PORT_API int oracle_is_sse2(void) { return 0; }

static int bad(const char *why)
{
	port_reason = why;
	return 0;
}

/* zig-zag -> natural, padded so runs past 63 land on 63 (jpeg.c:293-305) */
static const u8 ZZ[80] = {0,	1,	 8,  16, 9,	 2,  3,	10, 17, 24, 32, 25, 18, 11, 4,  5,	12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13,
								  6,	7,	 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31,
								  39, 46, 53, 60, 61, 54, 47, 55, 62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

/* ---------------------------------------------------------------------------------------------- */
/* Huffman tables: canonical codes per T.81 Annex C; lookup = linear walk over lengths            */
/* (results equal stbi__jpeg_huff_decode, jpeg.c:193, which accelerates the same codes)           */
/* ---------------------------------------------------------------------------------------------- */
typedef struct
{
	int defined;
	int count[17];
	unsigned first[17]; /* first code of each length */
	int index[17];		  /* index of its symbol */
	u8 sym[256];
} htab;

static int htab_build(htab *t, const u8 counts[16], const u8 *syms, int n)
{
	unsigned code = 0;
	int l, k = 0;
	for (l = 1; l <= 16; ++l)
	{
		t->count[l] = counts[l - 1];
		t->first[l] = code;
		t->index[l] = k;
		if (counts[l - 1])
		{
			code += counts[l - 1];
			if (code - 1 >= (1u << l))
				return bad("bad code lengths"); /* jpeg.c:109 */
		}
		k += counts[l - 1];
		code <<= 1;
	}
	memcpy(t->sym, syms, (size_t)n);
	t->defined = 1;
	return 1;
}

/* ---------------------------------------------------------------------------------------------- */
/* decoder state                                                                                  */
/* ---------------------------------------------------------------------------------------------- */
typedef struct
{
	int id, h, v, tq, hd, ha;
	int x, y, w2, h2, bw, bh; /* jpeg.c:1627-1637 */
	int dc_pred;
	short *coef; /* bw*bh*64, natural order */
	u8 *plane;	 /* w2*h2 */
	u8 *have;	 /* per block: IDCT'd by the reference */
} pcomp;

typedef struct
{
	const u8 *buf;
	size_t len, pos;
	/* bit reader (stbi__grow_buffer_unsafe semantics, jpeg.c:167) */
	uint64_t bits;
	int nbits;
	int stop;	 /* marker reached: zeros from now on */
	int marker; /* the marker that stopped us, 0xff = none */

	unsigned short dq[4][64];
	htab dc[4], ac[4];
	pcomp c[4];
	int n, w, h, hmax, vmax, mcux, mcuy;
	int progressive, jfif, app14, rgb, ri;
	int scan_n, order[4], ss, se, ah, al, eobrun;
} dec;

static int get8(dec *d) { return d->pos < d->len ? d->buf[d->pos++] : (d->pos++, 0); } /* common.c:30-39 */
static int get16(dec *d)
{
	int a = get8(d);
	return (a << 8) + get8(d);
}
static int at_eof(dec *d) { return d->pos >= d->len; }

static void fill_bits(dec *d)
{
	while (d->nbits <= 56)
	{
		unsigned b = 0;
		if (!d->stop)
		{
			b = (unsigned)get8(d);
			if (b == 0xff)
			{
				int c = get8(d);
				while (c == 0xff)
					c = get8(d);
				if (c != 0)
				{
					d->marker = c;
					d->stop = 1;
					b = 0;
				}
			}
		}
		d->bits |= (uint64_t)b << (56 - d->nbits);
		d->nbits += 8;
	}
}
static unsigned peek(dec *d, int n)
{
	if (d->nbits < n)
		fill_bits(d);
	return (unsigned)(d->bits >> (64 - n));
}
static void drop(dec *d, int n)
{
	d->bits <<= n;
	d->nbits -= n;
}
static int take(dec *d, int n)
{
	unsigned v;
	if (n == 0)
		return 0;
	v = peek(d, n);
	drop(d, n);
	return (int)v;
}
static void reset_entropy(dec *d) /* stbi__jpeg_reset, jpeg.c:1142 */
{
	d->bits = 0;
	d->nbits = 0;
	d->stop = 0;
	d->marker = 0xff;
	d->c[0].dc_pred = d->c[1].dc_pred = d->c[2].dc_pred = d->c[3].dc_pred = 0;
	d->eobrun = 0;
}

static int decode_sym(dec *d, const htab *t)
{
	unsigned code = 0;
	int l;
	unsigned window = peek(d, 16);
	for (l = 1; l <= 16; ++l)
	{
		code = window >> (16 - l);
		if (t->count[l] && code >= t->first[l] && code - t->first[l] < (unsigned)t->count[l])
		{
			drop(d, l);
			return t->sym[t->index[l] + (int)(code - t->first[l])];
		}
	}
	drop(d, 16);
	return -1;
}

static int extend(dec *d, int n) /* stbi__extend_receive, jpeg.c:250 */
{
	int v;
	if (n == 0 || n > 16)
		return 0;
	v = take(d, n);
	if (v < (1 << (n - 1)))
		v += (int)(~0u << n) + 1;
	return v;
}
