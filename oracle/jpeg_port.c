/*
 * oracle/jpeg_port.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * A plain-C, serial CPU restatement of the reference's JPEG decode path, written for this
 * repository (no reference source is copied).  The preferred checker is the reference's own code
 * compiled from /root/reference into oracle/_ref/ (oracle/ref_unity.c); this port is what is left
 * when that build is absent.  It is PINNED: tests/test_oracle.py requires byte-identical output
 * against every golden fixture under tests/golden/ (produced by the reference build) and, whenever
 * oracle/_ref is present, against that build on the whole stream matrix for req_comp 0..4.
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
	/* bit reader, kept exactly like the reference's (jpeg.c:42-47, 167-187): a 32-bit left-aligned buffer that is
	   refilled at the same places to the same level.  On damaged streams WHERE a refill happens decides what the
	   decoder does: the refill that meets a marker returns without adding a single bit (jpeg.c:176-182), so a Huffman
	   code that straddles the end of the data fails with "bad huffman code" (jpeg.c:218, 237) unless an earlier refill
	   has already met the marker; and how far the reader stands ahead of the decoder at the end of a restart interval
	   decides whether the RSTn is seen (jpeg.c:1183-1187). */
	uint32_t cbuf;
	int cbits;	 /* valid bits at the top of cbuf; may go negative behind a marker, as in the reference */
	int stop;	 /* marker reached: zeros from now on ("nomore") */
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

static const uint32_t BMASK[17] = {0, 1, 3, 7, 15, 31, 63, 127, 255, 511, 1023, 2047, 4095, 8191, 16383, 32767, 65535};

/* stbi__grow_buffer_unsafe, jpeg.c:167-187: bytes until more than 24 bits are buffered; FF 00 -> FF, FF FF.. fill,
   FF xx -> remember the marker and return AT ONCE (no bits added by this call; later calls add zero bytes) */
static void grow(dec *d)
{
	do
	{
		unsigned b = d->stop ? 0u : (unsigned)get8(d);
		if (b == 0xff)
		{
			int c = get8(d);
			while (c == 0xff)
				c = get8(d);
			if (c != 0)
			{
				d->marker = c;
				d->stop = 1;
				return;
			}
		}
		/* cbits < 0 only behind a marker, where b == 0; the count is masked like the x86 shift the reference compiles to */
		d->cbuf |= b << ((24 - d->cbits) & 31);
		d->cbits += 8;
	} while (d->cbits <= 24);
}
/* n raw bits: stbi__jpeg_get_bits, jpeg.c:268-278 (n in 1..16) */
static int take(dec *d, int n)
{
	uint32_t k;
	if (n == 0)
		return 0;
	if (d->cbits < n)
		grow(d);
	k = (d->cbuf << n) | (d->cbuf >> (32 - n));
	d->cbuf = k & ~BMASK[n];
	d->cbits -= n;
	return (int)(k & BMASK[n]);
}
/* one bit: stbi__jpeg_get_bit, jpeg.c:280-289 */
static int take1(dec *d)
{
	uint32_t k;
	if (d->cbits < 1)
		grow(d);
	k = d->cbuf;
	d->cbuf <<= 1;
	--d->cbits;
	return (k & 0x80000000u) != 0;
}
static void reset_entropy(dec *d) /* stbi__jpeg_reset, jpeg.c:1142 */
{
	d->cbuf = 0;
	d->cbits = 0;
	d->stop = 0;
	d->marker = 0xff;
	d->c[0].dc_pred = d->c[1].dc_pred = d->c[2].dc_pred = d->c[3].dc_pred = 0;
	d->eobrun = 0;
}

/* stbi__jpeg_huff_decode, jpeg.c:193-243.  The reference finds the code through a 9-bit table and a maxcode walk; the
   canonical search below finds the same (length, symbol).  What matters on damaged data is kept literally: refill
   below 16 bits, "no such code" eats 16 bits, and a code longer than the bits actually buffered fails. */
static int decode_sym(dec *d, const htab *t)
{
	unsigned window, code;
	int l;
	if (d->cbits < 16)
		grow(d);
	window = d->cbuf >> 16;
	for (l = 1; l <= 16; ++l)
	{
		code = window >> (16 - l);
		if (t->count[l] && code >= t->first[l] && code - t->first[l] < (unsigned)t->count[l])
		{
			if (l > d->cbits)
				return -1; /* jpeg.c:218, 237 */
			d->cbuf <<= l;
			d->cbits -= l;
			return t->sym[t->index[l] + (int)(code - t->first[l])];
		}
	}
	d->cbits -= 16; /* jpeg.c:229 */
	return -1;
}

/* stbi__extend_receive, jpeg.c:250-265.  n > 16 (a damaged DC table) still triggers the refill and then yields 0
   without consuming anything; n == 16 reads past the reference's bias table (undefined there, 0 here). */
static int extend(dec *d, int n)
{
	static const int BIAS[17] = {0, -1, -3, -7, -15, -31, -63, -127, -255, -511, -1023, -2047, -4095, -8191, -16383, -32767, 0};
	uint32_t k;
	int sgn;
	if (d->cbits < n)
		grow(d);
	if (n < 0 || n >= 17)
		return 0;
	sgn = (int32_t)d->cbuf >> 31;
	if (n == 0)
		return 0;
	k = (d->cbuf << n) | (d->cbuf >> (32 - n));
	d->cbuf = k & ~BMASK[n];
	k &= BMASK[n];
	d->cbits -= n;
	return (int)k + (BIAS[n] & ~sgn);
}

/* The reference's fast-AC table (stbi__build_fast_ac, jpeg.c:138-165; used at jpeg.c:335-346 and 427-437): when the top
   9 bits hold a code of length l AND its s magnitude bits (s > 0, l + s <= 9) and the value fits -128..127, both are
   consumed in one step WITHOUT the "code longer than the buffer" check.  Returns 1 and fills run / value, else 0. */
static int fast_ac(dec *d, const htab *t, int *run, int *value)
{
	const unsigned top9 = d->cbuf >> 23;
	int l;
	for (l = 1; l <= 9; ++l)
	{
		const unsigned code = top9 >> (9 - l);
		if (t->count[l] && code >= t->first[l] && code - t->first[l] < (unsigned)t->count[l])
		{
			const int rs = t->sym[t->index[l] + (int)(code - t->first[l])];
			const int s = rs & 15;
			int k;
			if (s == 0 || l + s > 9)
				return 0;
			k = (int)((top9 << l) & 511u) >> (9 - s);
			if (k < (1 << (s - 1)))
				k += (int)(~0u << s) + 1;
			if (k < -128 || k > 127)
				return 0;
			d->cbuf <<= l + s;
			d->cbits -= l + s;
			*run = (rs >> 4) & 15;
			*value = k;
			return 1;
		}
	}
	return 0;
}

/* ---------------------------------------------------------------------------------------------- */
/* block decoders (jpeg.c:308-558).  Coefficient blocks are stored in natural order.              */
/* ---------------------------------------------------------------------------------------------- */

/* baseline: jpeg.c:308-370 -- dequantised at decode time, truncated to short */
static int block_baseline(dec *d, short *blk, pcomp *c)
{
	const unsigned short *q = d->dq[c->tq];
	int t, k;
	if (d->cbits < 16)
		grow(d);
	t = decode_sym(d, &d->dc[c->hd]);
	if (t < 0)
		return bad("bad huffman code");
	memset(blk, 0, 64 * sizeof(short));
	c->dc_pred += t ? extend(d, t) : 0;
	blk[0] = (short)(c->dc_pred * q[0]);
	for (k = 1; k < 64;)
	{
		int rs, r, s;
		if (d->cbits < 16)
			grow(d);
		if (fast_ac(d, &d->ac[c->ha], &r, &s))
		{
			int z;
			k += r;
			z = ZZ[k++];
			blk[z] = (short)(s * q[z]);
			continue;
		}
		rs = decode_sym(d, &d->ac[c->ha]);
		if (rs < 0)
			return bad("bad huffman code");
		r = rs >> 4;
		s = rs & 15;
		if (s == 0)
		{
			if (rs != 0xf0)
				break; /* end of block */
			k += 16;
		}
		else
		{
			int z;
			k += r;
			z = ZZ[k++];
			blk[z] = (short)(extend(d, s) * q[z]);
		}
	}
	return 1;
}

/* progressive DC, first and refinement passes: jpeg.c:372-402 */
static int block_prog_dc(dec *d, short *blk, pcomp *c)
{
	if (d->se != 0)
		return bad("can't merge dc and ac");
	if (d->cbits < 16)
		grow(d);
	if (d->ah == 0)
	{
		int t;
		memset(blk, 0, 64 * sizeof(short));
		t = decode_sym(d, &d->dc[c->hd]);
		if (t < 0)
			return bad("can't merge dc and ac"); /* the reference's text for this case, jpeg.c:388 */
		c->dc_pred += t ? extend(d, t) : 0;
		blk[0] = (short)(c->dc_pred * (1 << d->al));
	}
	else if (take1(d))
		blk[0] = (short)(blk[0] + (short)(1 << d->al));
	return 1;
}

/* progressive AC: jpeg.c:406-558 */
static int block_prog_ac(dec *d, short *blk, const htab *h)
{
	int k;
	if (d->ss == 0)
		return bad("can't merge dc and ac");
	if (d->ah == 0)
	{
		const int shift = d->al;
		if (d->eobrun)
		{
			--d->eobrun;
			return 1;
		}
		k = d->ss;
		do
		{
			int rs, r, s;
			if (d->cbits < 16)
				grow(d);
			if (fast_ac(d, h, &r, &s))
			{
				int z;
				k += r;
				z = ZZ[k++];
				blk[z] = (short)(s * (1 << shift));
				continue;
			}
			rs = decode_sym(d, h);
			if (rs < 0)
				return bad("bad huffman code");
			r = rs >> 4;
			s = rs & 15;
			if (s == 0)
			{
				if (r < 15)
				{
					d->eobrun = 1 << r;
					if (r)
						d->eobrun += take(d, r);
					--d->eobrun;
					break;
				}
				k += 16;
			}
			else
			{
				int z;
				k += r;
				z = ZZ[k++];
				blk[z] = (short)(extend(d, s) * (1 << shift));
			}
		} while (k <= d->se);
	}
	else
	{
		const short bit = (short)(1 << d->al);
		if (d->eobrun)
		{
			--d->eobrun;
			for (k = d->ss; k <= d->se; ++k)
			{
				short *p = &blk[ZZ[k]];
				if (*p != 0 && take1(d) && (*p & bit) == 0)
					*p = (short)(*p > 0 ? *p + bit : *p - bit);
			}
		}
		else
		{
			k = d->ss;
			do
			{
				int rs = decode_sym(d, h), r, s;
				if (rs < 0)
					return bad("bad huffman code");
				r = rs >> 4;
				s = rs & 15;
				if (s == 0)
				{
					if (r < 15)
					{
						d->eobrun = (1 << r) - 1;
						if (r)
							d->eobrun += take(d, r);
						r = 64; /* run to the end of the band */
					}
					/* r == 15: sixteen zero-history coefficients are skipped, none is set */
				}
				else
				{
					if (s != 1)
						return bad("bad huffman code");
					s = take1(d) ? bit : -bit;
				}
				while (k <= d->se)
				{
					short *p = &blk[ZZ[k++]];
					if (*p != 0)
					{
						if (take1(d) && (*p & bit) == 0)
							*p = (short)(*p > 0 ? *p + bit : *p - bit);
					}
					else
					{
						if (r == 0)
						{
							*p = (short)s;
							break;
						}
						--r;
					}
				}
			} while (k <= d->se);
		}
	}
	return 1;
}

/* ---------------------------------------------------------------------------------------------- */
/* one scan: stbi__parse_entropy_coded_data, jpeg.c:1155-1317                                     */
/* ---------------------------------------------------------------------------------------------- */
static int one_block(dec *d, pcomp *c, int bx, int by)
{
	short *blk = c->coef + 64 * ((size_t)by * c->bw + bx);
	if (!d->progressive)
		return block_baseline(d, blk, c);
	if (d->scan_n > 1 || d->ss == 0)
		return block_prog_dc(d, blk, c);
	return block_prog_ac(d, blk, &d->ac[c->ha]);
}

static int scan_decode(dec *d)
{
	int todo = d->ri ? d->ri : 0x7fffffff;
	int ux, uy, nx, ny;
	reset_entropy(d);
	if (d->scan_n == 1)
	{
		pcomp *c = &d->c[d->order[0]];
		nx = (c->x + 7) >> 3; /* blocks covering the component's real size (jpeg.c:1169) */
		ny = (c->y + 7) >> 3;
	}
	else
	{
		nx = d->mcux;
		ny = d->mcuy;
	}
	for (uy = 0; uy < ny; ++uy)
		for (ux = 0; ux < nx; ++ux)
		{
			if (d->scan_n == 1)
			{
				if (!one_block(d, &d->c[d->order[0]], ux, uy))
					return 0;
			}
			else
			{
				int k, x, y;
				for (k = 0; k < d->scan_n; ++k)
				{
					pcomp *c = &d->c[d->order[k]];
					for (y = 0; y < c->v; ++y)
						for (x = 0; x < c->h; ++x)
							if (!one_block(d, c, ux * c->h + x, uy * c->v + y))
								return 0;
				}
			}
			if (--todo <= 0)
			{
				/* restart interval over (jpeg.c:1183-1191): the next marker must be RSTn, else the scan ends */
				if (d->cbits < 24)
					grow(d);
				if (!(d->marker >= 0xd0 && d->marker <= 0xd7))
					return 1;
				reset_entropy(d);
				todo = d->ri ? d->ri : 0x7fffffff;
			}
		}
	return 1;
}

/* ---------------------------------------------------------------------------------------------- */
/* markers (jpeg.c:1119-1134, 1349-1468, 1471-1521, 1549-1659)                                    */
/* ---------------------------------------------------------------------------------------------- */
#define M_NONE 0xff

static int next_marker(dec *d) /* stbi__get_marker */
{
	int x;
	if (d->marker != M_NONE)
	{
		x = d->marker;
		d->marker = M_NONE;
		return x;
	}
	x = get8(d);
	if (x != 0xff)
		return M_NONE;
	while (x == 0xff)
		x = get8(d);
	return x;
}

static void skip(dec *d, int n)
{
	if (n > 0)
		d->pos += (size_t)n;
}

static int table_marker(dec *d, int m) /* stbi__process_marker */
{
	int L;
	switch (m)
	{
	case M_NONE:
		return bad("expected marker");
	case 0xdd: /* DRI */
		if (get16(d) != 4)
			return bad("bad DRI len");
		d->ri = get16(d);
		return 1;
	case 0xdb: /* DQT: 8- or 16-bit entries in zig-zag order, kept in natural order */
		L = get16(d) - 2;
		while (L > 0)
		{
			int q = get8(d), p = q >> 4, t = q & 15, i;
			if (p != 0 && p != 1)
				return bad("bad DQT type");
			if (t > 3)
				return bad("bad DQT table");
			for (i = 0; i < 64; ++i)
				d->dq[t][ZZ[i]] = (unsigned short)(p ? get16(d) : get8(d));
			L -= p ? 129 : 65;
		}
		return L == 0;
	case 0xc4: /* DHT */
		L = get16(d) - 2;
		while (L > 0)
		{
			u8 counts[16], syms[256];
			int q = get8(d), tc = q >> 4, th = q & 15, i, n = 0;
			if (tc > 1 || th > 3)
				return bad("bad DHT header");
			for (i = 0; i < 16; ++i)
			{
				counts[i] = (u8)get8(d);
				n += counts[i];
			}
			if (n > 256) /* the reference has no such check: stbi__build_huffman overruns its size[257] array (undefined) and in practice ends in this string */
				return bad("bad code lengths");
			L -= 17;
			for (i = 0; i < n; ++i)
				syms[i] = (u8)get8(d);
			if (!htab_build(tc ? &d->ac[th] : &d->dc[th], counts, syms, n))
				return 0;
			L -= n;
		}
		return L == 0;
	default:
		break;
	}
	if ((m >= 0xe0 && m <= 0xef) || m == 0xfe)
	{
		L = get16(d);
		if (L < 2)
			return bad(m == 0xfe ? "bad COM len" : "bad APP len");
		L -= 2;
		if (m == 0xe0 && L >= 5) /* JFIF */
		{
			static const u8 tag[5] = {'J', 'F', 'I', 'F', 0};
			int ok = 1, i;
			for (i = 0; i < 5; ++i)
				if (get8(d) != tag[i])
					ok = 0;
			L -= 5;
			if (ok)
				d->jfif = 1;
		}
		else if (m == 0xee && L >= 12) /* Adobe: version, flags0, flags1, colour transform */
		{
			static const u8 tag[6] = {'A', 'd', 'o', 'b', 'e', 0};
			int ok = 1, i;
			for (i = 0; i < 6; ++i)
				if (get8(d) != tag[i])
					ok = 0;
			L -= 6;
			if (ok)
			{
				get8(d);
				get16(d);
				get16(d);
				d->app14 = get8(d);
				L -= 6;
			}
		}
		skip(d, L);
		return 1;
	}
	return bad("unknown marker");
}

static int frame_header(dec *d) /* stbi__process_frame_header */
{
	int Lf = get16(d), i;
	if (Lf < 11)
		return bad("bad SOF len");
	if (get8(d) != 8)
		return bad("only 8-bit");
	d->h = get16(d);
	d->w = get16(d);
	if (d->h == 0)
		return bad("no header height");
	if (d->w == 0)
		return bad("0 width");
	d->n = get8(d);
	if (d->n != 1 && d->n != 3 && d->n != 4)
		return bad("bad component count");
	if (Lf != 8 + 3 * d->n)
		return bad("bad SOF len");
	d->rgb = 0;
	d->hmax = d->vmax = 1;
	for (i = 0; i < d->n; ++i)
	{
		static const u8 rgb[3] = {'R', 'G', 'B'};
		pcomp *c = &d->c[i];
		int q;
		c->id = get8(d);
		if (d->n == 3 && c->id == rgb[i])
			++d->rgb;
		q = get8(d);
		c->h = q >> 4;
		c->v = q & 15;
		if (c->h < 1 || c->h > 4)
			return bad("bad H");
		if (c->v < 1 || c->v > 4)
			return bad("bad V");
		c->tq = get8(d);
		if (c->tq > 3)
			return bad("bad TQ");
		if (c->h > d->hmax) d->hmax = c->h;
		if (c->v > d->vmax) d->vmax = c->v;
	}
	for (i = 0; i < d->n; ++i)
		if (d->hmax % d->c[i].h || d->vmax % d->c[i].v)
			return bad(d->hmax % d->c[i].h ? "bad H" : "bad V");
	d->mcux = (d->w + 8 * d->hmax - 1) / (8 * d->hmax);
	d->mcuy = (d->h + 8 * d->vmax - 1) / (8 * d->vmax);
	return 1;
}

static int alloc_planes(dec *d)
{
	int i;
	for (i = 0; i < d->n; ++i)
	{
		pcomp *c = &d->c[i];
		c->x = (d->w * c->h + d->hmax - 1) / d->hmax; /* jpeg.c:1627-1637 */
		c->y = (d->h * c->v + d->vmax - 1) / d->vmax;
		c->w2 = d->mcux * c->h * 8;
		c->h2 = d->mcuy * c->v * 8;
		c->bw = c->w2 / 8;
		c->bh = c->h2 / 8;
		c->coef = (short *)calloc((size_t)c->bw * c->bh * 64, sizeof(short));
		c->plane = (u8 *)calloc((size_t)c->w2 * c->h2 + 15, 1);
		if (!c->coef || !c->plane)
			return bad("outofmem");
	}
	return 1;
}

static int scan_header(dec *d) /* stbi__process_scan_header */
{
	int Ls = get16(d), i, q;
	d->scan_n = get8(d);
	if (d->scan_n < 1 || d->scan_n > 4 || d->scan_n > d->n)
		return bad("bad SOS component count");
	if (Ls != 6 + 2 * d->scan_n)
		return bad("bad SOS len");
	for (i = 0; i < d->scan_n; ++i)
	{
		int id = get8(d), which;
		q = get8(d);
		for (which = 0; which < d->n; ++which)
			if (d->c[which].id == id)
				break;
		if (which == d->n)
			return 0;
		d->c[which].hd = q >> 4;
		d->c[which].ha = q & 15;
		if (d->c[which].hd > 3)
			return bad("bad DC huff"); /* jpeg.c:1491 */
		if (d->c[which].ha > 3)
			return bad("bad AC huff"); /* jpeg.c:1494 */
		d->order[i] = which;
	}
	d->ss = get8(d);
	d->se = get8(d);
	q = get8(d);
	d->ah = q >> 4;
	d->al = q & 15;
	if (d->progressive)
	{
		if (d->ss > 63 || d->se > 63 || d->ss > d->se || d->ah > 13 || d->al > 13)
			return bad("bad SOS");
	}
	else
	{
		if (d->ss != 0 || d->ah != 0 || d->al != 0)
			return bad("bad SOS");
		d->se = 63;
	}
	return 1;
}

/* header up to and including SOF (stbi__decode_jpeg_header), then every scan (stbi__decode_jpeg_image) */
static int decode_markers(dec *d, int header_only)
{
	int m;
	d->marker = M_NONE;
	d->jfif = 0;
	d->app14 = -1;
	d->ri = 0; /* before the header: a DRI ahead of SOF stays in force (jpeg.c:1707) */
	m = next_marker(d);
	if (m != 0xd8)
		return bad("no SOI");
	m = next_marker(d);
	while (m != 0xc0 && m != 0xc1 && m != 0xc2)
	{
		if (!table_marker(d, m))
			return 0;
		m = next_marker(d);
		while (m == M_NONE)
		{
			if (at_eof(d))
				return bad("no SOF");
			m = next_marker(d);
		}
	}
	d->progressive = m == 0xc2;
	if (!frame_header(d))
		return 0;
	if (header_only)
		return 1;
	if (!alloc_planes(d))
		return 0;
	for (m = next_marker(d); m != 0xd9; m = next_marker(d))
	{
		if (m == 0xda)
		{
			if (!scan_header(d) || !scan_decode(d))
				return 0;
			if (d->marker == M_NONE)
				while (!at_eof(d)) /* bytes between the entropy data and the next marker (jpeg.c:1722-1735) */
					if (get8(d) == 0xff)
					{
						d->marker = get8(d);
						break;
					}
		}
		else if (m == 0xdc)
		{
			int Ld = get16(d), NL = get16(d);
			if (Ld != 4)
				return bad("bad DNL len");
			if (NL != d->h)
				return bad("bad DNL height");
		}
		else if (!table_marker(d, m))
			return 0;
	}
	return 1;
}

/* ---------------------------------------------------------------------------------------------- */
/* integer IDCT (stbi__idct_block, jpeg.c:561-679): 12-bit constants, column pass >> 10, row pass  */
/* >> 17 with the +128 level shift folded into the rounding term, clamp to 0..255                  */
/* ---------------------------------------------------------------------------------------------- */
static u8 clamp8(int x) { return (u8)(x < 0 ? 0 : x > 255 ? 255 : x); }

/* one 1-D pass: even part into e[0..3], odd part into o[0..3]; outputs are e[i] +- o[3-i].
   Constants = (int)(x * 4096 + 0.5), which truncates toward zero for the negative ones (jpeg.c:563) */
static void idct_1d(const int s[8], int e[4], int o[4])
{
	int p1, p2, p3, p4, p5, t0, t1, t2, t3;
	p1 = (s[2] + s[6]) * 2217;
	t2 = p1 + s[6] * -7567;
	t3 = p1 + s[2] * 3135;
	t0 = (s[0] + s[4]) * 4096;
	t1 = (s[0] - s[4]) * 4096;
	e[0] = t0 + t3;
	e[3] = t0 - t3;
	e[1] = t1 + t2;
	e[2] = t1 - t2;
	t0 = s[7];
	t1 = s[5];
	t2 = s[3];
	t3 = s[1];
	p3 = t0 + t2;
	p4 = t1 + t3;
	p1 = t0 + t3;
	p2 = t1 + t2;
	p5 = (p3 + p4) * 4816;
	t0 *= 1223;
	t1 *= 8410;
	t2 *= 12586;
	t3 *= 6149;
	p1 = p5 + p1 * -3685;
	p2 = p5 + p2 * -10497;
	p3 = p3 * -8034;
	p4 = p4 * -1597;
	o[3] = t3 + p1 + p4;
	o[2] = t2 + p2 + p3;
	o[1] = t1 + p2 + p4;
	o[0] = t0 + p1 + p3;
}

static void idct_block(u8 *out, int stride, const short in[64])
{
	int v[64], i, k, s[8], e[4], o[4];
	for (i = 0; i < 8; ++i) /* columns; the all-zero-AC shortcut of jpeg.c:624 gives the same values */
	{
		for (k = 0; k < 8; ++k)
			s[k] = in[i + 8 * k];
		idct_1d(s, e, o);
		for (k = 0; k < 4; ++k)
		{
			v[i + 8 * k] = (e[k] + 512 + o[3 - k]) >> 10;
			v[i + 8 * (7 - k)] = (e[k] + 512 - o[3 - k]) >> 10;
		}
	}
	for (i = 0; i < 8; ++i) /* rows */
	{
		const int bias = 65536 + (128 << 17);
		idct_1d(v + 8 * i, e, o);
		for (k = 0; k < 4; ++k)
		{
			out[i * stride + k] = clamp8((e[k] + bias + o[3 - k]) >> 17);
			out[i * stride + 7 - k] = clamp8((e[k] + bias - o[3 - k]) >> 17);
		}
	}
}

/* pass 3: coefficient planes -> sample planes.  Baseline blocks are already dequantised; progressive
   blocks are dequantised here in short arithmetic over the component's real size (jpeg.c:1319-1347) */
static void planes_from_coefficients(dec *d, int ncomp)
{
	int n, bx, by, i;
	for (n = 0; n < ncomp; ++n)
	{
		pcomp *c = &d->c[n];
		const int nx = d->progressive ? (c->x + 7) >> 3 : c->bw, ny = d->progressive ? (c->y + 7) >> 3 : c->bh;
		for (by = 0; by < ny; ++by)
			for (bx = 0; bx < nx; ++bx)
			{
				short *blk = c->coef + 64 * ((size_t)by * c->bw + bx);
				if (d->progressive)
					for (i = 0; i < 64; ++i)
						blk[i] = (short)(blk[i] * d->dq[c->tq][i]);
				idct_block(c->plane + (size_t)by * 8 * c->w2 + bx * 8, c->w2, blk);
			}
	}
}

/* ---------------------------------------------------------------------------------------------- */
/* upsampling rows (jpeg.c:1765-1840, 1962-1971)                                                  */
/* ---------------------------------------------------------------------------------------------- */
static u8 *up_v2(u8 *out, u8 *nr, u8 *fr, int w)
{
	int i;
	for (i = 0; i < w; ++i)
		out[i] = (u8)((3 * nr[i] + fr[i] + 2) >> 2);
	return out;
}
static u8 *up_h2(u8 *out, u8 *in, int w)
{
	int i;
	if (w == 1)
	{
		out[0] = out[1] = in[0];
		return out;
	}
	out[0] = in[0];
	out[1] = (u8)((in[0] * 3 + in[1] + 2) >> 2);
	for (i = 1; i < w - 1; ++i)
	{
		int n = 3 * in[i] + 2;
		out[2 * i] = (u8)((n + in[i - 1]) >> 2);
		out[2 * i + 1] = (u8)((n + in[i + 1]) >> 2);
	}
	out[2 * i] = (u8)((in[w - 2] * 3 + in[w - 1] + 2) >> 2);
	out[2 * i + 1] = in[w - 1];
	return out;
}
static u8 *up_hv2(u8 *out, u8 *nr, u8 *fr, int w)
{
	int i, t0, t1;
	if (w == 1)
	{
		out[0] = out[1] = (u8)((3 * nr[0] + fr[0] + 2) >> 2);
		return out;
	}
	t1 = 3 * nr[0] + fr[0];
	out[0] = (u8)((t1 + 2) >> 2);
	for (i = 1; i < w; ++i)
	{
		t0 = t1;
		t1 = 3 * nr[i] + fr[i];
		out[2 * i - 1] = (u8)((3 * t0 + t1 + 8) >> 4);
		out[2 * i] = (u8)((3 * t1 + t0 + 8) >> 4);
	}
	out[2 * w - 1] = (u8)((t1 + 2) >> 2);
	return out;
}
static u8 *up_generic(u8 *out, u8 *nr, int w, int hs)
{
	int i, j;
	for (i = 0; i < w; ++i)
		for (j = 0; j < hs; ++j)
			out[i * hs + j] = nr[i];
	return out;
}

/* ---------------------------------------------------------------------------------------------- */
/* colour (jpeg.c:1975-2018, 2218-2222; common.c:173)                                             */
/* ---------------------------------------------------------------------------------------------- */
#define FIX20(x) (((int)((x)*4096.0f + 0.5f)) << 8)
static void ycc_row(u8 *out, const u8 *y, const u8 *pcb, const u8 *pcr, int count, int step)
{
	int i;
	for (i = 0; i < count; ++i)
	{
		const int yf = (y[i] << 20) + (1 << 19);
		const int cr = pcr[i] - 128, cb = pcb[i] - 128;
		const int r = (yf + cr * FIX20(1.40200f)) >> 20;
		const int g = (yf + cr * -FIX20(0.71414f) + (int)((unsigned)(cb * -FIX20(0.34414f)) & 0xffff0000u)) >> 20;
		const int b = (yf + cb * FIX20(1.77200f)) >> 20;
		out[0] = clamp8(r);
		out[1] = clamp8(g);
		out[2] = clamp8(b);
		out[3] = 255;
		out += step;
	}
}
static u8 blinn(u8 x, u8 y)
{
	unsigned t = (unsigned)x * y + 128;
	return (u8)((t + (t >> 8)) >> 8);
}
static u8 luma(int r, int g, int b) { return (u8)(((r * 77) + (g * 150) + (29 * b)) >> 8); }

/* ---------------------------------------------------------------------------------------------- */
/* pass 4 + driver: load_jpeg_image, jpeg.c:2224-2440                                             */
/* ---------------------------------------------------------------------------------------------- */
typedef struct
{
	int hs, vs, w_lores, ystep, ypos;
	u8 *line0, *line1, *buf;
} upsampler;

static void dec_free(dec *d)
{
	int i;
	for (i = 0; i < 4; ++i)
	{
		free(d->c[i].coef);
		free(d->c[i].plane);
	}
}

static u8 *port_load(const u8 *buf, int len, int *px, int *py, int *pcomp_out, int req_comp)
{
	dec *d;
	u8 *out = NULL, *rows[4] = {0, 0, 0, 0};
	upsampler up[4];
	int n, decode_n, is_rgb, k, i, j;
	port_reason = NULL;
	if (len < 2 || buf[0] != 0xff || buf[1] != 0xd8)
	{
		/* stbi__load_main: only JPEG is recognised (image_api.c:55) */
		int p = 0;
		while (p < len && buf[p] == 0xff)
			++p;
		if (!(len >= 2 && buf[0] == 0xff && p < len && buf[p] == 0xd8))
		{
			bad("unknown image type");
			return NULL;
		}
	}
	if (req_comp < 0 || req_comp > 4)
	{
		bad("bad req_comp");
		return NULL;
	}
	d = (dec *)calloc(1, sizeof(dec));
	if (!d)
		return NULL;
	d->buf = buf;
	d->len = (size_t)(len > 0 ? len : 0);
	memset(up, 0, sizeof(up));
	if (!decode_markers(d, 0))
		goto done;
	n = req_comp ? req_comp : (d->n >= 3 ? 3 : 1);
	is_rgb = d->n == 3 && (d->rgb == 3 || (d->app14 == 0 && !d->jfif));
	decode_n = (d->n == 3 && n < 3 && !is_rgb) ? 1 : d->n;
	planes_from_coefficients(d, d->n);

	for (k = 0; k < decode_n; ++k)
	{
		upsampler *r = &up[k];
		r->buf = (u8 *)malloc((size_t)d->w + 3);
		if (!r->buf)
		{
			bad("outofmem");
			goto done;
		}
		r->hs = d->hmax / d->c[k].h;
		r->vs = d->vmax / d->c[k].v;
		r->ystep = r->vs >> 1;
		r->w_lores = (d->w + r->hs - 1) / r->hs;
		r->ypos = 0;
		r->line0 = r->line1 = d->c[k].plane;
	}
	out = (u8 *)malloc((size_t)n * d->w * d->h + 1);
	if (!out)
	{
		bad("outofmem");
		goto done;
	}
	for (j = 0; j < d->h; ++j)
	{
		u8 *o = out + (size_t)n * d->w * j;
		for (k = 0; k < decode_n; ++k)
		{
			upsampler *r = &up[k];
			const int bot = r->ystep >= (r->vs >> 1);
			u8 *nr = bot ? r->line1 : r->line0, *fr = bot ? r->line0 : r->line1;
			if (r->hs == 1 && r->vs == 1)
				rows[k] = nr;
			else if (r->hs == 1 && r->vs == 2)
				rows[k] = up_v2(r->buf, nr, fr, r->w_lores);
			else if (r->hs == 2 && r->vs == 1)
				rows[k] = up_h2(r->buf, nr, r->w_lores);
			else if (r->hs == 2 && r->vs == 2)
				rows[k] = up_hv2(r->buf, nr, fr, r->w_lores);
			else
				rows[k] = up_generic(r->buf, nr, r->w_lores, r->hs);
			if (++r->ystep >= r->vs)
			{
				r->ystep = 0;
				r->line0 = r->line1;
				if (++r->ypos < d->c[k].y)
					r->line1 += d->c[k].w2;
			}
		}
		if (n >= 3)
		{
			const u8 *y = rows[0];
			if (d->n == 3)
			{
				if (is_rgb)
					for (i = 0; i < d->w; ++i, o += n)
					{
						o[0] = y[i];
						o[1] = rows[1][i];
						o[2] = rows[2][i];
						o[3] = 255;
					}
				else
					ycc_row(o, y, rows[1], rows[2], d->w, n);
			}
			else if (d->n == 4)
			{
				if (d->app14 == 0) /* CMYK */
					for (i = 0; i < d->w; ++i, o += n)
					{
						const u8 m = rows[3][i];
						o[0] = blinn(rows[0][i], m);
						o[1] = blinn(rows[1][i], m);
						o[2] = blinn(rows[2][i], m);
						o[3] = 255;
					}
				else if (d->app14 == 2) /* YCCK */
				{
					ycc_row(o, y, rows[1], rows[2], d->w, n);
					for (i = 0; i < d->w; ++i, o += n)
					{
						const u8 m = rows[3][i];
						o[0] = blinn((u8)(255 - o[0]), m);
						o[1] = blinn((u8)(255 - o[1]), m);
						o[2] = blinn((u8)(255 - o[2]), m);
					}
				}
				else
					ycc_row(o, y, rows[1], rows[2], d->w, n);
			}
			else
				for (i = 0; i < d->w; ++i, o += n)
				{
					o[0] = o[1] = o[2] = y[i];
					o[3] = 255;
				}
		}
		else
		{
			if (is_rgb)
				for (i = 0; i < d->w; ++i, o += n)
				{
					o[0] = luma(rows[0][i], rows[1][i], rows[2][i]);
					if (n == 2)
						o[1] = 255;
				}
			else if (d->n == 4 && d->app14 == 0)
				for (i = 0; i < d->w; ++i, o += n)
				{
					const u8 m = rows[3][i];
					o[0] = luma(blinn(rows[0][i], m), blinn(rows[1][i], m), blinn(rows[2][i], m));
					o[1] = 255;
				}
			else if (d->n == 4 && d->app14 == 2)
				for (i = 0; i < d->w; ++i, o += n)
				{
					o[0] = blinn((u8)(255 - rows[0][i]), rows[3][i]);
					o[1] = 255;
				}
			else
			{
				const u8 *y = rows[0];
				if (n == 1)
					for (i = 0; i < d->w; ++i)
						o[i] = y[i];
				else
					for (i = 0; i < d->w; ++i)
					{
						*o++ = y[i];
						*o++ = 255;
					}
			}
		}
	}
	*px = d->w;
	*py = d->h;
	if (pcomp_out)
		*pcomp_out = d->n >= 3 ? 3 : 1;
done:
	for (k = 0; k < 4; ++k)
		free(up[k].buf);
	if (port_reason && out)
	{
		free(out);
		out = NULL;
	}
	dec_free(d);
	free(d);
	return out;
}

PORT_API u8 *stbi_load_from_memory(const u8 *buffer, int len, int *x, int *y, int *comp, int req_comp)
{
	return port_load(buffer, len, x, y, comp, req_comp);
}

PORT_API int stbi_info_from_memory(const u8 *buffer, int len, int *x, int *y, int *comp)
{
	dec *d = (dec *)calloc(1, sizeof(dec));
	int ok;
	port_reason = NULL;
	if (!d)
		return 0;
	d->buf = buffer;
	d->len = (size_t)(len > 0 ? len : 0);
	ok = decode_markers(d, 1);
	if (ok)
	{
		if (x) *x = d->w;
		if (y) *y = d->h;
		if (comp) *comp = d->n >= 3 ? 3 : 1;
	}
	free(d);
	return ok;
}

/* stage-level hooks, same interface as oracle/ref_unity.c */
PORT_API void oracle_idct_block(u8 *out, int out_stride, short data[64]) { idct_block(out, out_stride, data); }
PORT_API void oracle_ycbcr_row(u8 *out, const u8 *y, const u8 *cb, const u8 *cr, int count, int step)
{
	ycc_row(out, y, cb, cr, count, step);
}
PORT_API void oracle_resample_row(int kind, u8 *out, u8 *near_row, u8 *far_row, int w, int hs)
{
	switch (kind)
	{
	case 0: up_v2(out, near_row, far_row, w); break;
	case 1: up_h2(out, near_row, w); break;
	case 2: up_hv2(out, near_row, far_row, w); break;
	default: up_generic(out, near_row, w, hs); break;
	}
}

/* ---- timed multi-image decode (bench.py cpu_baseline leg when oracle/_ref is absent) ---- */
typedef struct
{
	const u8 *const *bufs;
	const int *lens;
	int count, req_comp, reps, tid, nthreads, failures;
	long long pixels;
} worker;

static void *worker_main(void *arg)
{
	worker *w = (worker *)arg;
	int r, i;
	for (r = 0; r < w->reps; ++r)
		for (i = w->tid; i < w->count; i += w->nthreads)
		{
			int x, y, c;
			u8 *p = port_load(w->bufs[i], w->lens[i], &x, &y, &c, w->req_comp);
			if (p)
			{
				w->pixels += (long long)x * y;
				free(p);
			}
			else
				w->failures++;
		}
	return NULL;
}

PORT_API double oracle_time_decode(const u8 *const *bufs, const int *lens, int count, int req_comp, int reps, int nthreads,
											  long long *pixels, int *failures)
{
	pthread_t th[256];
	worker w[256];
	struct timespec t0, t1;
	int t;
	if (nthreads < 1) nthreads = 1;
	if (nthreads > 256) nthreads = 256;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (t = 0; t < nthreads; ++t)
	{
		worker v = {bufs, lens, count, req_comp, reps, t, nthreads, 0, 0};
		w[t] = v;
		pthread_create(&th[t], NULL, worker_main, &w[t]);
	}
	*pixels = 0;
	*failures = 0;
	for (t = 0; t < nthreads; ++t)
	{
		pthread_join(th[t], NULL);
		*pixels += w[t].pixels;
		*failures += w[t].failures;
	}
	clock_gettime(CLOCK_MONOTONIC, &t1);
	return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
