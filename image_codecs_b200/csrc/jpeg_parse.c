/*
 * jpeg_parse.c -- host-side JPEG marker/header parser.  See jpeg_parse.h.
 *
 * Accept/reject behaviour and failure strings follow the reference;
 * each function cites the reference routine whose behaviour it reproduces.  The structure is
 * this repo's own: a cursor over an in-memory file, a parser state holding the *raw* DQT/DHT
 * contents (the lookup tables are built on the device), and an entropy-coded-segment walker that
 * replaces stbi__parse_entropy_coded_data (jpeg.c:1155) on the host side by locating markers only.
 */
#include "jpeg_parse.h"

#include <limits.h>
#include <stdlib.h>
#include <string.h>

#define MARKER_NONE 0xff

/* zig-zag position -> row-major index (T.81 Figure A.6; reference table jpeg.c:293) */
static const unsigned char jp_natural_order[64] = {
	 0,  1,  8,  16, 9,  2,	 3,  10, 17, 24, 32, 25, 18, 11, 4,	 5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,	7,	 14, 21, 28,
	 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

typedef struct
{
	const unsigned char *buf;
	size_t len, pos;
} cursor;

/* reading past the end yields 0, as the reference's memory reader does (common.c:39) */
static int cur_u8(cursor *c) { return c->pos < c->len ? c->buf[c->pos++] : (c->pos++, 0); }
static int cur_u16(cursor *c)
{
	int hi = cur_u8(c);
	return (hi << 8) + cur_u8(c);
}
static int cur_eof(const cursor *c) { return c->pos >= c->len; }
/* stbi__skip (common.c:64): negative skips jump to the end */
static void cur_skip(cursor *c, int n)
{
	if (n < 0)
		c->pos = c->len;
	else
		c->pos += (size_t)n;
}

typedef struct
{
	cursor cur;
	jp_result *r;
	int pending; /* marker seen while walking entropy-coded data (reference: j->marker) */

	unsigned short dequant[4][64]; /* natural order */
	jcu_huff_spec dc[4], ac[4];
	int dc_pool[4], ac_pool[4]; /* index of the current contents in r->huff, -1 = not pushed yet */
	int dc_defined[4], ac_defined[4];

	int id[4], tq[4];
	int progressive, jfif, app14, rgb, restart_interval;
	int ncomp;
} parser;

static int fail(parser *p, const char *why)
{
	p->r->ok = 0;
	p->r->reason = why;
	return 0;
}

/* ---- growable arrays: on allocation failure the old array (and capacity) stay and 0 is returned ---- */
static int grow(void **ptr, int *cap, int need, size_t elem)
{
	if (need > *cap)
	{
		int nc = *cap ? *cap * 2 : 16;
		void *np;
		while (nc < need)
			nc *= 2;
		np = realloc(*ptr, (size_t)nc * elem);
		if (!np)
			return 0;
		*ptr = np;
		*cap = nc;
	}
	return 1;
}

/* index of the new entry, or -1 (with the result marked "outofmem") */
static int push_huff(parser *p, const jcu_huff_spec *s)
{
	jp_result *r = p->r;
	if (!grow((void **)&r->huff, &r->cap_huff, r->nhuff + 1, sizeof(jcu_huff_spec)))
	{
		fail(p, "outofmem");
		return -1;
	}
	r->huff[r->nhuff] = *s;
	return r->nhuff++;
}

/* stbi__get_marker (jpeg.c:1119) */
static int next_marker(parser *p)
{
	int x;
	if (p->pending != MARKER_NONE)
	{
		x = p->pending;
		p->pending = MARKER_NONE;
		return x;
	}
	x = cur_u8(&p->cur);
	if (x != 0xff)
		return MARKER_NONE;
	while (x == 0xff)
		x = cur_u8(&p->cur);
	return x;
}

/* the canonical-code validity test of stbi__build_huffman (jpeg.c:101-115), without building tables */
static int code_lengths_valid(const unsigned char bits[16])
{
	unsigned code = 0;
	int j;
	for (j = 1; j <= 16; ++j)
	{
		if (bits[j - 1])
		{
			code += bits[j - 1];
			if (code - 1 >= (1u << j))
				return 0;
		}
		code <<= 1;
	}
	return 1;
}

/* stbi__process_marker (jpeg.c:1349) */
static int table_or_misc_marker(parser *p, int m)
{
	cursor *c = &p->cur;
	int L;
	switch (m)
	{
	case MARKER_NONE:
		return fail(p, "expected marker");
	case 0xDD:
		if (cur_u16(c) != 4)
			return fail(p, "bad DRI len");
		p->restart_interval = cur_u16(c);
		return 1;
	case 0xDB:
		L = cur_u16(c) - 2;
		while (L > 0)
		{
			int q = cur_u8(c), prec = q >> 4, t = q & 15, i;
			if (prec != 0 && prec != 1)
				return fail(p, "bad DQT type");
			if (t > 3)
				return fail(p, "bad DQT table");
			for (i = 0; i < 64; ++i)
				p->dequant[t][jp_natural_order[i]] = (unsigned short)(prec ? cur_u16(c) : cur_u8(c));
			L -= prec ? 129 : 65;
		}
		if (L != 0)
		{
			p->r->ok = 0; /* reference returns L==0 without setting a reason (jpeg.c:1379) */
			return 0;
		}
		return 1;
	case 0xC4:
		L = cur_u16(c) - 2;
		while (L > 0)
		{
			jcu_huff_spec spec;
			int q = cur_u8(c), tc = q >> 4, th = q & 15, i, n = 0;
			if (tc > 1 || th > 3)
				return fail(p, "bad DHT header");
			memset(&spec, 0, sizeof(spec));
			for (i = 0; i < 16; ++i)
			{
				spec.bits[i] = (unsigned char)cur_u8(c);
				n += spec.bits[i];
			}
			L -= 17;
			/* n > 256 overruns the reference's 256-entry arrays (undefined there); rejected here */
			if (n > 256 || !code_lengths_valid(spec.bits))
				return fail(p, "bad code lengths");
			for (i = 0; i < n; ++i)
				spec.vals[i] = (unsigned char)cur_u8(c);
			if (tc == 0)
			{
				p->dc[th] = spec;
				p->dc_pool[th] = -1;
				p->dc_defined[th] = 1;
			}
			else
			{
				p->ac[th] = spec;
				p->ac_pool[th] = -1;
				p->ac_defined[th] = 1;
			}
			L -= n;
		}
		if (L != 0)
		{
			p->r->ok = 0; /* jpeg.c:1416: no reason string */
			return 0;
		}
		return 1;
	default:
		break;
	}
	if ((m >= 0xE0 && m <= 0xEF) || m == 0xFE)
	{
		L = cur_u16(c);
		if (L < 2)
			return fail(p, m == 0xFE ? "bad COM len" : "bad APP len");
		L -= 2;
		if (m == 0xE0 && L >= 5)
		{
			static const unsigned char tag[5] = {'J', 'F', 'I', 'F', 0};
			int ok = 1, i;
			for (i = 0; i < 5; ++i)
				if (cur_u8(c) != tag[i])
					ok = 0;
			L -= 5;
			if (ok)
				p->jfif = 1;
		}
		else if (m == 0xEE && L >= 12)
		{
			static const unsigned char tag[6] = {'A', 'd', 'o', 'b', 'e', 0};
			int ok = 1, i;
			for (i = 0; i < 6; ++i)
				if (cur_u8(c) != tag[i])
					ok = 0;
			L -= 6;
			if (ok)
			{
				cur_u8(c);	/* version */
				cur_u16(c); /* flags0 */
				cur_u16(c); /* flags1 */
				p->app14 = cur_u8(c);
				L -= 6;
			}
		}
		cur_skip(c, L);
		return 1;
	}
	return fail(p, "unknown marker");
}

/* stbi__process_frame_header (jpeg.c:1549) */
static int frame_header(parser *p, int mode)
{
	cursor *c = &p->cur;
	jcu_image *im = &p->r->img;
	int Lf, prec, i, q, n, hmax = 1, vmax = 1;
	unsigned w, h;
	Lf = cur_u16(c);
	if (Lf < 11)
		return fail(p, "bad SOF len");
	prec = cur_u8(c);
	if (prec != 8)
		return fail(p, "only 8-bit");
	h = (unsigned)cur_u16(c);
	if (h == 0)
		return fail(p, "no header height");
	w = (unsigned)cur_u16(c);
	if (w == 0)
		return fail(p, "0 width");
	n = cur_u8(c);
	if (n != 3 && n != 1 && n != 4)
		return fail(p, "bad component count");
	if (Lf != 8 + 3 * n)
		return fail(p, "bad SOF len");
	p->ncomp = n;
	p->rgb = 0;
	for (i = 0; i < n; ++i)
	{
		static const unsigned char rgb_ids[3] = {'R', 'G', 'B'};
		p->id[i] = cur_u8(c);
		if (n == 3 && p->id[i] == rgb_ids[i])
			++p->rgb;
		q = cur_u8(c);
		im->comp[i].h = q >> 4;
		if (!im->comp[i].h || im->comp[i].h > 4)
			return fail(p, "bad H");
		im->comp[i].v = q & 15;
		if (!im->comp[i].v || im->comp[i].v > 4)
			return fail(p, "bad V");
		p->tq[i] = cur_u8(c);
		if (p->tq[i] > 3)
			return fail(p, "bad TQ");
	}
	im->width = (int)w;
	im->height = (int)h;
	im->ncomp = n;
	p->r->comps_in_file = n >= 3 ? 3 : 1;
	if (mode != JP_MODE_LOAD)
		return 1;

	/* stbi__mad3sizes_valid(img_x, img_y, img_n, 0), jpeg.c:1604 */
	if ((unsigned long long)w * h * (unsigned)n > (unsigned long long)INT_MAX)
		return fail(p, "too large");

	for (i = 0; i < n; ++i)
	{
		if (im->comp[i].h > hmax) hmax = im->comp[i].h;
		if (im->comp[i].v > vmax) vmax = im->comp[i].v;
	}
	im->hmax = hmax;
	im->vmax = vmax;
	im->mcu_x = ((int)w + hmax * 8 - 1) / (hmax * 8);
	im->mcu_y = ((int)h + vmax * 8 - 1) / (vmax * 8);
	for (i = 0; i < n; ++i)
	{
		jcu_comp *k = &im->comp[i];
		k->cx = ((int)w * k->h + hmax - 1) / hmax;
		k->cy = ((int)h * k->v + vmax - 1) / vmax;
		k->bw = im->mcu_x * k->h;
		k->bh = im->mcu_y * k->v;
		k->cbw = (k->cx + 7) >> 3;
		k->cbh = (k->cy + 7) >> 3;
		k->hs = hmax / k->h;
		k->vs = vmax / k->v;
		k->w_lores = ((int)w + k->hs - 1) / k->hs;
		/* stbi__malloc_mad2(w2, h2, 15), jpeg.c:1641 */
		if ((unsigned long long)k->bw * 8ull * (unsigned long long)k->bh * 8ull + 15ull > (unsigned long long)INT_MAX)
			return fail(p, "outofmem");
		/* progressive: the coefficient plane, stbi__malloc_mad3(w2, h2, sizeof(short), 15), jpeg.c:1651 */
		if (p->progressive &&
			 (unsigned long long)k->bw * 8ull * (unsigned long long)k->bh * 8ull * 2ull + 15ull > (unsigned long long)INT_MAX)
			return fail(p, "outofmem");
	}
	return 1;
}

/* stbi__process_scan_header (jpeg.c:1471); fills sc->ncomp/comp/ss/se/ah/al and table selectors */
static int scan_header(parser *p, jcu_scan *sc, int hd[4], int ha[4])
{
	cursor *c = &p->cur;
	int i, Ls = cur_u16(c), aa;
	sc->ncomp = cur_u8(c);
	if (sc->ncomp < 1 || sc->ncomp > 4 || sc->ncomp > p->ncomp)
		return fail(p, "bad SOS component count");
	if (Ls != 6 + 2 * sc->ncomp)
		return fail(p, "bad SOS len");
	for (i = 0; i < sc->ncomp; ++i)
	{
		int id = cur_u8(c), q = cur_u8(c), which;
		for (which = 0; which < p->ncomp; ++which)
			if (p->id[which] == id)
				break;
		if (which == p->ncomp)
		{
			p->r->ok = 0; /* jpeg.c:1488 returns 0 without a reason */
			return 0;
		}
		hd[i] = q >> 4;
		if (hd[i] > 3)
			return fail(p, "bad DC huff");
		ha[i] = q & 15;
		if (ha[i] > 3)
			return fail(p, "bad AC huff");
		sc->comp[i] = which;
	}
	sc->ss = cur_u8(c);
	sc->se = cur_u8(c);
	aa = cur_u8(c);
	sc->ah = aa >> 4;
	sc->al = aa & 15;
	if (p->progressive)
	{
		if (sc->ss > 63 || sc->se > 63 || sc->ss > sc->se || sc->ah > 13 || sc->al > 13)
			return fail(p, "bad SOS");
	}
	else
	{
		if (sc->ss != 0)
			return fail(p, "bad SOS");
		if (sc->ah != 0 || sc->al != 0)
			return fail(p, "bad SOS");
		sc->se = 63;
	}
	return 1;
}

/*
 * Walk the entropy-coded bytes of one scan, starting at the cursor.  This is the host's share of
 * stbi__parse_entropy_coded_data (jpeg.c:1155): it does not decode, it only reproduces how the
 * reference's bit reader (stbi__grow_buffer_unsafe, jpeg.c:167) classifies 0xFF sequences:
 *   FF 00        -> a data byte                      FF FF..      -> fill bytes
 *   FF D0..D7    -> restart marker: when a restart interval is in force and more intervals are
 *                   expected, the next segment starts after it (jpeg.c:1180-1189; the marker's
 *                   number is not checked, jpeg.c:1138)
 *   FF xx        -> any other marker ends the scan and becomes the pending marker
 * Reaching the end of the file leaves no pending marker (the reference then fails with
 * "expected marker", jpeg.c:1722-1735 + 1355).
 */
static int walk_entropy_segments(parser *p, jcu_scan *sc)
{
	jp_result *r = p->r;
	const unsigned char *buf = p->cur.buf;
	size_t len = p->cur.len, pos = p->cur.pos, scan_begin = pos;
	int ri = sc->restart_interval;
	unsigned expected = ri ? (unsigned)((sc->total_units + ri - 1) / ri) : 1u;
	size_t marker_bytes = 0;

	sc->seg_first = (uint32_t)r->nsegs;
	sc->nseg = 1;
	if (!grow((void **)&r->segs, &r->cap_segs, r->nsegs + 1, sizeof(uint32_t)))
		return fail(p, "outofmem");
	r->segs[r->nsegs++] = (uint32_t)(pos < len ? pos : len);
	p->pending = MARKER_NONE;

	for (;;)
	{
		const unsigned char *ff = pos < len ? (const unsigned char *)memchr(buf + pos, 0xff, len - pos) : NULL;
		size_t at, q;
		int code;
		if (!ff)
		{
			pos = len;
			sc->data_end = (uint32_t)len;
			break;
		}
		at = (size_t)(ff - buf);
		q = at + 1;
		while (q < len && buf[q] == 0xff)
			++q;
		code = q < len ? buf[q] : 0;
		if (code == 0)
		{
			pos = q + 1;
			if (pos > len)
			{
				pos = len;
				sc->data_end = (uint32_t)len;
				break;
			}
			continue;
		}
		if (ri && code >= 0xd0 && code <= 0xd7 && sc->nseg < expected)
		{
			if (!grow((void **)&r->segs, &r->cap_segs, r->nsegs + 1, sizeof(uint32_t)))
				return fail(p, "outofmem");
			r->segs[r->nsegs++] = (uint32_t)(q + 1);
			sc->nseg++;
			marker_bytes += (q + 1) - at;
			pos = q + 1;
			continue;
		}
		sc->data_end = (uint32_t)at;
		pos = q + 1;
		if (ri && code >= 0xd0 && code <= 0xd7 && sc->total_units % ri == 0)
		{
			/* One RSTn too many behind a scan whose last interval is complete: the restart check at the end of that
			   interval (jpeg.c:1180-1189) takes it for a restart and clears it, the scan returns without a pending
			   marker, and the marker loop skips to the next 0xFF and takes the byte behind it as the marker
			   (jpeg.c:1722-1735), whatever lies in between. */
			const unsigned char *nf = pos < len ? (const unsigned char *)memchr(buf + pos, 0xff, len - pos) : NULL;
			sc->end_rst = 1;
			if (!nf)
			{
				pos = len;
				break;
			}
			pos = (size_t)(nf - buf) + 2;
			code = pos - 1 < len ? buf[pos - 1] : 0;
			if (pos > len)
				pos = len + 1; /* the missing byte read as 0 (cur_u8 past the end) */
		}
		p->pending = code;
		break;
	}
	p->cur.pos = pos;
	r->entropy_bytes += (double)(sc->data_end - scan_begin) - (double)marker_bytes;
	/* fewer restart intervals than the scan needs (a lost RSTn followed by a marker): the reference's scan ends
	   "successfully" and the blocks behind it keep what malloc left there (jpeg.c:1184-1187); here they must not
	   keep what an EARLIER BATCH left in the reused coefficient slab, so the batch builder clears it */
	if (sc->nseg < expected)
		r->needs_clear = 1;
	return 1;
}

static int pool_index(parser *p, int is_ac, int t)
{
	if (is_ac)
	{
		if (!p->ac_defined[t])
			return -1;
		if (p->ac_pool[t] < 0)
			p->ac_pool[t] = push_huff(p, &p->ac[t]);
		return p->ac_pool[t];
	}
	if (!p->dc_defined[t])
		return -1;
	if (p->dc_pool[t] < 0)
		p->dc_pool[t] = push_huff(p, &p->dc[t]);
	return p->dc_pool[t];
}

/* one SOS: header, scan descriptor, segment walk (reference jpeg.c:1716-1736) */
static int one_scan(parser *p)
{
	jp_result *r = p->r;
	jcu_image *im = &r->img;
	jcu_scan sc;
	int hd[4], ha[4], i, needs_dc, needs_ac;
	memset(&sc, 0, sizeof(sc));
	if (!scan_header(p, &sc, hd, ha))
		return 0;

	if (!p->progressive)
		sc.kind = JCU_SCAN_BASELINE;
	else if (sc.ncomp > 1 || sc.ss == 0)
	{
		/* interleaved progressive scans and Ss==0 scans take the DC decoder (jpeg.c:1253,1297),
		   which rejects Se != 0 on its first block (jpeg.c:376) */
		if (sc.se != 0)
			return fail(p, "can't merge dc and ac");
		sc.kind = sc.ah == 0 ? JCU_SCAN_PROG_DC_FIRST : JCU_SCAN_PROG_DC_REFINE;
	}
	else
		sc.kind = sc.ah == 0 ? JCU_SCAN_PROG_AC_FIRST : JCU_SCAN_PROG_AC_REFINE;

	needs_dc = sc.kind == JCU_SCAN_BASELINE || sc.kind == JCU_SCAN_PROG_DC_FIRST;
	needs_ac = sc.kind == JCU_SCAN_BASELINE || sc.kind == JCU_SCAN_PROG_AC_FIRST || sc.kind == JCU_SCAN_PROG_AC_REFINE;
	for (i = 0; i < 4; ++i)
		sc.dc_spec[i] = sc.ac_spec[i] = -1;
	for (i = 0; i < sc.ncomp; ++i)
	{
		if (needs_dc)
		{
			sc.dc_spec[i] = pool_index(p, 0, hd[i]);
			if (sc.dc_spec[i] < 0)
				return p->r->ok ? fail(p, "bad huffman code") : 0; /* table never defined: undefined in the reference */
		}
		if (needs_ac)
		{
			sc.ac_spec[i] = pool_index(p, 1, ha[i]);
			if (sc.ac_spec[i] < 0)
				return p->r->ok ? fail(p, "bad huffman code") : 0;
		}
	}
	sc.restart_interval = p->restart_interval;
	if (sc.ncomp == 1)
	{
		const jcu_comp *k = &im->comp[sc.comp[0]];
		sc.units_per_row = k->cbw; /* jpeg.c:1169: blocks covering the real component size */
		sc.total_units = k->cbw * k->cbh;
	}
	else
	{
		sc.units_per_row = im->mcu_x;
		sc.total_units = im->mcu_x * im->mcu_y;
	}
	/* Wave = dependency level: a scan must follow every earlier scan of the image that touches one of its
	   components in an overlapping coefficient band (the reference runs scans strictly in file order,
	   jpeg.c:1716; scans on disjoint data commute).  Baseline and first DC scans own the whole block (the
	   latter clears it, jpeg.c:380). */
	{
		/* write set in zig-zag positions.  A first AC pass can land up to 15 positions past its band: the
		   reference does not check `k += r` against spec_end before it stores (jpeg.c:461-463), so on a damaged
		   stream the stray coefficient is part of its result and later scans of those positions must see it */
#define SCAN_LO(s) (((s)->kind == JCU_SCAN_BASELINE || (s)->kind == JCU_SCAN_PROG_DC_FIRST) ? 0 : (s)->ss)
#define SCAN_HI(s) (((s)->kind == JCU_SCAN_BASELINE || (s)->kind == JCU_SCAN_PROG_DC_FIRST) ? 63 \
						  : (s)->kind == JCU_SCAN_PROG_AC_FIRST ? ((s)->se + 15 > 63 ? 63 : (s)->se + 15) : (s)->se)
		const int lo = SCAN_LO(&sc), hi = SCAN_HI(&sc);
		int j, a, b, level = 0;
		for (j = 0; j < r->nscans; ++j)
		{
			const jcu_scan *o = &r->scans[j];
			const int olo = SCAN_LO(o), ohi = SCAN_HI(o);
			int shared = 0;
			for (a = 0; a < sc.ncomp; ++a)
				for (b = 0; b < o->ncomp; ++b)
					shared |= sc.comp[a] == o->comp[b];
			if (shared && lo <= ohi && olo <= hi && o->wave + 1 > level)
				level = o->wave + 1;
		}
		sc.wave = level;
	}
	/* baseline dequantises at decode time with the table in force at this scan (jpeg.c:1176) */
	if (!p->progressive)
		for (i = 0; i < sc.ncomp; ++i)
			memcpy(im->qt[sc.comp[i]], p->dequant[p->tq[sc.comp[i]]], sizeof(im->qt[0]));

	if (!walk_entropy_segments(p, &sc))
		return 0;

	if (!grow((void **)&r->scans, &r->cap_scans, r->nscans + 1, sizeof(jcu_scan)))
		return fail(p, "outofmem");
	r->scans[r->nscans++] = sc;
	return 1;
}

/* output-stage selection: the case analysis of load_jpeg_image (jpeg.c:2241-2249, 2320-2431) */
static void choose_output(parser *p, int req_comp)
{
	jcu_image *im = &p->r->img;
	int n = req_comp ? req_comp : (im->ncomp >= 3 ? 3 : 1);
	int is_rgb = im->ncomp == 3 && (p->rgb == 3 || (p->app14 == 0 && !p->jfif));
	int i, mode;
	im->n_out = n;
	im->decode_n = (im->ncomp == 3 && n < 3 && !is_rgb) ? 1 : im->ncomp;
	if (n >= 3)
	{
		if (im->ncomp == 3)
			mode = is_rgb ? JCU_CM_RGB : JCU_CM_YCBCR;
		else if (im->ncomp == 4)
			mode = p->app14 == 0 ? JCU_CM_CMYK : p->app14 == 2 ? JCU_CM_YCCK : JCU_CM_YCBCR;
		else
			mode = JCU_CM_Y;
	}
	else
	{
		if (is_rgb)
			mode = JCU_CM_RGB_LUMA;
		else if (im->ncomp == 4 && p->app14 == 0)
			mode = JCU_CM_CMYK_LUMA;
		else if (im->ncomp == 4 && p->app14 == 2)
			mode = JCU_CM_YCCK_LUMA;
		else
			mode = JCU_CM_Y;
	}
	im->color_mode = mode;

	im->path = JCU_PATH_GENERIC;
	if (mode == JCU_CM_Y && im->comp[0].h == im->hmax && im->comp[0].v == im->vmax)
		im->path = JCU_PATH_FUSED_GRAY; /* component 0 is at full resolution: resample_row_1 (jpeg.c:2270) */
	else if (mode == JCU_CM_YCBCR && im->ncomp == 3)
	{
		const jcu_comp *c = im->comp;
		if (c[0].h == 2 && c[0].v == 2 && c[1].h == 1 && c[1].v == 1 && c[2].h == 1 && c[2].v == 1)
			im->path = JCU_PATH_FUSED_420;
		else if (c[0].h == 1 && c[0].v == 1 && c[1].h == 1 && c[1].v == 1 && c[2].h == 1 && c[2].v == 1)
			im->path = JCU_PATH_FUSED_444;
		else if (c[0].h == 2 && c[0].v == 1 && c[1].h == 1 && c[1].v == 1 && c[2].h == 1 && c[2].v == 1)
			im->path = JCU_PATH_FUSED_422;
	}
	p->r->coef_blocks = 0;
	for (i = 0; i < im->decode_n; ++i)
		p->r->coef_blocks += (long long)im->comp[i].bw * im->comp[i].bh;
	/* a component used by the output stage but sampled with hmax % h != 0 makes the reference read
	   outside its planes (jpeg.c:2273-2276 uses integer division); such files are not decodable here */
}

void jp_result_free(jp_result *r)
{
	free(r->scans);
	free(r->segs);
	free(r->huff);
	memset(r, 0, sizeof(*r));
}

int jp_parse(const unsigned char *buf, size_t len, int req_comp, int mode, jp_result *r)
{
	parser P, *p = &P;
	int m, i;
	memset(p, 0, sizeof(*p));
	memset(r, 0, sizeof(*r));
	p->cur.buf = buf;
	p->cur.len = len;
	p->r = r;
	p->pending = MARKER_NONE;
	p->app14 = -1;
	for (i = 0; i < 4; ++i)
		p->dc_pool[i] = p->ac_pool[i] = -1;
	r->ok = 1;

	/* stbi__decode_jpeg_header (jpeg.c:1670) */
	m = next_marker(p);
	if (m != 0xd8)
		return fail(p, "no SOI");
	if (mode == JP_MODE_TYPE)
		return 1;
	m = next_marker(p);
	while (!(m == 0xc0 || m == 0xc1 || m == 0xc2))
	{
		if (!table_or_misc_marker(p, m))
			return 0;
		m = next_marker(p);
		while (m == MARKER_NONE)
		{
			if (cur_eof(&p->cur))
				return fail(p, "no SOF");
			m = next_marker(p);
		}
	}
	p->progressive = (m == 0xc2);
	r->img.progressive = p->progressive;
	if (!frame_header(p, mode))
		return 0;
	if (mode == JP_MODE_HEADER)
		return 1;

	/* load_jpeg_image validates req_comp before decoding (jpeg.c:2230); done by the caller before
	   jp_parse so that "bad req_comp" wins over parse errors, as in the reference */
	(void)req_comp;

	/* marker loop of stbi__decode_jpeg_image (jpeg.c:1713-1752) */
	m = next_marker(p);
	while (m != 0xd9)
	{
		if (m == 0xda)
		{
			if (!one_scan(p))
				return 0;
		}
		else if (m == 0xdc)
		{
			int Ld = cur_u16(&p->cur);
			unsigned NL = (unsigned)cur_u16(&p->cur);
			if (Ld != 4)
				return fail(p, "bad DNL len");
			if (NL != (unsigned)r->img.height)
				return fail(p, "bad DNL height");
		}
		else
		{
			if (!table_or_misc_marker(p, m))
				return 0;
		}
		m = next_marker(p);
	}
	/* progressive dequantises after the last scan with the tables then in force (jpeg.c:1341) */
	if (p->progressive)
		for (i = 0; i < p->ncomp; ++i)
			memcpy(r->img.qt[i], p->dequant[p->tq[i]], sizeof(r->img.qt[0]));

	/* a frame component that no scan carried: its blocks are never written by the entropy stage (the reference
	   shows uninitialised heap there); the builder clears the slab so the result is deterministic */
	for (i = 0; i < p->ncomp; ++i)
	{
		int s_, k_, seen = 0;
		for (s_ = 0; s_ < r->nscans; ++s_)
			for (k_ = 0; k_ < r->scans[s_].ncomp; ++k_)
				seen |= r->scans[s_].comp[k_] == i;
		if (!seen)
			r->needs_clear = 1;
	}

	choose_output(p, req_comp);
	r->consumed = p->cur.pos < p->cur.len ? p->cur.pos : p->cur.len;
	r->img.ok = 1;
	return 1;
}
