/*
 * oracle/ref_unity.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Unity build of the UNMODIFIED reference sources, #included from where they lie under
 * REF_ROOT (= /root/reference, passed by oracle/Makefile).  Nothing from the reference is
 * copied into this repository; the built library lands in oracle/_ref/ (git-ignored).
 *
 * Include order follows SURVEY.md Appendix A:
 *   prelude -> common.c -> codec/jpeg.c -> image_api.c -> convert.c -> codec/write_bmp.c
 *   -> codec/jpeg_write.c
 *
 * Besides the reference's own public entry points (stbi_load_from_memory, stbi_info_from_memory,
 * stbi_write_jpg_to_func, stbi_failure_reason, stbi_image_free) this file exports small
 * `oracle_*` hooks that call the reference's static kernels directly, for stage-level
 * known-answer tests:
 *   oracle_idct_block      -> stbi__idct_block          (jpeg.c:615)
 *   oracle_ycbcr_row       -> stbi__YCbCr_to_RGB_row    (jpeg.c:1976)
 *   oracle_resample_*      -> stbi__resample_row_*      (jpeg.c:1774-1840,1962)
 *   oracle_decode_capture  -> load_jpeg_image (jpeg.c:2224) with idct_block_kernel hooked so
 *                             the dequantised coefficient blocks and the component planes
 *                             can be inspected.
 */
#define STBI_NO_PNG
#define STBI_NO_BMP
#define STBI_NO_GIF
#define STBI_NO_PSD
#define STBI_NO_PIC
#define STBI_NO_PNM
#define STBI_NO_HDR
#define STBI_NO_TGA
#ifndef ORACLE_LINEAR /* the _linear build keeps stbi_loadf* (convert.c:286-337, common.c:391-424) */
#define STBI_NO_LINEAR
#endif
#define STBI_NO_STDIO
#define STBI_NO_ZLIB
#define STBI_WRITE_NO_STDIO

#include "ref_prelude.h"

#define ORACLE_STR2(x) #x
#define ORACLE_STR(x) ORACLE_STR2(x)
#define ORACLE_REF(file) ORACLE_STR(REF_ROOT/file)

#include ORACLE_REF(common.c)
#include ORACLE_REF(codec/jpeg.c)

/* image_api.c uses stbi__info_main (image_api.c:123) but the reference never defines it. */
static int stbi__info_main(stbi__context *s, int *x, int *y, int *comp);
#include ORACLE_REF(image_api.c)
static int stbi__info_main(stbi__context *s, int *x, int *y, int *comp)
{
	if (stbi__jpeg_info(s, x, y, comp))
		return 1;
	return stbi__err("unknown image type", "Image not of any known type, or corrupt");
}

#include ORACLE_REF(convert.c)
#include ORACLE_REF(codec/write_bmp.c)
#include ORACLE_REF(codec/jpeg_write.c)

/* ------------------------------------------------------------------------------------ */
/* stage-level hooks                                                                     */
/* ------------------------------------------------------------------------------------ */

#define ORACLE_API __attribute__((visibility("default")))

/* the reference's flip flag (convert.c:97-101 reads it; its setter lived in the missing header) */
ORACLE_API void oracle_set_flip(int flag) { stbi__vertically_flip_on_load = flag; }

ORACLE_API int oracle_is_sse2(void)
{
#ifdef ORACLE_USE_SSE2
	return 1;
#else
	return 0;
#endif
}

ORACLE_API void oracle_idct_block(unsigned char *out, int out_stride, short data[64])
{
	stbi__idct_block(out, out_stride, data);
}

ORACLE_API void oracle_ycbcr_row(unsigned char *out, const unsigned char *y, const unsigned char *cb,
											const unsigned char *cr, int count, int step)
{
	stbi__YCbCr_to_RGB_row(out, y, cb, cr, count, step);
}

/* kind: 0 = v2 (jpeg.c:1774), 1 = h2 (:1784), 2 = hv2 (:1816), 3 = generic (:1962) */
ORACLE_API void oracle_resample_row(int kind, unsigned char *out, unsigned char *near_row, unsigned char *far_row,
												int w, int hs)
{
	switch (kind)
	{
	case 0: stbi__resample_row_v_2(out, near_row, far_row, w, hs); break;
	case 1: stbi__resample_row_h_2(out, near_row, far_row, w, hs); break;
	case 2: stbi__resample_row_hv_2(out, near_row, far_row, w, hs); break;
	default: stbi__resample_row_generic(out, near_row, far_row, w, hs); break;
	}
}

/* ---- capture of IDCT inputs / outputs during a full decode ---- */
typedef struct
{
	stbi__jpeg *z;
	short *coef[4];			 /* per component: (h2/8)*(w2/8)*64 shorts, natural order, dequantised */
	unsigned char *plane[4]; /* per component: h2*w2 samples */
	unsigned char *seen[4];	 /* per component: (h2/8)*(w2/8) flags: block was IDCT'd */
} oracle_capture;

static __thread oracle_capture *oracle__cap;

static void oracle__capture_idct(stbi_uc *out, int out_stride, short data[64])
{
	oracle_capture *c = oracle__cap;
	stbi__jpeg *z = c->z;
	int n, r;
	stbi__idct_block(out, out_stride, data);
	for (n = 0; n < z->s->img_n; ++n)
	{
		stbi_uc *base = z->img_comp[n].data;
		size_t span = (size_t)z->img_comp[n].w2 * z->img_comp[n].h2;
		if (out >= base && out < base + span)
		{
			size_t off = (size_t)(out - base);
			int by = (int)(off / z->img_comp[n].w2) / 8;
			int bx = (int)(off % z->img_comp[n].w2) / 8;
			int bw = z->img_comp[n].w2 / 8;
			if (c->coef[n])
				memcpy(c->coef[n] + 64 * ((size_t)by * bw + bx), data, 64 * sizeof(short));
			if (c->seen[n])
				c->seen[n][(size_t)by * bw + bx] = 1;
			if (c->plane[n])
				for (r = 0; r < 8; ++r)
					memcpy(c->plane[n] + (size_t)(by * 8 + r) * z->img_comp[n].w2 + bx * 8, out + (size_t)r * out_stride, 8);
			return;
		}
	}
}

/*
 * Geometry query: fills w2[4], h2[4] (MCU-padded plane sizes, jpeg.c:1636-1637) and
 * returns img_n, or 0 on failure.  Runs the reference header parse (jpeg.c:1670).
 */
ORACLE_API int oracle_plane_geometry(const unsigned char *buf, int len, int *img_x, int *img_y, int w2[4], int h2[4],
												 int cx[4], int cy[4])
{
	stbi__context s;
	stbi__jpeg *j = (stbi__jpeg *)malloc(sizeof(stbi__jpeg));
	int i, n = 0, h_max = 1, v_max = 1;
	stbi__start_mem(&s, buf, len);
	j->s = &s;
	if (stbi__decode_jpeg_header(j, STBI__SCAN_header))
	{
		n = s.img_n;
		*img_x = (int)s.img_x;
		*img_y = (int)s.img_y;
		for (i = 0; i < n; ++i)
		{
			if (j->img_comp[i].h > h_max) h_max = j->img_comp[i].h;
			if (j->img_comp[i].v > v_max) v_max = j->img_comp[i].v;
		}
		for (i = 0; i < n; ++i)
		{
			int mcu_x = ((int)s.img_x + h_max * 8 - 1) / (h_max * 8);
			int mcu_y = ((int)s.img_y + v_max * 8 - 1) / (v_max * 8);
			w2[i] = mcu_x * j->img_comp[i].h * 8;
			h2[i] = mcu_y * j->img_comp[i].v * 8;
			cx[i] = ((int)s.img_x * j->img_comp[i].h + h_max - 1) / h_max;
			cy[i] = ((int)s.img_y * j->img_comp[i].v + v_max - 1) / v_max;
		}
	}
	free(j);
	return n;
}

/*
 * Full decode through load_jpeg_image (jpeg.c:2224) with the IDCT hook installed.
 * coef[n] / plane[n] / seen[n] may each be NULL.  Buffers must be sized from
 * oracle_plane_geometry and zero-filled by the caller.  Returns the malloc'ed pixel
 * buffer (free with stbi_image_free) or NULL.
 */
ORACLE_API unsigned char *oracle_decode_capture(const unsigned char *buf, int len, int *x, int *y, int *comp,
																int req_comp, short *coef[4], unsigned char *plane[4],
																unsigned char *seen[4])
{
	stbi__context s;
	oracle_capture cap;
	unsigned char *result;
	int i;
	stbi__jpeg *j = (stbi__jpeg *)malloc(sizeof(stbi__jpeg));
	stbi__start_mem(&s, buf, len);
	j->s = &s;
	stbi__setup_jpeg(j);
	j->idct_block_kernel = oracle__capture_idct;
	cap.z = j;
	for (i = 0; i < 4; ++i)
	{
		cap.coef[i] = coef ? coef[i] : NULL;
		cap.plane[i] = plane ? plane[i] : NULL;
		cap.seen[i] = seen ? seen[i] : NULL;
	}
	oracle__cap = &cap;
	result = load_jpeg_image(j, x, y, comp, req_comp);
	oracle__cap = NULL;
	free(j);
	return result;
}

/* ---- jpeg_write.c into a growable memory buffer ---- */
typedef struct
{
	unsigned char *data;
	size_t size, cap;
} oracle_membuf;

static void oracle__mem_write(void *ctx, void *data, int size)
{
	oracle_membuf *m = (oracle_membuf *)ctx;
	if (m->size + (size_t)size > m->cap)
	{
		size_t ncap = m->cap ? m->cap * 2 : 1 << 16;
		while (ncap < m->size + (size_t)size)
			ncap *= 2;
		m->data = (unsigned char *)realloc(m->data, ncap);
		m->cap = ncap;
	}
	memcpy(m->data + m->size, data, (size_t)size);
	m->size += (size_t)size;
}

/* stbi_write_jpg_to_func (jpeg_write.c:368) into memory; returns malloc'ed stream, *out_len set. */
ORACLE_API unsigned char *oracle_write_jpg(int w, int h, int comp, const void *pixels, int quality, int *out_len)
{
	oracle_membuf m = {0, 0, 0};
	if (!stbi_write_jpg_to_func(oracle__mem_write, &m, w, h, comp, pixels, quality))
	{
		free(m.data);
		return NULL;
	}
	*out_len = (int)m.size;
	return m.data;
}

ORACLE_API void oracle_free(void *p) { free(p); }

/* ---- timed multi-image decode for the CPU baseline (bench.py cpu_baseline leg) ---- */
#include <pthread.h>
#include <time.h>

typedef struct
{
	const unsigned char *const *bufs;
	const int *lens;
	int count, req_comp, reps;
	int tid, nthreads;
	long long pixels;
	int failures;
} oracle_worker;

static void *oracle__worker_main(void *arg)
{
	oracle_worker *w = (oracle_worker *)arg;
	int r, i;
	for (r = 0; r < w->reps; ++r)
		for (i = w->tid; i < w->count; i += w->nthreads)
		{
			int x, y, c;
			unsigned char *p = stbi_load_from_memory(w->bufs[i], w->lens[i], &x, &y, &c, w->req_comp);
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

/*
 * Decode `count` streams `reps` times with `nthreads` worker threads, each thread decoding
 * whole images (the reference has no intra-image parallelism, jpeg.c:5).  Returns wall
 * seconds; *pixels = total decoded pixels; *failures = failed decodes.
 */
ORACLE_API double oracle_time_decode(const unsigned char *const *bufs, const int *lens, int count, int req_comp,
												 int reps, int nthreads, long long *pixels, int *failures)
{
	pthread_t th[256];
	oracle_worker w[256];
	struct timespec t0, t1;
	int t;
	if (nthreads < 1) nthreads = 1;
	if (nthreads > 256) nthreads = 256;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (t = 0; t < nthreads; ++t)
	{
		w[t].bufs = bufs;
		w[t].lens = lens;
		w[t].count = count;
		w[t].req_comp = req_comp;
		w[t].reps = reps;
		w[t].tid = t;
		w[t].nthreads = nthreads;
		w[t].pixels = 0;
		w[t].failures = 0;
		pthread_create(&th[t], NULL, oracle__worker_main, &w[t]);
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
