/*
 * image_api.c -- the stbi_* entry points of include/image_api.h (host C).
 *
 * Mirrors the reference's public layer for the JPEG leg:
 *   stbi_load* / stbi_info*          convert.c:188-266, image_api.c:74-131
 *   stbi__load_main (JPEG probe)     image_api.c:3-13, :55
 *   stbi__load_and_postprocess_8bit  convert.c:78-104 (vertical flip)
 *   stbi__jpeg_test / _info          jpeg.c:2454-2490
 * and adds the batched decoder.  All pixel work happens on the GPU through
 * include/jpeg_cuda_abi.h; there is no CPU decode path in this library.
 */
#include "../../include/image_api.h"

#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <math.h>
#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_parse.h"

#define API __attribute__((visibility("default")))

/* ---- failure reason, flip flag (names from the reference's missing header) ---- */
static __thread const char *tl_reason;
static int g_flip_on_load;

static void *err_null(const char *why)
{
	if (why)
		tl_reason = why;
	return NULL;
}
static int err_zero(const char *why)
{
	if (why)
		tl_reason = why;
	return 0;
}

API const char *stbi_failure_reason(void) { return tl_reason; }
API void stbi_image_free(void *p) { free(p); }
API void stbi_set_flip_vertically_on_load(int flag) { g_flip_on_load = flag; }
/* PNG-only switches of the reference (convert.c / png.c): accepted so that callers link unchanged; a JPEG has no
   alpha to un-premultiply and no CgBI chunk */
API void stbi_set_unpremultiply_on_load(int flag) { (void)flag; }
API void stbi_convert_iphone_png_to_rgb(int flag) { (void)flag; }

static double now_ms(void)
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return 1e3 * (double)ts.tv_sec + 1e-6 * (double)ts.tv_nsec;
}

/* ===================================================================================== */
/* batch object                                                                          */
/* ===================================================================================== */
/* persistent marker-parser workers, shared by a batch object and the further objects of its chunk pipeline (one
   chunk is parsed at a time); created on first use, joined when the owning object is destroyed */
struct parse_job;
typedef struct
{
	pthread_t th[64];
	int n;
	pthread_mutex_t mu;
	pthread_cond_t cv_work, cv_done;
	unsigned gen;
	int pending, stop, active; /* active: workers that take part in the current job (tids 0..active-1) */
	const struct parse_job *job;
} parse_pool;

struct stbi_jpeg_batch
{
	stbi_jpeg_batch_params params;
	jcu_ctx *ctx;
	int nthreads;
	int force_generic; /* params.force_generic: route every image through the two-pass path (tests) */
	parse_pool *pool;	 /* owned by the root object; pipeline objects point at their root's */
	int pool_owner;
	size_t mem_budget; /* device bytes one prepared chunk may need (coefficients + planes + pixels + bitstreams) */

	/* prepared batch */
	int count, req_comp;
	jp_result *res; /* per image */
	int cap_res;
	int *dev_err;
	int dev_err_valid;

	/* merged descriptors (host) */
	jcu_image *images;
	jcu_scan *scans;
	uint32_t *segs;
	jcu_huff_spec *huff;
	size_t cap_images, cap_scans, cap_segs, cap_huff;
	int n_scans, n_huff, n_waves;
	uint32_t n_segs;

	unsigned char *slab; /* pinned bitstream staging */
	size_t slab_cap, slab_bytes;
	size_t coef_elems, plane_bytes, out_bytes;
	int need_clear;
	int flip; /* stbi_set_flip_vertically_on_load as it stood when the batch was prepared: the kernels store bottom-up */

	stbi_jpeg_batch_timings tm;

#define STBI_JPEG_PIPE_MAX 4
	struct stbi_jpeg_batch *pipe[STBI_JPEG_PIPE_MAX - 1]; /* further objects (own stream, staging and device buffers) so that the host-buffer
												call can parse / upload chunk k+1 while chunk k computes and downloads */
};

static int online_cpus(void)
{
	long n = sysconf(_SC_NPROCESSORS_ONLN);
	if (n < 1) n = 1;
	if (n > 64) n = 64;
	return (int)n;
}

/* failure word of the device decoders -> the reference's string */
static const char *dev_reason(int word)
{
	return JCU_ERR_CODE(word) == JCU_ERR_UNKNOWN_MARKER ? "unknown marker" /* jpeg.c:1469 */
																		 : "bad huffman code"; /* jpeg.c:317,351,444,499,521 */
}

API int stbi_jpeg_cuda_device_count(void) { return jcu_device_count(); }

static stbi_jpeg_batch *batch_create_with_pool(const stbi_jpeg_batch_params *params, parse_pool *shared)
{
	stbi_jpeg_batch *b;
	jcu_ctx_options o;
	size_t free_b = 0, total_b = 0;
	int rc;
	if (jcu_device_count() < 1)
		return (stbi_jpeg_batch *)err_null("no cuda device");
	b = (stbi_jpeg_batch *)calloc(1, sizeof(*b));
	if (!b)
		return (stbi_jpeg_batch *)err_null("outofmem");
	if (params)
		b->params = *params;
	b->nthreads = b->params.host_threads > 0 ? b->params.host_threads : online_cpus();
	if (b->nthreads > 64) b->nthreads = 64;
	b->force_generic = b->params.force_generic != 0;
	b->pool = shared;
	memset(&o, 0, sizeof(o));
	o.device = b->params.device;
	o.profile = b->params.profile;
	o.entropy_mode = b->params.entropy_mode;
	o.force_generic = b->force_generic;
	o.rows_per_band = b->params.rows_per_band;
	o.guard = b->params.debug_guards;
	o.selfsync_sub_bytes = b->params.selfsync_sub_bytes;
	o.host_threads = b->nthreads;
	rc = jcu_ctx_create_opts(&o, &b->ctx);
	if (rc != 0)
	{
		free(b);
		return (stbi_jpeg_batch *)err_null("cuda context");
	}
	/* what one prepared chunk may ask of the device: the caller's figure, else most of the device for the staged
	   calls (stbi_jpeg_decode_batch narrows it to a share per pipeline object) */
	if (b->params.device_mem_mb > 0)
		b->mem_budget = (size_t)b->params.device_mem_mb << 20;
	else if (jcu_mem_info(b->params.device, &free_b, &total_b) == 0 && total_b)
		b->mem_budget = total_b / 10 * 9;
	else
		b->mem_budget = (size_t)16 << 30;
	return b;
}

API stbi_jpeg_batch *stbi_jpeg_batch_create(const stbi_jpeg_batch_params *params) { return batch_create_with_pool(params, NULL); }

static void batch_release_results(stbi_jpeg_batch *b)
{
	int i;
	for (i = 0; i < b->count; ++i)
		jp_result_free(&b->res[i]);
	b->count = 0;
}

static void pool_destroy(parse_pool *pl);

API void stbi_jpeg_batch_destroy(stbi_jpeg_batch *b)
{
	if (!b)
		return;
	{
		int pi;
		for (pi = 0; pi < STBI_JPEG_PIPE_MAX - 1; ++pi)
			if (b->pipe[pi])
				stbi_jpeg_batch_destroy(b->pipe[pi]);
	}
	if (b->pool_owner)
		pool_destroy(b->pool);
	batch_release_results(b);
	jcu_ctx_destroy(b->ctx);
	jcu_host_free(b->slab);
	free(b->res);
	free(b->dev_err);
	free(b->images);
	free(b->scans);
	free(b->segs);
	free(b->huff);
	free(b);
}

/* ---- threaded parse: each worker copies its files into the pinned slab and parses them there ---- */
typedef struct parse_job
{
	stbi_jpeg_batch *b;
	const stbi_uc *const *bufs;
	const int *lens;
	const size_t *offs;
	int nthreads;
} parse_job;

static void parse_share(const parse_job *j, int tid)
{
	stbi_jpeg_batch *b = j->b;
	int i;
	for (i = tid; i < b->count; i += j->nthreads)
	{
		jp_result *r = &b->res[i];
		size_t len = j->lens[i] > 0 ? (size_t)j->lens[i] : 0;
		unsigned char *dst = b->slab + j->offs[i];
		jp_result probe;
		if (len)
			memcpy(dst, j->bufs[i], len);
		/* stbi__load_main: JPEG is recognised by SOI alone (image_api.c:10, jpeg.c:2460) */
		if (!jp_parse(dst, len, b->req_comp, JP_MODE_TYPE, &probe))
		{
			memset(r, 0, sizeof(*r));
			r->ok = 0;
			r->reason = "unknown image type";
			continue;
		}
		jp_parse(dst, len, b->req_comp, JP_MODE_LOAD, r);
		if (r->ok)
		{
			int k;
			if (b->force_generic)
				r->img.path = JCU_PATH_GENERIC;
			for (k = 0; k < r->img.decode_n; ++k)
				if (r->img.hmax % r->img.comp[k].h || r->img.vmax % r->img.comp[k].v)
				{
					r->ok = 0;
					r->img.ok = 0;
					r->reason = "unsupported sampling"; /* the reference reads outside its planes here */
				}
		}
	}
}

typedef struct
{
	parse_pool *pl;
	int tid;
} pool_arg;

static void *pool_main(void *arg)
{
	pool_arg a = *(pool_arg *)arg;
	parse_pool *pl = a.pl;
	unsigned seen = 0;
	free(arg);
	pthread_mutex_lock(&pl->mu);
	for (;;)
	{
		const parse_job *job;
		while (!pl->stop && pl->gen == seen)
			pthread_cond_wait(&pl->cv_work, &pl->mu);
		if (pl->stop)
			break;
		seen = pl->gen;
		if (a.tid >= pl->active)
			continue;
		job = pl->job;
		pthread_mutex_unlock(&pl->mu);
		parse_share(job, a.tid);
		pthread_mutex_lock(&pl->mu);
		if (--pl->pending == 0)
			pthread_cond_signal(&pl->cv_done);
	}
	pthread_mutex_unlock(&pl->mu);
	return NULL;
}

static parse_pool *pool_create(int n)
{
	parse_pool *pl = (parse_pool *)calloc(1, sizeof(*pl));
	int i;
	if (!pl)
		return NULL;
	pthread_mutex_init(&pl->mu, NULL);
	pthread_cond_init(&pl->cv_work, NULL);
	pthread_cond_init(&pl->cv_done, NULL);
	for (i = 0; i < n && i < 64; ++i)
	{
		pool_arg *a = (pool_arg *)malloc(sizeof(*a));
		if (!a)
			break;
		a->pl = pl;
		a->tid = i;
		if (pthread_create(&pl->th[i], NULL, pool_main, a) != 0)
		{
			free(a);
			break;
		}
		pl->n = i + 1;
	}
	return pl;
}

static void pool_destroy(parse_pool *pl)
{
	int i;
	if (!pl)
		return;
	pthread_mutex_lock(&pl->mu);
	pl->stop = 1;
	pthread_cond_broadcast(&pl->cv_work);
	pthread_mutex_unlock(&pl->mu);
	for (i = 0; i < pl->n; ++i)
		pthread_join(pl->th[i], NULL);
	pthread_cond_destroy(&pl->cv_work);
	pthread_cond_destroy(&pl->cv_done);
	pthread_mutex_destroy(&pl->mu);
	free(pl);
}

/* run one parse job on `nt` pool workers and wait for it */
static void pool_run(parse_pool *pl, parse_job *job, int nt)
{
	pthread_mutex_lock(&pl->mu);
	job->nthreads = nt;
	pl->job = job;
	pl->active = nt;
	pl->pending = nt;
	++pl->gen;
	pthread_cond_broadcast(&pl->cv_work);
	while (pl->pending > 0)
		pthread_cond_wait(&pl->cv_done, &pl->mu);
	pthread_mutex_unlock(&pl->mu);
}

/* NULL on allocation failure (the old array is freed by the caller's destroy path) */
static void *grow_buf(void *p, size_t *cap, size_t need, size_t elem)
{
	if (need > *cap)
	{
		size_t nc = *cap ? *cap : 64;
		void *np;
		while (nc < need)
			nc *= 2;
		np = realloc(p, nc * elem);
		if (!np)
			return NULL;
		p = np;
		*cap = nc;
	}
	return p;
}

#define ALIGN_UP(x, a) (((x) + ((a)-1)) & ~((size_t)(a)-1))

/* device bytes image r asks for: coefficient blocks of all frame components, sample planes on the generic path,
   packed pixels, and its file twice (bitstream slab + unstuffed copy) */
static size_t image_device_bytes(const jp_result *r, size_t file_len)
{
	const jcu_image *im = &r->img;
	size_t n = 0;
	int k;
	for (k = 0; k < im->ncomp; ++k)
	{
		const size_t blocks = (size_t)im->comp[k].bw * (size_t)im->comp[k].bh;
		n += blocks * 128;
		if (k < im->decode_n && im->path == JCU_PATH_GENERIC)
			n += blocks * 64;
	}
	n += ALIGN_UP((size_t)im->n_out * (size_t)im->width * (size_t)im->height, 256);
	return n + 2 * file_len + 4096;
}

/* Parse `count` streams into b->res (threaded) and stage them in the pinned slab.  The chunk is cut where its
   images would ask the device for more than b->mem_budget bytes: with allow_defer the images behind the cut are
   left for the caller's next chunk (b->count says how many were taken, at least one), without it they fail
   alone with "outofmem"; an image that exceeds the budget all by itself fails alone either way -- the reference
   fails it through its own malloc (jpeg.c:1643, 2297), the others of the batch decode. */
static int batch_parse(stbi_jpeg_batch *b, const stbi_uc *const *bufs, const int *lens, int count, int req_comp, int allow_defer)
{
	size_t *offs, total = 0;
	int i, nt;
	double t0 = now_ms();
	/* the staging slab, the descriptor block and the device buffers are about to be reused: whatever the previous batch
	   of this object still has in flight (a staged caller may not have waited) must be done first */
	if (b->count > 0)
		jcu_sync(b->ctx);
	batch_release_results(b);
	if (count > b->cap_res)
	{
		jp_result *nr = (jp_result *)realloc(b->res, sizeof(jp_result) * (size_t)count);
		int *ne;
		if (nr)
			b->res = nr;
		ne = (int *)realloc(b->dev_err, sizeof(int) * (size_t)count);
		if (ne)
			b->dev_err = ne;
		if (!nr || !ne)
			return err_zero("outofmem");
		b->cap_res = count;
	}
	if (count > 0)
		memset(b->res, 0, sizeof(jp_result) * (size_t)count);
	b->count = count;
	b->req_comp = req_comp;
	b->dev_err_valid = 0;
	offs = (size_t *)malloc(sizeof(size_t) * (size_t)(count + 1));
	if (!offs)
	{
		b->count = 0;
		return err_zero("outofmem");
	}
	for (i = 0; i < count; ++i)
	{
		offs[i] = total;
		total += ALIGN_UP((size_t)(lens[i] > 0 ? lens[i] : 0) + 16, 16);
	}
	offs[count] = total;
	if (total + 64 > b->slab_cap)
	{
		jcu_host_free(b->slab);
		b->slab_cap = total + total / 4 + 4096;
		b->slab = (unsigned char *)jcu_host_alloc(b->slab_cap);
		if (!b->slab)
		{
			b->slab_cap = 0;
			free(offs);
			b->count = 0;
			return err_zero("outofmem");
		}
	}
	b->slab_bytes = total;

	nt = b->nthreads;
	if (nt > count) nt = count;
	{
		parse_job j = {b, bufs, lens, offs, 1};
		if (nt > 1 && !b->pool)
		{
			b->pool = pool_create(b->nthreads);
			b->pool_owner = b->pool != NULL;
		}
		if (nt > 1 && b->pool && b->pool->n > 1)
			pool_run(b->pool, &j, nt < b->pool->n ? nt : b->pool->n);
		else
			parse_share(&j, 0);
	}
	if (req_comp < 0 || req_comp > 4)
		for (i = 0; i < count; ++i)
			if (!(b->res[i].reason && !strcmp(b->res[i].reason, "unknown image type")))
			{
				b->res[i].ok = 0; /* load_jpeg_image checks req_comp before decoding (jpeg.c:2230) */
				b->res[i].img.ok = 0;
				b->res[i].reason = "bad req_comp";
			}

	/* ---- device memory budget: cut the chunk, or fail the images that do not fit ---- */
	{
		size_t used = 0;
		int taken = 0;
		for (i = 0; i < count; ++i)
		{
			jp_result *r = &b->res[i];
			size_t need;
			if (!r->ok)
				continue;
			need = image_device_bytes(r, (size_t)(lens[i] > 0 ? lens[i] : 0));
			if (need > b->mem_budget || (!allow_defer && used + need > b->mem_budget))
			{
				jp_result_free(r);
				r->ok = 0;
				r->reason = "outofmem";
				continue;
			}
			if (used + need > b->mem_budget && taken > 0)
				break; /* allow_defer: this image and everything behind it go into the caller's next chunk */
			used += need;
			++taken;
		}
		if (i < count)
		{
			int k;
			for (k = i; k < count; ++k)
				jp_result_free(&b->res[k]);
			count = i;
			b->count = count;
			b->slab_bytes = offs[count];
		}
	}

	/* ---- merge into batch-wide descriptor arrays ---- */
	{
		size_t ns = 0, ng = 0, nh = 0, coef = 0, planes = 0, outb = 0;
		double bits = 0, cbytes = 0;
		int waves = 0, clear = 0;
		void *g0, *g1, *g2, *g3;
		for (i = 0; i < count; ++i)
			if (b->res[i].ok)
			{
				ns += (size_t)b->res[i].nscans;
				ng += (size_t)b->res[i].nsegs;
				nh += (size_t)b->res[i].nhuff;
			}
		g0 = grow_buf(b->images, &b->cap_images, (size_t)count + 1, sizeof(jcu_image));
		if (g0) b->images = (jcu_image *)g0;
		g1 = grow_buf(b->scans, &b->cap_scans, ns + 1, sizeof(jcu_scan));
		if (g1) b->scans = (jcu_scan *)g1;
		g2 = grow_buf(b->segs, &b->cap_segs, ng + 1, sizeof(uint32_t));
		if (g2) b->segs = (uint32_t *)g2;
		g3 = grow_buf(b->huff, &b->cap_huff, nh + 1, sizeof(jcu_huff_spec));
		if (g3) b->huff = (jcu_huff_spec *)g3;
		if (!g0 || !g1 || !g2 || !g3)
		{
			free(offs);
			batch_release_results(b);
			return err_zero("outofmem");
		}
		b->flip = g_flip_on_load != 0;
		ns = ng = nh = 0;
		for (i = 0; i < count; ++i)
		{
			jp_result *r = &b->res[i];
			jcu_image *im = &b->images[i];
			int k, s;
			if (!r->ok)
			{
				memset(im, 0, sizeof(*im));
				continue;
			}
			*im = r->img;
			im->flip = (uint32_t)b->flip;
			{
				int q;
				for (k = 0; k < 4; ++k)
					for (q = 0; q < 64; ++q)
						im->qt_out[k][q] = im->progressive ? im->qt[k][q] : 1;
			}
			im->data_off = offs[i];
			im->data_len = (uint32_t)(lens[i] > 0 ? lens[i] : 0);
			for (k = 0; k < im->ncomp; ++k)
			{
				im->comp[k].coef_off = coef;
				coef += (size_t)im->comp[k].bw * im->comp[k].bh * 64;
				if (k < im->decode_n)
				{
					im->comp[k].plane_off = planes;
					if (im->path == JCU_PATH_GENERIC)
						planes += ALIGN_UP((size_t)im->comp[k].bw * im->comp[k].bh * 64, 16);
				}
			}
			im->out_off = outb;
			outb += ALIGN_UP((size_t)im->n_out * (size_t)im->width * (size_t)im->height, 256);
			for (s = 0; s < r->nscans; ++s)
			{
				jcu_scan sc = r->scans[s];
				sc.image = i;
				for (k = 0; k < 4; ++k)
				{
					if (sc.dc_spec[k] >= 0) sc.dc_spec[k] += (int32_t)nh;
					if (sc.ac_spec[k] >= 0) sc.ac_spec[k] += (int32_t)nh;
				}
				sc.seg_first += (uint32_t)ng;
				b->scans[ns++] = sc;
			}
			memcpy(b->segs + ng, r->segs, sizeof(uint32_t) * (size_t)r->nsegs);
			ng += (size_t)r->nsegs;
			memcpy(b->huff + nh, r->huff, sizeof(jcu_huff_spec) * (size_t)r->nhuff);
			nh += (size_t)r->nhuff;
			for (s = 0; s < r->nscans; ++s)
				if (r->scans[s].wave + 1 > waves) waves = r->scans[s].wave + 1;
			/* the slab is reused from batch to batch: blocks that no scan of this image writes (progressive files
			   accumulate; baseline files that lost a restart marker or a whole component) must start at zero */
			if (im->progressive || r->needs_clear) clear = 1;
			bits += r->entropy_bytes * 8.0;
			cbytes += 128.0 * (double)r->coef_blocks;
		}
		b->n_scans = (int)ns;
		b->n_segs = (uint32_t)ng;
		b->n_huff = (int)nh;
		b->n_waves = waves;
		b->coef_elems = coef;
		b->plane_bytes = planes;
		b->out_bytes = outb;
		b->need_clear = clear;
		memset(&b->tm, 0, sizeof(b->tm));
		b->tm.entropy_bits = bits;
		b->tm.coef_bytes = cbytes;
		for (i = 0; i < count; ++i)
			if (b->res[i].ok)
				b->tm.out_bytes += (double)b->images[i].n_out * b->images[i].width * b->images[i].height;
	}
	free(offs);
	b->tm.parse_ms = (float)(now_ms() - t0);
	return 1;
}

static void batch_report(stbi_jpeg_batch *b, int *xs, int *ys, int *comps, int *status)
{
	int i;
	for (i = 0; i < b->count; ++i)
	{
		const jp_result *r = &b->res[i];
		int ok = r->ok && !(b->dev_err_valid && b->dev_err[i]);
		if (xs) xs[i] = r->ok ? r->img.width : 0;
		if (ys) ys[i] = r->ok ? r->img.height : 0;
		if (comps) comps[i] = r->ok ? r->comps_in_file : 0;
		if (status) status[i] = ok ? STBI_JPEG_OK : STBI_JPEG_FAILED;
	}
}

static int batch_upload(stbi_jpeg_batch *b)
{
	jcu_batch jb;
	memset(&jb, 0, sizeof(jb));
	jb.n_images = b->count;
	jb.images = b->images;
	jb.n_scans = b->n_scans;
	jb.scans = b->scans;
	jb.n_segs = b->n_segs;
	jb.seg_start = b->segs;
	jb.n_huff = b->n_huff;
	jb.huff = b->huff;
	jb.bitstream = b->slab;
	jb.bitstream_bytes = b->slab_bytes;
	jb.coef_elems = b->coef_elems;
	jb.plane_bytes = b->plane_bytes;
	jb.out_bytes = b->out_bytes;
	jb.n_waves = b->n_waves;
	jb.need_coef_clear = b->need_clear;
	if (jcu_upload(b->ctx, &jb) != 0)
		return err_zero("outofmem"); /* device allocation or copy failed: see jcu_error_string */
	return 1;
}

API int stbi_jpeg_batch_prepare(stbi_jpeg_batch *b, const stbi_uc *const *bufs, const int *lens, int count,
										  int req_comp, int *xs, int *ys, int *comps, int *status)
{
	int i, nok = 0;
	if (!b || count < 0)
		return -1;
	if (!batch_parse(b, bufs, lens, count, req_comp, 0))
		return -1;
	if (!batch_upload(b))
		return -1;
	batch_report(b, xs, ys, comps, status);
	for (i = 0; i < count; ++i)
		nok += b->res[i].ok;
	return nok;
}

API int stbi_jpeg_batch_run(stbi_jpeg_batch *b)
{
	if (!b)
		return 0;
	b->dev_err_valid = 0;
	if (jcu_run(b->ctx) != 0)
		return err_zero("cuda launch");
	return 1;
}

API int stbi_jpeg_batch_sync(stbi_jpeg_batch *b)
{
	jcu_timings t;
	if (!b)
		return 0;
	if (jcu_sync(b->ctx) != 0)
		return err_zero("cuda sync");
	if (!b->dev_err_valid && b->count > 0)
	{
		if (jcu_image_errors(b->ctx, b->dev_err, b->count) != 0)
			return err_zero("cuda sync");
		b->dev_err_valid = 1;
	}
	if (jcu_get_timings(b->ctx, &t) == 0)
	{
		b->tm.h2d_ms = t.h2d_ms;
		b->tm.entropy_ms = t.entropy_ms;
		b->tm.idct_ms = t.idct_ms;
		b->tm.d2h_ms = t.d2h_ms;
		b->tm.total_ms = t.total_ms;
		b->tm.launches = t.launches;
		b->tm.selfsync_scans = t.ss_scans;
		b->tm.selfsync_rounds = t.ss_rounds;
	}
	return 1;
}

API int stbi_jpeg_batch_status(stbi_jpeg_batch *b, int image)
{
	if (!b || image < 0 || image >= b->count)
		return STBI_JPEG_FAILED;
	return (b->res[image].ok && !(b->dev_err_valid && b->dev_err[image])) ? STBI_JPEG_OK : STBI_JPEG_FAILED;
}

API const char *stbi_jpeg_batch_reason(stbi_jpeg_batch *b, int image)
{
	if (!b || image < 0 || image >= b->count)
		return "bad image index";
	if (!b->res[image].ok)
		return b->res[image].reason ? b->res[image].reason : "corrupt jpeg";
	if (b->dev_err_valid && b->dev_err[image])
		return dev_reason(b->dev_err[image]);
	return NULL;
}

API void *stbi_jpeg_batch_device_pixels(stbi_jpeg_batch *b, int image, size_t *nbytes)
{
	const jcu_image *im;
	if (!b || image < 0 || image >= b->count || !b->res[image].ok)
		return NULL;
	im = &b->images[image];
	if (nbytes)
		*nbytes = (size_t)im->n_out * (size_t)im->width * (size_t)im->height;
	return (unsigned char *)jcu_device_out(b->ctx) + im->out_off;
}

/* enqueue the download of one image (no sync) */
static int batch_enqueue_fetch(stbi_jpeg_batch *b, int image, stbi_uc *out, size_t out_cap)
{
	const jcu_image *im = &b->images[image];
	size_t n = (size_t)im->n_out * (size_t)im->width * (size_t)im->height;
	if (!b->res[image].ok)
		return 0;
	if (n > out_cap)
	{
		b->res[image].ok = 0;
		b->res[image].reason = "output too small";
		return 0;
	}
	return jcu_download(b->ctx, im->out_off, n, out) == 0;
}

/* enqueue the downloads of images [0, n) of the batch into outs[0..n): neighbours that are contiguous both in
   the device pixel slab and in host memory travel as one copy (up to 64 MB), which takes the per-copy launch
   gaps off the PCIe link (one 6 MB copy per 1080p image otherwise) */
static void batch_enqueue_fetch_all(stbi_jpeg_batch *b, int n, stbi_uc *const *outs, const size_t *out_caps)
{
	size_t run_src = 0, run_bytes = 0;
	stbi_uc *run_dst = NULL;
	int i;
	const int merge = !(b->params.no_download_merge & 1);
	for (i = 0; i < n; ++i)
	{
		const jcu_image *im = &b->images[i];
		const size_t bytes = (size_t)im->n_out * (size_t)im->width * (size_t)im->height;
		if (!b->res[i].ok)
			continue;
		if (bytes > out_caps[i])
		{
			b->res[i].ok = 0;
			b->res[i].reason = "output too small";
			continue;
		}
		if (merge && run_bytes && im->out_off == run_src + run_bytes && outs[i] == run_dst + run_bytes &&
			 run_bytes < ((size_t)64 << 20))
		{
			run_bytes += bytes;
			continue;
		}
		if (run_bytes)
			jcu_download(b->ctx, run_src, run_bytes, run_dst);
		run_src = im->out_off;
		run_dst = outs[i];
		run_bytes = bytes;
	}
	if (run_bytes)
		jcu_download(b->ctx, run_src, run_bytes, run_dst);
}

API int stbi_jpeg_batch_fetch(stbi_jpeg_batch *b, int image, stbi_uc *out, size_t out_cap)
{
	if (!b || image < 0 || image >= b->count)
		return 0;
	if (!batch_enqueue_fetch(b, image, out, out_cap))
		return err_zero(b->res[image].reason);
	if (!stbi_jpeg_batch_sync(b))
		return 0;
	if (b->dev_err[image])
		return err_zero(dev_reason(b->dev_err[image]));
	return 1;
}

API int stbi_jpeg_batch_fetch_all(stbi_jpeg_batch *b, stbi_uc *const *outs, const size_t *out_caps)
{
	int i, n = 0;
	if (!b)
		return -1;
	batch_enqueue_fetch_all(b, b->count, outs, out_caps);
	if (!stbi_jpeg_batch_sync(b))
		return -1;
	for (i = 0; i < b->count; ++i)
		if (stbi_jpeg_batch_status(b, i) == STBI_JPEG_OK)
			++n;
	return n;
}

API int stbi_jpeg_batch_debug_guards(stbi_jpeg_batch *b)
{
	int r;
	if (!b)
		return -1;
	r = jcu_debug_guards_ok(b->ctx);
	{
		int pi;
		for (pi = 0; pi < STBI_JPEG_PIPE_MAX - 1; ++pi)
			if (r == 1 && b->pipe[pi])
				r = jcu_debug_guards_ok(b->pipe[pi]->ctx);
	}
	return r;
}

API int stbi_jpeg_batch_timings_get(stbi_jpeg_batch *b, stbi_jpeg_batch_timings *t)
{
	if (!b || !t)
		return 0;
	*t = b->tm;
	return 1;
}

/* ===================================================================================== */
/* one-call batched decode with host buffers                                             */
/* ===================================================================================== */
/* marker walk up to the first scan: a progressive frame (SOF2) with no restart interval in force */
static int looks_serial(const stbi_uc *buf, int len)
{
	int i = 2, progressive = 0, dri = 0;
	if (!buf || len < 4 || buf[0] != 0xff || buf[1] != 0xd8)
		return 0;
	while (i + 4 <= len)
	{
		int m, l;
		if (buf[i] != 0xff)
			return 0;
		m = buf[i + 1];
		if (m == 0xff)
		{
			++i;
			continue;
		}
		l = (buf[i + 2] << 8) | buf[i + 3];
		if (m == 0xda)
			break;
		if (m == 0xc2)
			progressive = 1;
		if (m == 0xdd && l >= 4 && i + 6 <= len)
			dri = (buf[i + 4] << 8) | buf[i + 5];
		i += 2 + l;
	}
	return progressive && dri == 0;
}

/* a chunk that could not be prepared or launched: its images fail alone, the batch goes on */
static void report_chunk_failed(int first, int n, int *xs, int *ys, int *comps, int *status)
{
	int i;
	for (i = first; i < first + n; ++i)
	{
		if (xs) xs[i] = 0;
		if (ys) ys[i] = 0;
		if (comps) comps[i] = 0;
		if (status) status[i] = STBI_JPEG_FAILED;
	}
}

API int stbi_jpeg_decode_batch(stbi_jpeg_batch *b, const stbi_uc *const *bufs, const int *lens, int count,
										 int req_comp, stbi_uc *const *outs, const size_t *out_caps, int *xs, int *ys,
										 int *comps, int *status)
{
	int done = 0, nok = 0, chunk, depth, k0, serial, broken = 0;
	stbi_jpeg_batch_timings acc;
	if (!b || count < 0)
		return -1;
	memset(&acc, 0, sizeof(acc));
	if (!outs)
	{
		/* sizing pass: host parse only (no device memory is asked for, so nothing is cut or failed for its size) */
		const size_t keep = b->mem_budget;
		int ok;
		b->mem_budget = (size_t)-1;
		ok = batch_parse(b, bufs, lens, count, req_comp, 0);
		b->mem_budget = keep;
		if (!ok)
			return -1;
		batch_report(b, xs, ys, comps, status);
		for (done = 0; done < count; ++done)
			nok += b->res[done].ok;
		return nok;
	}
	chunk = b->params.chunk_images > 0 ? b->params.chunk_images : 256;
	/* Software pipeline over chunks, `depth` objects deep (each with its own stream, staging and device
	   buffers): object k % depth parses + uploads + launches + enqueues the download of chunk k on its own
	   stream, then the host moves on to chunk k+1 on the next object and only waits for chunk k when that
	   object is needed again.  Host parse, H2D, kernels and D2H of neighbouring chunks overlap; D2H (3 B/px over
	   PCIe) is the floor.  Three deep, not two: with two, both downloads in flight finish together and the link
	   idles while the host prepares the next chunk (measured: 11 ms of 155 per 1024 x 1080p).  Each object
	   downloads on its own stream: one shared FIFO download stream measured 10-20 % slower. */
	/* Progressive files without restart markers decode one warp per scan (jpeg_entropy_prog.cuh): a chunk takes about
	   as long as its longest scan chain however few images it holds, so such batches are not ramped or split below
	   `chunk`.  Everything else is pipelined as soon as there are enough images to overlap anything. */
	serial = count > 0 && looks_serial(bufs[0], lens[0]) && looks_serial(bufs[count / 2], lens[count / 2]);
	depth = 1;
	if (count > chunk || (count >= 64 && !serial))
	{
		depth = b->params.pipeline_depth > 0 ? b->params.pipeline_depth : 3;
		if (depth > STBI_JPEG_PIPE_MAX) depth = STBI_JPEG_PIPE_MAX;
		for (k0 = 0; k0 < depth - 1; ++k0)
			if (!b->pipe[k0])
			{
				stbi_jpeg_batch_params pp = b->params;
				if (!b->pool && b->nthreads > 1)
				{
					b->pool = pool_create(b->nthreads);
					b->pool_owner = b->pool != NULL;
				}
				b->pipe[k0] = batch_create_with_pool(&pp, b->pool);
				if (!b->pipe[k0])
				{
					depth = k0 + 1;
					break;
				}
				if (b->params.no_download_merge & 2) /* tuning: one FIFO download stream for the whole pipeline */
					jcu_share_download_stream(b->pipe[k0]->ctx, b->ctx);
			}
	}
	{
		stbi_jpeg_batch *obj[STBI_JPEG_PIPE_MAX];
		size_t keep_budget[STBI_JPEG_PIPE_MAX];
		int pend_first[STBI_JPEG_PIPE_MAX] = {-1, -1, -1, -1}, pend_n[STBI_JPEG_PIPE_MAX] = {0, 0, 0, 0}, k = 0, pending = 0;
		const int trace = b->params.trace != 0; /* diagnostic: host timeline of the chunk pipeline on stderr */
		const double tr0 = now_ms();
		double tr_wait0 = 0, tr_wait1 = 0, tr_prep = 0, tr_run = 0;
		obj[0] = b;
		for (k0 = 1; k0 < depth; ++k0)
			obj[k0] = b->pipe[k0 - 1];
		/* the objects in flight share the device: each chunk may ask for its share of the budget */
		for (k0 = 0; k0 < depth; ++k0)
		{
			keep_budget[k0] = obj[k0]->mem_budget;
			if (b->params.device_mem_mb <= 0)
				obj[k0]->mem_budget = b->mem_budget / (size_t)(depth + 1);
		}
		while (done < count || pending > 0)
		{
			const int slot = k % depth;
			stbi_jpeg_batch *s = obj[slot];
			int i;
			if (pend_first[slot] >= 0)
			{
				const int first = pend_first[slot], n = pend_n[slot];
				tr_wait0 = now_ms() - tr0;
				if (!stbi_jpeg_batch_sync(s))
				{
					/* a device fault is sticky: nothing later can succeed.  Wait for the copies of the other objects (they
					   target the caller's buffers, which the caller may free once this call returns), then give up. */
					broken = 1;
					report_chunk_failed(first, n, xs, ys, comps, status);
				}
				else
				{
					tr_wait1 = now_ms() - tr0;
					batch_report(s, xs ? xs + first : NULL, ys ? ys + first : NULL, comps ? comps + first : NULL,
									 status ? status + first : NULL);
					for (i = 0; i < n; ++i)
						if (stbi_jpeg_batch_status(s, i) == STBI_JPEG_OK)
							++nok;
					acc.parse_ms += s->tm.parse_ms;
					acc.h2d_ms += s->tm.h2d_ms;
					acc.entropy_ms += s->tm.entropy_ms;
					acc.idct_ms += s->tm.idct_ms;
					acc.d2h_ms += s->tm.d2h_ms;
					acc.total_ms += s->tm.total_ms;
					acc.entropy_bits += s->tm.entropy_bits;
					acc.coef_bytes += s->tm.coef_bytes;
					acc.out_bytes += s->tm.out_bytes;
					acc.launches += s->tm.launches;
					acc.selfsync_scans += s->tm.selfsync_scans;
					if (s->tm.selfsync_rounds > acc.selfsync_rounds)
						acc.selfsync_rounds = s->tm.selfsync_rounds;
				}
				pend_first[slot] = -1;
				--pending;
			}
			if (done < count && broken)
			{
				report_chunk_failed(done, count - done, xs, ys, comps, status);
				done = count;
			}
			if (done < count)
			{
				/* ramp the first chunks (16, 32, 64, ...) so the first download starts early */
				const int want = (!serial && depth > 1 && k < 4 && (16 << k) < chunk) ? (16 << k) : chunk;
				int n = count - done < want ? count - done : want;
				if (!batch_parse(s, bufs + done, lens + done, n, req_comp, 1))
				{
					report_chunk_failed(done, n, xs, ys, comps, status); /* host allocation failure: these images only */
					done += n;
				}
				else
				{
					n = s->count; /* the memory budget may have cut the chunk; the rest comes next */
					tr_prep = now_ms() - tr0;
					if (!batch_upload(s) || !stbi_jpeg_batch_run(s))
					{
						report_chunk_failed(done, n, xs, ys, comps, status);
						done += n;
					}
					else
					{
						tr_run = now_ms() - tr0;
						batch_enqueue_fetch_all(s, n, outs + done, out_caps + done);
						if (trace)
							fprintf(stderr, "chunk %d slot %d n %d: wait %.2f..%.2f prepared %.2f launched %.2f downloads enqueued %.2f\n", k,
									  slot, n, tr_wait0, tr_wait1, tr_prep, tr_run, now_ms() - tr0);
						pend_first[slot] = done;
						pend_n[slot] = n;
						++pending;
						done += n;
					}
				}
			}
			++k;
		}
		for (k0 = 0; k0 < depth; ++k0)
			obj[k0]->mem_budget = keep_budget[k0];
	}
	b->tm = acc;
	if (broken)
	{
		err_zero("cuda sync");
		return -1;
	}
	return nok;
}

/* ===================================================================================== */
/* all devices of the box behind one call                                                */
/* ===================================================================================== */
#define STBI_JPEG_MULTI_MAX 16
struct stbi_jpeg_multi
{
	int n;
	stbi_jpeg_batch *b[STBI_JPEG_MULTI_MAX];
	int first[STBI_JPEG_MULTI_MAX + 1];
};

API stbi_jpeg_multi *stbi_jpeg_multi_create(const int *devices, int n_devices, const stbi_jpeg_batch_params *params)
{
	stbi_jpeg_multi *m;
	int d, visible = jcu_device_count();
	if (visible < 1)
		return (stbi_jpeg_multi *)err_null("no cuda device");
	if (n_devices <= 0)
	{
		n_devices = visible;
		devices = NULL;
	}
	if (n_devices > STBI_JPEG_MULTI_MAX)
		n_devices = STBI_JPEG_MULTI_MAX;
	m = (stbi_jpeg_multi *)calloc(1, sizeof(*m));
	if (!m)
		return (stbi_jpeg_multi *)err_null("outofmem");
	for (d = 0; d < n_devices; ++d)
	{
		stbi_jpeg_batch_params p;
		memset(&p, 0, sizeof(p));
		if (params)
			p = *params;
		p.device = devices ? devices[d] : d;
		if (p.host_threads <= 0)
		{
			/* the devices share the host cores */
			p.host_threads = online_cpus() / n_devices;
			if (p.host_threads < 2) p.host_threads = 2;
		}
		m->b[d] = stbi_jpeg_batch_create(&p);
		if (!m->b[d])
		{
			stbi_jpeg_multi_destroy(m);
			return NULL; /* reason set by stbi_jpeg_batch_create */
		}
		m->n = d + 1;
	}
	return m;
}

API void stbi_jpeg_multi_destroy(stbi_jpeg_multi *m)
{
	int d;
	if (!m)
		return;
	for (d = 0; d < m->n; ++d)
		stbi_jpeg_batch_destroy(m->b[d]);
	free(m);
}

API int stbi_jpeg_multi_device_count(const stbi_jpeg_multi *m) { return m ? m->n : 0; }

API int stbi_jpeg_multi_last_split(const stbi_jpeg_multi *m, int *first, int cap)
{
	int d;
	if (!m)
		return 0;
	for (d = 0; d <= m->n && d < cap; ++d)
		first[d] = m->first[d];
	return m->n;
}

typedef struct
{
	stbi_jpeg_batch *b;
	const stbi_uc *const *bufs;
	const int *lens;
	int count, req_comp;
	stbi_uc *const *outs;
	const size_t *out_caps;
	int *xs, *ys, *comps, *status;
	int rc;
	const char *reason;
} multi_job;

static void *multi_main(void *arg)
{
	multi_job *j = (multi_job *)arg;
	j->rc = stbi_jpeg_decode_batch(j->b, j->bufs, j->lens, j->count, j->req_comp, j->outs, j->out_caps, j->xs, j->ys, j->comps, j->status);
	j->reason = j->rc < 0 ? tl_reason : NULL; /* the failure string is thread-local: hand it to the caller's thread */
	return NULL;
}

API int stbi_jpeg_decode_batch_multi(stbi_jpeg_multi *m, const stbi_uc *const *bufs, const int *lens, int count, int req_comp,
												 stbi_uc *const *outs, const size_t *out_caps, int *xs, int *ys, int *comps, int *status)
{
	multi_job jobs[STBI_JPEG_MULTI_MAX];
	pthread_t th[STBI_JPEG_MULTI_MAX];
	int started[STBI_JPEG_MULTI_MAX];
	double total = 0, run = 0;
	int d, i, nok = 0, failed = 0;
	if (!m || m->n < 1 || count < 0)
		return -1;
	/* contiguous ranges, balanced by compressed bytes: range d ends where the running sum passes (d+1)/n of the total */
	for (i = 0; i < count; ++i)
		total += lens[i] > 0 ? lens[i] : 0;
	m->first[0] = 0;
	d = 1;
	for (i = 0; i < count && d < m->n; ++i)
	{
		run += lens[i] > 0 ? lens[i] : 0;
		while (d < m->n && run >= total * d / m->n)
			m->first[d++] = i + 1;
	}
	while (d <= m->n)
		m->first[d++] = count;
	for (d = 0; d < m->n; ++d)
	{
		const int f = m->first[d], n = m->first[d + 1] - f;
		multi_job *j = &jobs[d];
		memset(j, 0, sizeof(*j));
		j->b = m->b[d];
		j->bufs = bufs + f;
		j->lens = lens + f;
		j->count = n;
		j->req_comp = req_comp;
		j->outs = outs ? outs + f : NULL;
		j->out_caps = out_caps ? out_caps + f : NULL;
		j->xs = xs ? xs + f : NULL;
		j->ys = ys ? ys + f : NULL;
		j->comps = comps ? comps + f : NULL;
		j->status = status ? status + f : NULL;
		started[d] = 0;
		if (n <= 0)
			continue;
		if (m->n > 1 && pthread_create(&th[d], NULL, multi_main, j) == 0)
			started[d] = 1;
		else
			multi_main(j); /* one device, or no thread to be had: the caller's thread does it */
	}
	for (d = 0; d < m->n; ++d)
	{
		if (started[d])
			pthread_join(th[d], NULL);
		if (jobs[d].count <= 0)
			continue;
		if (jobs[d].rc < 0)
		{
			failed = 1;
			err_zero(jobs[d].reason ? jobs[d].reason : "cuda sync");
		}
		else
			nok += jobs[d].rc;
	}
	return failed ? -1 : nok;
}

/* ===================================================================================== */
/* per-image entry points                                                                */
/* ===================================================================================== */
/* The reference allocates fresh decoder state per call (jpeg.c:2445-2450) and is re-entrant apart from the failure
   string; so are these: every call borrows a decoder object (own stream, own device buffers, pre-sized so that an
   ordinary image meets no cudaMalloc) from a pool that grows with the number of threads inside the library at
   once, up to PI_POOL_MAX; beyond that callers queue for a free object. */
#define PI_POOL_MAX 16
static pthread_mutex_t g_pi_lock = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_pi_cv = PTHREAD_COND_INITIALIZER;
static stbi_jpeg_batch *g_pi_free[PI_POOL_MAX];
static int g_pi_nfree, g_pi_created;
static int g_default_device;

API void stbi_jpeg_set_default_device(int device) { g_default_device = device < 0 ? 0 : device; }

static stbi_jpeg_batch *pi_acquire(void)
{
	stbi_jpeg_batch *b = NULL;
	pthread_mutex_lock(&g_pi_lock);
	for (;;)
	{
		if (g_pi_nfree > 0)
		{
			b = g_pi_free[--g_pi_nfree];
			if (b->params.device == g_default_device)
				break;
			/* the default device changed since this object was made */
			--g_pi_created;
			pthread_mutex_unlock(&g_pi_lock);
			stbi_jpeg_batch_destroy(b);
			pthread_mutex_lock(&g_pi_lock);
			b = NULL;
			continue;
		}
		if (g_pi_created < PI_POOL_MAX)
		{
			stbi_jpeg_batch_params p;
			++g_pi_created;
			memset(&p, 0, sizeof(p));
			p.device = g_default_device;
			p.host_threads = 1;
			pthread_mutex_unlock(&g_pi_lock);
			b = stbi_jpeg_batch_create(&p);
			if (b) /* room for a 2048 x 2048 image in any sampling before the first call needs it */
				jcu_reserve(b->ctx, (size_t)4 << 20, (size_t)3 * 2048 * 2048, (size_t)4 * 2048 * 2048);
			else
			{
				pthread_mutex_lock(&g_pi_lock);
				--g_pi_created;
				pthread_cond_signal(&g_pi_cv);
				pthread_mutex_unlock(&g_pi_lock);
			}
			return b;
		}
		pthread_cond_wait(&g_pi_cv, &g_pi_lock);
	}
	pthread_mutex_unlock(&g_pi_lock);
	return b;
}

static void pi_release(stbi_jpeg_batch *b)
{
	pthread_mutex_lock(&g_pi_lock);
	g_pi_free[g_pi_nfree++] = b;
	pthread_cond_signal(&g_pi_cv);
	pthread_mutex_unlock(&g_pi_lock);
}

/* 16-bit / float variants: state of the reference's setters, convert.c:402-411 */
static float g_l2h_gamma = 2.2f, g_l2h_scale = 1.0f;
enum { WIDE_NONE = 0, WIDE_16 = 1, WIDE_FLOAT = 2 };

/* widen = WIDE_NONE: 8-bit samples; WIDE_16 / WIDE_FLOAT: the widening kernel of jpeg_widen.cuh runs behind the
   output stage and the wide samples are what crosses the bus */
static void *load_from_buffer_as(const stbi_uc *buffer, size_t len, int *x, int *y, int *comp, int req_comp, int widen)
{
	jp_result probe;
	stbi_uc *out = NULL;
	stbi_jpeg_batch *g_default;
	int ilen, xs = 0, ys = 0, cs = 0, st = 0;
	/* stbi__load_main (image_api.c:10-13,55): only JPEG is recognised, by its SOI */
	if (!jp_parse(buffer, len, req_comp, JP_MODE_TYPE, &probe))
		return (stbi_uc *)err_null("unknown image type");
	if (req_comp < 0 || req_comp > 4)
		return (stbi_uc *)err_null("bad req_comp"); /* jpeg.c:2230 */
	if (len > 0x7fffffffu)
		return (stbi_uc *)err_null("too large");
	ilen = (int)len;

	g_default = pi_acquire();
	if (g_default)
	{
		const stbi_uc *bufs[1];
		bufs[0] = buffer;
		if (stbi_jpeg_batch_prepare(g_default, bufs, &ilen, 1, req_comp, &xs, &ys, &cs, &st) == 1 &&
			 stbi_jpeg_batch_run(g_default))
		{
			int n = req_comp ? req_comp : cs;
			size_t bytes = (size_t)n * (size_t)xs * (size_t)ys;
			if (widen == WIDE_NONE)
			{
				out = (stbi_uc *)malloc(bytes + 1); /* +1 as stbi__malloc_mad3(n, x, y, 1), jpeg.c:2293 */
				if (!out)
					err_null("outofmem");
				else if (!stbi_jpeg_batch_fetch(g_default, 0, out, bytes))
				{
					free(out);
					out = NULL;
				}
			}
			else
			{
				const int elem = widen == WIDE_16 ? 2 : 4;
				const int nonalpha = (n & 1) ? n : n - 1; /* common.c:404-407 */
				float lut[512];
				int k;
				if (widen == WIDE_FLOAT)
					for (k = 0; k < 256; ++k)
					{
						/* the reference's per-sample expressions (common.c:412, :417) depend on the byte alone */
						lut[k] = (float)(pow((unsigned char)k / 255.0f, g_l2h_gamma) * g_l2h_scale);
						lut[256 + k] = (unsigned char)k / 255.0f;
					}
				out = (stbi_uc *)malloc(bytes > 0 ? bytes * (size_t)elem : 1);
				if (!out)
					err_null("outofmem");
				else if (jcu_widen(g_default->ctx, g_default->images[0].out_off, bytes, elem, n, nonalpha, lut) != 0 ||
							jcu_download_wide(g_default->ctx, bytes * (size_t)elem, out) != 0 || !stbi_jpeg_batch_sync(g_default))
				{
					err_null("cuda widen");
					free(out);
					out = NULL;
				}
				else if (g_default->dev_err[0])
				{
					err_null(dev_reason(g_default->dev_err[0]));
					free(out);
					out = NULL;
				}
			}
		}
		else if (g_default->count == 1 && !g_default->res[0].ok)
			err_null(g_default->res[0].reason); /* NULL reason: the reference left it unchanged too */
	}
	if (g_default)
		pi_release(g_default);
	if (!out)
		return NULL;
	if (x) *x = xs;
	if (y) *y = ys;
	if (comp) *comp = cs;
	return out;
}

static stbi_uc *load_from_buffer(const stbi_uc *buffer, size_t len, int *x, int *y, int *comp, int req_comp)
{
	return (stbi_uc *)load_from_buffer_as(buffer, len, x, y, comp, req_comp, WIDE_NONE);
}

API stbi_uc *stbi_load_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp, int req_comp)
{
	return load_from_buffer(buffer, len > 0 ? (size_t)len : 0, x, y, comp, req_comp);
}

/* The GPU path needs the whole image in memory, the reference refills a 128-byte window (common.c:10) and so reads
   at most 128 bytes past the image.  A callback source may be a stream with more behind the image (or a socket
   that blocks), so it is read in growing steps only until the marker walk reaches the EOI of the image: whenever
   the buffer has doubled, the host parser is tried on what is there, and a parse that succeeds never looks at
   later bytes.  A file that is damaged or cut short is read to its end, like the reference would. */
static stbi_uc *drain_callbacks(stbi_io_callbacks const *c, void *user, size_t *len, int mode)
{
	size_t cap = 1 << 12, n = 0;
	stbi_uc *buf = (stbi_uc *)malloc(cap);
	while (buf)
	{
		int want = (int)(cap - n > 0x40000000u ? 0x40000000u : cap - n);
		int got = c->read(user, (char *)buf + n, want);
		if (got <= 0)
			break;
		n += (size_t)got;
		if (n == cap)
		{
			stbi_uc *nb;
			jp_result r;
			int complete;
			memset(&r, 0, sizeof(r));
			complete = jp_parse(buf, n, 0, mode, &r);
			jp_result_free(&r);
			if (complete)
				break;
			nb = (stbi_uc *)realloc(buf, cap * 2);
			if (!nb)
			{
				free(buf);
				return NULL;
			}
			buf = nb;
			cap *= 2;
		}
	}
	*len = n;
	return buf;
}

API stbi_uc *stbi_load_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp, int req_comp)
{
	size_t len = 0;
	stbi_uc *buf = drain_callbacks(clbk, user, &len, JP_MODE_LOAD), *r;
	if (!buf)
		return (stbi_uc *)err_null("outofmem");
	r = load_from_buffer(buf, len, x, y, comp, req_comp);
	free(buf);
	return r;
}

static stbi_uc *drain_file(FILE *f, size_t *len)
{
	size_t cap = 1 << 16, n = 0;
	stbi_uc *buf = (stbi_uc *)malloc(cap);
	while (buf)
	{
		size_t got = fread(buf + n, 1, cap - n, f);
		n += got;
		if (n < cap)
			break;
		{
			stbi_uc *nb = (stbi_uc *)realloc(buf, cap * 2);
			if (!nb)
			{
				free(buf);
				return NULL;
			}
			buf = nb;
			cap *= 2;
		}
	}
	*len = n;
	return buf;
}

/* The reference reads the FILE through a 128-byte window and, after a successful load, seeks back by what it had
   buffered but not used (convert.c:208, :222): the stream is left just behind the image's EOI marker, so that
   images stored back to back can be loaded one after the other.  Here the whole rest of the file was drained. */
static void seek_behind_image(FILE *f, long start, const stbi_uc *buf, size_t len)
{
	jp_result r;
	memset(&r, 0, sizeof(r));
	if (start >= 0 && jp_parse(buf, len, 0, JP_MODE_LOAD, &r))
		fseek(f, start + (long)r.consumed, SEEK_SET);
	jp_result_free(&r);
}

API stbi_uc *stbi_load_from_file(FILE *f, int *x, int *y, int *comp, int req_comp)
{
	size_t len = 0;
	const long start = ftell(f);
	stbi_uc *buf = drain_file(f, &len), *r;
	if (!buf)
		return (stbi_uc *)err_null("outofmem");
	r = load_from_buffer(buf, len, x, y, comp, req_comp);
	if (r)
		seek_behind_image(f, start, buf, len);
	free(buf);
	return r;
}

API stbi_uc *stbi_load(char const *filename, int *x, int *y, int *comp, int req_comp)
{
	FILE *f = fopen(filename, "rb");
	stbi_uc *r;
	if (!f)
		return (stbi_uc *)err_null("can't fopen");
	r = stbi_load_from_file(f, x, y, comp, req_comp);
	fclose(f);
	return r;
}

/* ---- 16-bit and float variants: the JPEG codec is 8-bit and the reference widens its (already
 *      flipped) result on the host -- stbi__load_and_postprocess_16bit convert.c:108-133 with stbi__convert_8_to_16
 *      :18-35, stbi__loadf_main :286-304 with stbi__ldr_to_hdr common.c:391-424.  Here both are store epilogues on
 *      the device (jpeg_widen.cuh) behind the output kernels, which also apply the flip. ---- */
static float g_h2l_gamma_i = 1.0f / 2.2f, g_h2l_scale_i = 1.0f; /* convert.c:408; no HDR source format on this path */
API void stbi_ldr_to_hdr_gamma(float gamma) { g_l2h_gamma = gamma; }
API void stbi_ldr_to_hdr_scale(float scale) { g_l2h_scale = scale; }
API void stbi_hdr_to_ldr_gamma(float gamma) { g_h2l_gamma_i = 1 / gamma; }
API void stbi_hdr_to_ldr_scale(float scale) { g_h2l_scale_i = 1 / scale; }

static void *wide_from_memory(stbi_uc const *buffer, size_t len, int *x, int *y, int *comp, int req_comp, int widen)
{
	void *r = load_from_buffer_as(buffer, len, x, y, comp, req_comp, widen);
	if (!r && widen == WIDE_FLOAT)
		err_null("unknown image type"); /* stbi__loadf_main replaces the codec's reason, convert.c:303 */
	return r;
}
static void *wide_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp, int req_comp, int widen)
{
	size_t len = 0;
	stbi_uc *buf = drain_callbacks(clbk, user, &len, JP_MODE_LOAD);
	void *r;
	if (!buf)
		return err_null("outofmem");
	r = wide_from_memory(buf, len, x, y, comp, req_comp, widen);
	free(buf);
	return r;
}
static void *wide_from_file(FILE *f, int *x, int *y, int *comp, int req_comp, int widen)
{
	size_t len = 0;
	const long start = ftell(f);
	stbi_uc *buf = drain_file(f, &len);
	void *r;
	if (!buf)
		return err_null("outofmem");
	r = wide_from_memory(buf, len, x, y, comp, req_comp, widen);
	if (r && widen == WIDE_16)
		seek_behind_image(f, start, buf, len); /* convert.c:222; stbi_loadf_from_file leaves the stream where the reader stopped */
	free(buf);
	return r;
}
static void *wide_from_path(char const *filename, int *x, int *y, int *comp, int req_comp, int widen)
{
	FILE *f = fopen(filename, "rb");
	void *r;
	if (!f)
		return err_null("can't fopen");
	r = wide_from_file(f, x, y, comp, req_comp, widen);
	fclose(f);
	return r;
}

API stbi_us *stbi_load_16_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp, int req_comp)
{
	return (stbi_us *)wide_from_memory(buffer, len > 0 ? (size_t)len : 0, x, y, comp, req_comp, WIDE_16);
}
API stbi_us *stbi_load_16_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp, int req_comp)
{
	return (stbi_us *)wide_from_callbacks(clbk, user, x, y, comp, req_comp, WIDE_16);
}
API stbi_us *stbi_load_from_file_16(FILE *f, int *x, int *y, int *comp, int req_comp)
{
	return (stbi_us *)wide_from_file(f, x, y, comp, req_comp, WIDE_16);
}
API stbi_us *stbi_load_16(char const *filename, int *x, int *y, int *comp, int req_comp)
{
	return (stbi_us *)wide_from_path(filename, x, y, comp, req_comp, WIDE_16);
}
API float *stbi_loadf_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp, int req_comp)
{
	return (float *)wide_from_memory(buffer, len > 0 ? (size_t)len : 0, x, y, comp, req_comp, WIDE_FLOAT);
}
API float *stbi_loadf_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp, int req_comp)
{
	return (float *)wide_from_callbacks(clbk, user, x, y, comp, req_comp, WIDE_FLOAT);
}
API float *stbi_loadf_from_file(FILE *f, int *x, int *y, int *comp, int req_comp)
{
	return (float *)wide_from_file(f, x, y, comp, req_comp, WIDE_FLOAT);
}
API float *stbi_loadf(char const *filename, int *x, int *y, int *comp, int req_comp)
{
	return (float *)wide_from_path(filename, x, y, comp, req_comp, WIDE_FLOAT);
}
/* a JPEG is neither HDR nor 16-bit (convert.c:345-400, image_api.c:96-145) */
API int stbi_is_hdr_from_memory(stbi_uc const *buffer, int len) { (void)buffer; (void)len; return 0; }
API int stbi_is_hdr_from_callbacks(stbi_io_callbacks const *clbk, void *user) { (void)clbk; (void)user; return 0; }
API int stbi_is_hdr_from_file(FILE *f) { (void)f; return 0; }
API int stbi_is_hdr(char const *filename) { (void)filename; return 0; }
API int stbi_is_16_bit_from_callbacks(stbi_io_callbacks const *clbk, void *user) { (void)clbk; (void)user; return 0; }
API int stbi_is_16_bit_from_file(FILE *f) { (void)f; return 0; }
API int stbi_is_16_bit(char const *filename) { (void)filename; return 0; }

/* ===================================================================================== */
/* encoder: host side of jpeg_encode.cuh; replaces stbi_write_jpg_core,  */
/* codec/jpeg_write.c:172-366, with the same byte stream                                   */
/* ===================================================================================== */
/* ITU-T T.81 Annex K.3 Huffman tables (the ones jpeg_write.c:175-200 writes into every file): code counts
   per length 1..16, then the symbols in code order */
static const unsigned char k_dc_lum_bits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const unsigned char k_dc_chr_bits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const unsigned char k_dc_vals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const unsigned char k_ac_lum_bits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const unsigned char k_ac_chr_bits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const unsigned char k_ac_lum_vals[162] = {
	0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32, 0x81, 0x91,
	0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a,
	0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53,
	0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79,
	0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5,
	0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9,
	0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2,
	0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
static const unsigned char k_ac_chr_vals[162] = {
	0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81, 0x08, 0x14,
	0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17,
	0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a,
	0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78,
	0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3,
	0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7,
	0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2,
	0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
/* luminance / chrominance quantisers of Annex K.1, natural order, and the raster -> zig-zag index map */
static const int k_qt_lum[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57,
											69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64,
											81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const int k_qt_chr[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
											99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
											99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
static const unsigned char k_zigzag[64] = {0,	1,	 5,  6,	14, 15, 27, 28, 2,  4,	7,	 13, 16, 26, 29, 42, 3,	 8,  12, 17, 25, 30,
														 41, 43, 9,  11, 18, 24, 31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38,
														 46, 51, 55, 60, 21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63};

static int g_flip_on_write;
API void stbi_flip_vertically_on_write(int flag) { g_flip_on_write = flag; }

/* canonical codes (T.81 Annex C) of one table: out[symbol] = code | length << 16 */
static void enc_codes(const unsigned char bits[16], const unsigned char *vals, uint32_t out[256])
{
	unsigned code = 0;
	int l, i, k = 0;
	memset(out, 0, 256 * sizeof(uint32_t));
	for (l = 1; l <= 16; ++l)
	{
		for (i = 0; i < bits[l - 1]; ++i)
			out[vals[k++]] = code++ | ((uint32_t)l << 16);
		code <<= 1;
	}
}

/* quality -> quantiser tables (zig-zag order, as stored in DQT) and the float scale tables (jpeg_write.c:210-235) */
static void enc_tables(int quality, int *subsample, unsigned char ytab[64], unsigned char uvtab[64], float fd_y[64], float fd_uv[64])
{
	static const float aasf[8] = {1.0f * 2.828427125f, 1.387039845f * 2.828427125f, 1.306562965f * 2.828427125f, 1.175875602f * 2.828427125f,
											1.0f * 2.828427125f, 0.785694958f * 2.828427125f, 0.541196100f * 2.828427125f, 0.275899379f * 2.828427125f};
	int i, row, col, k;
	quality = quality ? quality : 90;
	*subsample = quality <= 90 ? 1 : 0;
	quality = quality < 1 ? 1 : quality > 100 ? 100 : quality;
	quality = quality < 50 ? 5000 / quality : 200 - quality * 2;
	for (i = 0; i < 64; ++i)
	{
		int yti = (k_qt_lum[i] * quality + 50) / 100, uvti = (k_qt_chr[i] * quality + 50) / 100;
		ytab[k_zigzag[i]] = (unsigned char)(yti < 1 ? 1 : yti > 255 ? 255 : yti);
		uvtab[k_zigzag[i]] = (unsigned char)(uvti < 1 ? 1 : uvti > 255 ? 255 : uvti);
	}
	for (row = 0, k = 0; row < 8; ++row)
		for (col = 0; col < 8; ++col, ++k)
		{
			fd_y[k] = 1 / (ytab[k_zigzag[k]] * aasf[row] * aasf[col]);
			fd_uv[k] = 1 / (uvtab[k_zigzag[k]] * aasf[row] * aasf[col]);
		}
}

/* SOI, APP0, DQT, SOF0, DHT, SOS exactly as jpeg_write.c:238-262 writes them; returns the length */
static size_t enc_header(unsigned char *h, int width, int height, int subsample, const unsigned char ytab[64], const unsigned char uvtab[64])
{
	static const unsigned char head0[] = {0xFF, 0xD8, 0xFF, 0xE0, 0, 0x10, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0, 0xFF, 0xDB, 0, 0x84, 0};
	static const unsigned char head2[] = {0xFF, 0xDA, 0, 0xC, 3, 1, 0, 2, 0x11, 3, 0x11, 0, 0x3F, 0};
	const unsigned char head1[] = {0xFF, 0xC0, 0, 0x11, 8, (unsigned char)(height >> 8), (unsigned char)(height & 0xff), (unsigned char)(width >> 8),
											 (unsigned char)(width & 0xff), 3, 1, (unsigned char)(subsample ? 0x22 : 0x11), 0, 2, 0x11, 1, 3, 0x11, 1,
											 0xFF, 0xC4, 0x01, 0xA2, 0};
	unsigned char *p = h;
#define ENC_PUT(src, n) do { memcpy(p, (src), (n)); p += (n); } while (0)
	ENC_PUT(head0, sizeof(head0));
	ENC_PUT(ytab, 64);
	*p++ = 1;
	ENC_PUT(uvtab, 64);
	ENC_PUT(head1, sizeof(head1));
	ENC_PUT(k_dc_lum_bits, 16);
	ENC_PUT(k_dc_vals, 12);
	*p++ = 0x10;
	ENC_PUT(k_ac_lum_bits, 16);
	ENC_PUT(k_ac_lum_vals, 162);
	*p++ = 1;
	ENC_PUT(k_dc_chr_bits, 16);
	ENC_PUT(k_dc_vals, 12);
	*p++ = 0x11;
	ENC_PUT(k_ac_chr_bits, 16);
	ENC_PUT(k_ac_chr_vals, 162);
	ENC_PUT(head2, sizeof(head2));
#undef ENC_PUT
	return (size_t)(p - h);
}

static pthread_mutex_t g_enc_lock = PTHREAD_MUTEX_INITIALIZER;
static __thread int g_enc_pixels_on_device; /* set by stbi_jpeg_encode_batch_device around its call */
static jcu_ctx *g_enc_ctx[16];

/* New batched entry point: encodes `count` images (comps interleaved bytes per pixel, rows top-down) on `device`
   into malloc'ed files (caller frees each with stbi_image_free).  pixels_on_device: the pointers are device
   pointers into ONE allocation starting at pixels[0].  Returns the number of files produced; kernel_ms (optional)
   receives the device time of the kernels. */
API int stbi_jpeg_encode_batch(int device, const stbi_uc *const *pixels, const int *xs, const int *ys, const int *comps, int count,
										 int quality, stbi_uc **outs, int *out_lens, float *kernel_ms)
{
	jcu_enc_image *imgs;
	unsigned char *hdrs;
	uint32_t huff[4][256], *lens;
	size_t pix_bytes = 0, out_bytes = 0, nblocks = 0;
	int i, nok = 0, rc;
	jcu_ctx *ctx;
	if (count <= 0)
		return 0;
	if (device < 0 || device >= 16 || jcu_device_count() <= device)
		return err_zero("no cuda device");
	pthread_mutex_lock(&g_enc_lock);
	if (!g_enc_ctx[device] && jcu_ctx_create(device, 0, 0, &g_enc_ctx[device]) != 0)
		g_enc_ctx[device] = NULL;
	ctx = g_enc_ctx[device];
	if (!ctx)
	{
		pthread_mutex_unlock(&g_enc_lock);
		return err_zero("no cuda device");
	}
	imgs = (jcu_enc_image *)calloc((size_t)count, sizeof(*imgs));
	hdrs = (unsigned char *)calloc((size_t)count, JCU_ENC_HDR_STRIDE);
	lens = (uint32_t *)calloc((size_t)count, sizeof(uint32_t));
	enc_codes(k_dc_lum_bits, k_dc_vals, huff[0]);
	enc_codes(k_dc_chr_bits, k_dc_vals, huff[1]);
	enc_codes(k_ac_lum_bits, k_ac_lum_vals, huff[2]);
	enc_codes(k_ac_chr_bits, k_ac_chr_vals, huff[3]);
	for (i = 0; i < count; ++i)
	{
		jcu_enc_image *im = &imgs[i];
		unsigned char ytab[64], uvtab[64];
		int sub, m;
		outs[i] = NULL;
		out_lens[i] = 0;
		/* jpeg_write.c:204: !data || !width || !height || comp > 4 || comp < 1 -> 0; 16-bit SOF fields */
		if (!pixels[i] || xs[i] <= 0 || ys[i] <= 0 || comps[i] < 1 || comps[i] > 4 || xs[i] > 65535 || ys[i] > 65535)
			continue;
		enc_tables(quality, &sub, ytab, uvtab, im->fdtbl_y, im->fdtbl_uv);
		m = sub ? 16 : 8;
		im->width = xs[i];
		im->height = ys[i];
		im->comp = comps[i];
		im->subsample = sub;
		im->flip = g_flip_on_write;
		im->mcu_x = (xs[i] + m - 1) / m;
		im->mcu_y = (ys[i] + m - 1) / m;
		im->nblocks = im->mcu_x * im->mcu_y * (sub ? 6 : 3);
		im->pix_off = pix_bytes;
		im->blk_first = nblocks;
		im->hdr_len = (uint32_t)enc_header(hdrs + (size_t)i * JCU_ENC_HDR_STRIDE, xs[i], ys[i], sub, ytab, uvtab);
		/* worst case: every block full (224 bytes) and every byte stuffed */
		im->out_cap = (uint32_t)(im->hdr_len + 2 + (size_t)im->nblocks * 448 + 16);
		im->out_off = out_bytes;
		out_bytes += ALIGN_UP((size_t)im->out_cap, 256);
		pix_bytes += ALIGN_UP((size_t)xs[i] * (size_t)ys[i] * (size_t)comps[i], 16);
		nblocks += (size_t)im->nblocks;
	}
	if (g_enc_pixels_on_device)
	{
		/* device pixels: offsets are relative to pixels[0] */
		for (i = 0; i < count; ++i)
			if (imgs[i].nblocks)
				imgs[i].pix_off = (uint64_t)((uintptr_t)pixels[i] - (uintptr_t)pixels[0]);
		rc = jcu_encode(ctx, imgs, count, (const uint8_t *)pixels[0], NULL, pix_bytes, 1, hdrs, &huff[0][0], out_bytes, lens,
							 kernel_ms);
	}
	else
		rc = jcu_encode(ctx, imgs, count, NULL, (const uint8_t *const *)pixels, pix_bytes, 0, hdrs, &huff[0][0], out_bytes, lens, kernel_ms);
	if (rc == 0)
	{
		for (i = 0; i < count; ++i)
			if (imgs[i].nblocks && lens[i])
			{
				outs[i] = (stbi_uc *)malloc(lens[i]);
				if (!outs[i] || jcu_encode_fetch(ctx, imgs[i].out_off, lens[i], outs[i]) != 0)
				{
					free(outs[i]);
					outs[i] = NULL;
				}
			}
		if (jcu_encode_sync(ctx) == 0)
			for (i = 0; i < count; ++i)
				if (outs[i])
				{
					out_lens[i] = (int)lens[i];
					++nok;
				}
	}
	free(imgs); free(hdrs); free(lens);
	pthread_mutex_unlock(&g_enc_lock);
	return nok;
}

/* same, with the source pixels already in device memory (all inside one allocation, pixels[0] the lowest address) */
API int stbi_jpeg_encode_batch_device(int device, const stbi_uc *const *pixels, const int *xs, const int *ys, const int *comps, int count,
												  int quality, stbi_uc **outs, int *out_lens, float *kernel_ms)
{
	int r;
	g_enc_pixels_on_device = 1;
	r = stbi_jpeg_encode_batch(device, pixels, xs, ys, comps, count, quality, outs, out_lens, kernel_ms);
	g_enc_pixels_on_device = 0;
	return r;
}

/* stbi_write_jpg_to_func (jpeg_write.c:368): same bytes, handed to `func` in one call (the reference makes many
   small calls; their concatenation is what a caller can observe) */
API int stbi_write_jpg_to_func(stbi_write_func *func, void *context, int x, int y, int comp, const void *data, int quality)
{
	const stbi_uc *px[1];
	stbi_uc *out[1] = {NULL};
	int len[1] = {0};
	if (!data || !x || !y || comp > 4 || comp < 1 || !func)
		return 0;
	px[0] = (const stbi_uc *)data;
	if (stbi_jpeg_encode_batch(g_default_device, px, &x, &y, &comp, 1, quality, out, len, NULL) != 1)
		return 0;
	func(context, out[0], len[0]);
	free(out[0]);
	return 1;
}

static void enc_file_write(void *context, void *data, int size) { fwrite(data, 1, (size_t)size, (FILE *)context); }
API int stbi_write_jpg(char const *filename, int x, int y, int comp, const void *data, int quality)
{
	FILE *f = fopen(filename, "wb");
	int r;
	if (!f)
		return 0;
	r = stbi_write_jpg_to_func(enc_file_write, f, x, y, comp, data, quality);
	fclose(f);
	return r;
}

/* ---- info: header-only, no GPU (stbi__jpeg_info_raw, jpeg.c:2466) ---- */
API int stbi_info_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp)
{
	jp_result r;
	if (!jp_parse(buffer, len > 0 ? (size_t)len : 0, 0, JP_MODE_HEADER, &r))
		return err_zero("unknown image type"); /* stbi__info_main falls through every codec */
	if (x) *x = r.img.width;
	if (y) *y = r.img.height;
	if (comp) *comp = r.comps_in_file;
	return 1;
}

API int stbi_info_from_callbacks(stbi_io_callbacks const *c, void *user, int *x, int *y, int *comp)
{
	size_t len = 0;
	stbi_uc *buf = drain_callbacks(c, user, &len, JP_MODE_HEADER); /* stbi__jpeg_info_raw stops behind the frame header */
	int ok;
	if (!buf)
		return err_zero("outofmem");
	ok = stbi_info_from_memory(buf, len > 0x7fffffffu ? 0x7fffffff : (int)len, x, y, comp);
	free(buf);
	return ok;
}

API int stbi_info_from_file(FILE *f, int *x, int *y, int *comp)
{
	long pos = ftell(f);
	size_t len = 0;
	stbi_uc *buf = drain_file(f, &len);
	int ok;
	if (!buf)
		return err_zero("outofmem");
	ok = stbi_info_from_memory(buf, len > 0x7fffffffu ? 0x7fffffff : (int)len, x, y, comp);
	free(buf);
	fseek(f, pos, SEEK_SET); /* image_api.c:92 */
	return ok;
}

API int stbi_info(char const *filename, int *x, int *y, int *comp)
{
	FILE *f = fopen(filename, "rb");
	int ok;
	if (!f)
		return err_zero("can't fopen");
	ok = stbi_info_from_file(f, x, y, comp);
	fclose(f);
	return ok;
}

API int stbi_is_16_bit_from_memory(stbi_uc const *buffer, int len)
{
	(void)buffer;
	(void)len;
	return 0;
}

API int stbi_jpeg_inspect_scans(stbi_uc const *buffer, int len, int *waves, int cap)
{
	jp_result r;
	size_t n = len > 0 ? (size_t)len : 0;
	int i, ns = 0;
	if (jp_parse(buffer, n, 0, JP_MODE_LOAD, &r))
	{
		ns = r.nscans;
		for (i = 0; i < ns && i < cap; ++i)
			waves[i] = r.scans[i].wave;
	}
	else
		err_zero(r.reason);
	jp_result_free(&r);
	return ns;
}

API int stbi_jpeg_inspect(stbi_uc const *buffer, int len, int req_comp, long long info[16])
{
	jp_result r;
	size_t n = len > 0 ? (size_t)len : 0;
	int ok;
	memset(info, 0, sizeof(long long) * 16);
	if (!jp_parse(buffer, n, req_comp, JP_MODE_TYPE, &r))
		return err_zero("unknown image type");
	if (req_comp < 0 || req_comp > 4)
		return err_zero("bad req_comp");
	ok = jp_parse(buffer, n, req_comp, JP_MODE_LOAD, &r);
	if (ok)
	{
		info[0] = r.img.width;
		info[1] = r.img.height;
		info[2] = r.comps_in_file;
		info[3] = r.img.progressive;
		info[4] = r.nscans;
		info[5] = r.nsegs;
		info[6] = r.img.path;
		info[7] = r.img.color_mode;
		info[8] = r.img.n_out;
		info[9] = r.img.decode_n;
		info[10] = r.img.hmax;
		info[11] = r.img.vmax;
		info[12] = r.img.mcu_x;
		info[13] = r.img.mcu_y;
		info[14] = r.coef_blocks;
		info[15] = (long long)r.entropy_bytes;
	}
	else
		err_zero(r.reason);
	jp_result_free(&r);
	return ok;
}
