/*
 * jpeg_cuda.cu -- implementation of include/jpeg_cuda_abi.h: device context, uploads, kernel
 * launches (entropy waves, output stage), downloads.  sm_100a only; no CPU fallback: every
 * entry point fails with a CUDA error when no device is usable.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_entropy.cuh"
#include "jpeg_entropy_baseline.cuh"
#include "jpeg_selfsync.cuh"
#include "jpeg_entropy_prog.cuh"
#include "jpeg_encode.cuh"
#include "jpeg_widen.cuh"
#include "jpeg_output.cuh"
#include "jpeg_fused420.cuh"
#include "jpeg_fused1x1.cuh"
#include "jpeg_fused422.cuh"

struct DevBuf
{
	void *p = nullptr;
	size_t cap = 0;
	size_t guard_at = 0; /* guard mode: offset of the JCU_GUARD_BYTES canary behind the logical size */
};
#define JCU_GUARD_BYTES 4096
/* jcu_ctx_options.guard (test hook; compute-sanitizer is not available on every box): every device buffer of the
   context is allocated at exactly the size asked for plus a canary that jcu_debug_guards_ok() verifies */

/*
 * Downloads into PAGEABLE caller memory (malloc'ed buffers: what a user of stbi_load has).  A cudaMemcpyAsync to
 * pageable memory is staged by the driver and blocks the calling thread for its whole duration (measured: 17.7 GB/s
 * and no overlap with the next chunk's parse).  Instead the copy engine writes 2 MB pieces into a ring of pinned
 * slots and a few host threads move each piece on to the caller's buffer as its event completes; jcu_sync waits
 * until the pieces of ITS context have arrived.  ONE ring per device, shared by every context on it (the chunk
 * pipeline has three, the per-image pool up to sixteen), created on the first pageable download and kept for the
 * life of the process.
 */
#if defined(__x86_64__) || defined(_M_X64)
#include <emmintrin.h>
/* pinned slot -> caller's pageable buffer with streaming stores: the destination lines are written whole, so fetching
   them first (what a cached store does) only costs memory bandwidth the copier threads compete for */
static void bounce_copy(void *dst, const void *src, size_t n)
{
	uint8_t *d = (uint8_t *)dst;
	const uint8_t *s = (const uint8_t *)src;
	const size_t head = (16u - ((uintptr_t)d & 15u)) & 15u;
	if (n < 256 || head > n)
	{
		memcpy(d, s, n);
		return;
	}
	memcpy(d, s, head);
	d += head;
	s += head;
	n -= head;
	const size_t blocks = n / 64;
	for (size_t i = 0; i < blocks; ++i, d += 64, s += 64)
	{
		const __m128i a = _mm_loadu_si128((const __m128i *)s), b = _mm_loadu_si128((const __m128i *)(s + 16));
		const __m128i c = _mm_loadu_si128((const __m128i *)(s + 32)), e = _mm_loadu_si128((const __m128i *)(s + 48));
		_mm_stream_si128((__m128i *)d, a);
		_mm_stream_si128((__m128i *)(d + 16), b);
		_mm_stream_si128((__m128i *)(d + 32), c);
		_mm_stream_si128((__m128i *)(d + 48), e);
	}
	_mm_sfence();
	memcpy(d, s, n - blocks * 64);
}
#else
static void bounce_copy(void *dst, const void *src, size_t n) { memcpy(dst, src, n); }
#endif

struct BounceOwner /* per context; both fields are guarded by the ring's mutex */
{
	int pending = 0;
	cudaError_t failed = cudaSuccess;
};
struct Bounce
{
	enum
	{
		NSLOT = 32
	};
	/* copier threads: one moves ~6 GB/s; measured on a 16-core host, 1024 x 1080p: 3 / 6 / 10 / 16 threads ->
	   6.4 / 9.1 / 10.5 / 11.1 GMP/s (pinned buffers on the same box: 16.7) */
	static int copier_threads(int hint)
	{
		/* `hint` = host threads the first context that needs the ring was given: eight ranks on one box each get an
		   eighth of the cores for their parser pool, and the copiers must shrink with it (measured at 8 GPUs: ten copier
		   threads per rank on a 32-thread host lost 9 % against the driver's own staged copy) */
		const unsigned hc = hint > 0 ? (unsigned)hint : std::thread::hardware_concurrency();
		return (int)std::min(12u, std::max(2u, hc / 2u + 2u));
	}
	static constexpr size_t SLOT_BYTES = (size_t)2 << 20;
	struct Job
	{
		int slot;
		void *dst;
		size_t bytes;
		BounceOwner *owner;
	};
	int device = 0;
	uint8_t *buf[NSLOT] = {nullptr};
	cudaEvent_t ev[NSLOT] = {nullptr};
	std::mutex m;
	std::condition_variable cv_job, cv_free, cv_idle;
	std::deque<Job> jobs;
	std::vector<int> free_slots;
	std::vector<std::thread> threads;

	void worker()
	{
		cudaSetDevice(device);
		for (;;)
		{
			Job j;
			{
				std::unique_lock<std::mutex> lk(m);
				cv_job.wait(lk, [&] { return !jobs.empty(); });
				j = jobs.front();
				jobs.pop_front();
			}
			const cudaError_t e = j.bytes ? cudaEventSynchronize(ev[j.slot]) : cudaSuccess;
			if (e == cudaSuccess && j.bytes)
				bounce_copy(j.dst, buf[j.slot], j.bytes);
			{
				std::lock_guard<std::mutex> lk(m);
				if (e != cudaSuccess && j.owner->failed == cudaSuccess)
					j.owner->failed = e;
				free_slots.push_back(j.slot);
				--j.owner->pending;
			}
			cv_free.notify_one();
			cv_idle.notify_all();
		}
	}
	cudaError_t start(int dev, int threads_hint)
	{
		device = dev;
		for (int i = 0; i < NSLOT; ++i)
		{
			cudaError_t e = cudaHostAlloc((void **)&buf[i], SLOT_BYTES, cudaHostAllocDefault);
			if (e == cudaSuccess)
				e = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
			if (e != cudaSuccess)
				return e;
			free_slots.push_back(i);
		}
		const int nt = copier_threads(threads_hint);
		for (int t = 0; t < nt; ++t)
			threads.emplace_back([this] { worker(); }).detach(); /* they live as long as the process */
		return cudaSuccess;
	}
	int take_slot()
	{
		std::unique_lock<std::mutex> lk(m);
		cv_free.wait(lk, [&] { return !free_slots.empty(); });
		const int s = free_slots.back();
		free_slots.pop_back();
		return s;
	}
	void submit(int slot, void *dst, size_t bytes, BounceOwner *owner)
	{
		{
			std::lock_guard<std::mutex> lk(m);
			jobs.push_back(Job{slot, dst, bytes, owner});
			++owner->pending;
		}
		cv_job.notify_one();
	}
	cudaError_t drain(BounceOwner *owner)
	{
		std::unique_lock<std::mutex> lk(m);
		cv_idle.wait(lk, [&] { return owner->pending == 0; });
		const cudaError_t e = owner->failed;
		owner->failed = cudaSuccess;
		return e;
	}
};
#define JCU_MAX_DEVICES 64
static Bounce *g_bounce[JCU_MAX_DEVICES];
static std::mutex g_bounce_m;
/* the device's ring (the device is current); nullptr + *err when the pinned slots cannot be had */
static Bounce *bounce_for_device(int device, int threads_hint, cudaError_t *err)
{
	*err = cudaSuccess;
	if (device < 0 || device >= JCU_MAX_DEVICES)
	{
		*err = cudaErrorInvalidDevice;
		return nullptr;
	}
	std::lock_guard<std::mutex> lk(g_bounce_m);
	if (!g_bounce[device])
	{
		Bounce *b = new Bounce();
		*err = b->start(device, threads_hint);
		if (*err != cudaSuccess)
			return nullptr; /* a half-built ring is left behind rather than torn down under its threads: the error is sticky anyway */
		g_bounce[device] = b;
	}
	return g_bounce[device];
}

struct jcu_ctx
{
	int device = 0;
	Bounce *bounce = nullptr; /* the device's pinned ring + copier threads for pageable destinations (first use) */
	BounceOwner bounce_own;	  /* this context's pieces in flight through it */
	int profile = 0;
	int entropy_mode = 0;
	bool guard = false;
	int rows_per_band_forced = 0;
	uint32_t ss_sub_forced = 0;
	int host_threads_hint = 0;
	cudaStream_t stream = nullptr;
	/* downloads run on their own stream, behind an event of the compute stream */
	cudaStream_t dl_own = nullptr, dl_use = nullptr;
	cudaEvent_t ev_ready = nullptr, ev_dl = nullptr;
	bool dl_open = false;
	cudaEvent_t ev_h2d0 = nullptr, ev_h2d1 = nullptr, ev_run0 = nullptr, ev_ent1 = nullptr, ev_out1 = nullptr, ev_d2h0 = nullptr,
					ev_d2h1 = nullptr;
	bool d2h_open = false;
	DevBuf bits, images, scans, segs, huff, work, coef, planes, out, err, lists;
	DevBuf ss_dwork, ss_dscans; /* the decoding phases' work items / scans (subsequences of ss_sub_bytes) */
	DevBuf ss_work, ss_scans, ss_exit_a, ss_exit_b, ss_entry, ss_nblk, ss_first, ss_changed;
	DevBuf wide, wide_lut; /* 16-bit / float copy of one image and its look-up table (jpeg_widen.cuh) */
	DevBuf uslab, useg_len; /* unstuffed copy of the baseline segments (jpeg_entropy_baseline.cuh) */
	DevBuf enc_pix, enc_img, enc_coef, enc_nbits, enc_bitoff, enc_tot, enc_bitbuf, enc_ffcnt, enc_ffoff, enc_out, enc_len, enc_hdr; /* encoder */
	cudaEvent_t ev_enc0 = nullptr, ev_enc1 = nullptr;
	std::vector<SsWork> ss_work_h; /* self-synchronising decode: CTA work items over all waves */
	std::vector<SsScan> ss_scan_h;
	std::vector<SsWork> ss_dwork_h; /* the same scans cut into decoding subsequences */
	std::vector<SsScan> ss_dscan_h;
	uint32_t ss_nsub_total = 0, ss_max_nsub = 0; /* raw 256-byte chunks (unstuffing) */
	uint32_t ss_dsub_total = 0, ss_max_dsub = 0, ss_sub_bytes = SS_SUB_BYTES; /* decoding subsequences */
	int ss_rounds = 0; /* refine rounds the last run needed (read back with the per-image errors) */
	int ss_coop_ctas = 0, ss_coop_ctas2 = 0; /* CTAs of k_ss_refine_loop<4 + 4 tables / 2 + 2 tables> that are resident at once on this device */
	bool ss_two = true;	 /* every self-sync scan of the batch uses <= 2 distinct DC and <= 2 distinct AC tables */
	int pc_resident = 1024; /* CTAs of k_prog_coop that are resident at once */
	/* batch metadata kept on the host */
	int n_images = 0, n_waves = 0, n_scans = 0;
	size_t coef_elems = 0, out_bytes = 0;
	bool need_clear = false;
	std::vector<EntWork> work_h;
	std::vector<int> wave_off; /* n_waves + 1 offsets into work_h */
	std::vector<int> wave_nbase; /* per wave: leading work items that are baseline scans */
	std::vector<char> wave_two;	  /* per wave: every baseline scan uses <= 2 distinct DC and <= 2 distinct AC tables */
	std::vector<int> wave_ncoop;	  /* per wave: trailing work items that are single-segment progressive AC scans (jpeg_entropy_prog.cuh) */
	std::vector<int> wave_nsmall;	  /* per wave: work items after the baseline ones that are progressive scans of <= 32 segments */
	std::vector<EntWork> pipe_work_h; /* progressive scans without restart markers of images whose scans ALL take the warp decoder:
													  one pipelined launch for the whole dependency chain (jpeg_entropy_prog.cuh, PcPipe) */
	std::vector<int32_t> pipe_preds_h; /* per scan: PC_MAX_PREDS predecessor scans, -1 = none */
	DevBuf pipe_work, pipe_preds, pipe_progress;
	std::vector<int32_t> list_h; /* image lists per path, concatenated */
	int list_off[JCU_PATH_COUNT] = {0, 0, 0, 0, 0}, list_n[JCU_PATH_COUNT] = {0, 0, 0, 0, 0};
	int tiles[JCU_PATH_COUNT] = {0, 0, 0, 0, 0};
	int path_n[JCU_PATH_COUNT] = {3, 3, 3, 3, 3};
	int path_deq[JCU_PATH_COUNT] = {0, 0, 0, 0, 0};
	int path_maxtiles[JCU_PATH_COUNT] = {0, 0, 0, 0, 0};
	int max_blocks_generic = 0;
	long long max_pixels_generic = 0;
	int launches = 0;
	int force_generic = 0;
	int rows_per_band = 8; /* 4:2:0 strip kernel: MCU rows per CTA, chosen per batch */
	/* pinned staging of the descriptor arrays: a cudaMemcpyAsync from pageable memory blocks the host until the stream
	   has reached it (behind the bitstream upload: 4-8 ms per 512-image chunk), which starves the chunk pipeline */
	unsigned char *pin_desc = nullptr;
	size_t pin_desc_cap = 0;
	/* per-image failure words + the self-sync round count, published by a last tiny kernel into MAPPED pinned memory:
	   a device-to-host copy of a few bytes would queue behind the bulk pixel downloads of the other pipeline
	   objects on the copy engine (measured: 7-12 ms per chunk) */
	int32_t *err_host = nullptr, *err_host_dev = nullptr;
	size_t err_host_cap = 0; /* int32 words incl. the 2 trailing self-sync words */
	char errstr[256] = {0};
};

/* is this host address page-locked (cudaHostAlloc / cudaHostRegister) or otherwise known to the driver? */
static bool host_ptr_is_pinned(const void *p)
{
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess)
	{
		cudaGetLastError();
		return false;
	}
	return a.type != cudaMemoryTypeUnregistered;
}

static int set_err(jcu_ctx *c, cudaError_t e, const char *where)
{
	if (c)
		snprintf(c->errstr, sizeof(c->errstr), "%s: %s", where, cudaGetErrorString(e));
	return (int)e;
}
#define CK(call)                                   \
	do                                              \
	{                                               \
		cudaError_t e__ = (call);                    \
		if (e__ != cudaSuccess)                      \
			return set_err(c, e__, #call);            \
	} while (0)

#define ensure(buf, bytes) ensure_impl(c->guard, buf, bytes) /* every caller has its context in `c` */
static cudaError_t ensure_impl(bool guard_mode, DevBuf &b, size_t bytes)
{
	if (guard_mode)
	{
		if (b.p)
			cudaFree(b.p);
		b.p = nullptr;
		b.cap = 0;
		const size_t at = (bytes + 15) & ~(size_t)15;
		cudaError_t e = cudaMalloc(&b.p, at + JCU_GUARD_BYTES);
		if (e != cudaSuccess)
			return e;
		b.cap = bytes;
		b.guard_at = at;
		return cudaMemset((char *)b.p + at, 0xA5, JCU_GUARD_BYTES);
	}
	if (bytes <= b.cap)
		return cudaSuccess;
	if (b.p)
		cudaFree(b.p);
	b.p = nullptr;
	b.cap = 0;
	size_t want = bytes + bytes / 8 + 256;
	cudaError_t e = cudaMalloc(&b.p, want);
	if (e == cudaSuccess)
		b.cap = want;
	return e;
}

extern "C" int jcu_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess)
	{
		cudaGetLastError();
		return 0;
	}
	return n;
}

extern "C" int jcu_ctx_create(int device, int profile, int entropy_mode, jcu_ctx **out)
{
	jcu_ctx_options o;
	memset(&o, 0, sizeof(o));
	o.device = device;
	o.profile = profile;
	o.entropy_mode = entropy_mode;
	return jcu_ctx_create_opts(&o, out);
}

extern "C" int jcu_ctx_create_opts(const jcu_ctx_options *o, jcu_ctx **out)
{
	*out = nullptr;
	cudaError_t e = cudaSetDevice(o->device);
	if (e != cudaSuccess)
		return (int)e;
	jcu_ctx *c = new jcu_ctx();
	c->device = o->device;
	c->profile = o->profile;
	c->entropy_mode = o->entropy_mode;
	c->guard = o->guard != 0;
	c->force_generic = o->force_generic != 0;
	c->rows_per_band_forced = o->rows_per_band > 0 ? o->rows_per_band : 0;
	c->host_threads_hint = o->host_threads > 0 ? o->host_threads : 0;
	if (o->selfsync_sub_bytes == 32 || o->selfsync_sub_bytes == 64 || o->selfsync_sub_bytes == 128 || o->selfsync_sub_bytes == 256)
		c->ss_sub_forced = (uint32_t)o->selfsync_sub_bytes;
	e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
	if (e != cudaSuccess)
	{
		delete c;
		return (int)e;
	}
	e = cudaStreamCreateWithFlags(&c->dl_own, cudaStreamNonBlocking);
	if (e != cudaSuccess)
	{
		cudaStreamDestroy(c->stream);
		delete c;
		return (int)e;
	}
	c->dl_use = c->dl_own;
	cudaEventCreateWithFlags(&c->ev_ready, cudaEventDisableTiming);
	cudaEventCreateWithFlags(&c->ev_dl, cudaEventDisableTiming);
	cudaEvent_t *evs[] = {&c->ev_h2d0, &c->ev_h2d1, &c->ev_run0, &c->ev_ent1, &c->ev_out1, &c->ev_d2h0, &c->ev_d2h1};
	for (auto ev : evs)
		cudaEventCreate(ev);
	/* tables (static) + slots and rings (dynamic) exceed the 48 KB default */
	cudaFuncSetAttribute(k_entropy_baseline<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, ENTB_DYN_SMEM);
	cudaFuncSetAttribute(k_entropy_baseline<4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, ENTB_DYN_SMEM);
	{
		int per_sm = 0, sms = 0;
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ss_refine_loop<EntbShared>, SS_THREADS, 0);
		c->ss_coop_ctas = std::max(1, per_sm * sms);
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ss_refine_loop<EntbShared2>, SS_THREADS, 0);
		c->ss_coop_ctas2 = std::max(1, per_sm * sms);
		per_sm = 0;
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_prog_coop, PC_THREADS, 0);
		c->pc_resident = std::max(64, per_sm * sms);
	}
	*out = c;
	return 0;
}

extern "C" void jcu_ctx_destroy(jcu_ctx *c)
{
	if (!c)
		return;
	cudaSetDevice(c->device);
	cudaStreamSynchronize(c->stream);
	DevBuf *bufs[] = {&c->bits, &c->images, &c->scans, &c->segs, &c->huff, &c->work, &c->coef, &c->planes, &c->out, &c->err, &c->lists,
							&c->ss_work, &c->ss_scans, &c->ss_dwork, &c->ss_dscans, &c->ss_exit_a, &c->ss_exit_b, &c->ss_entry, &c->ss_nblk, &c->ss_first, &c->ss_changed, &c->uslab, &c->useg_len,
							&c->pipe_work, &c->pipe_preds, &c->pipe_progress, &c->enc_pix, &c->enc_img, &c->enc_coef, &c->enc_nbits, &c->enc_bitoff, &c->enc_tot, &c->enc_bitbuf, &c->enc_ffcnt, &c->enc_ffoff, &c->enc_out, &c->enc_len, &c->enc_hdr, &c->wide, &c->wide_lut};
	for (auto b : bufs)
		if (b->p)
			cudaFree(b->p);
	if (c->ev_enc0)
		cudaEventDestroy(c->ev_enc0);
	if (c->ev_enc1)
		cudaEventDestroy(c->ev_enc1);
	cudaEvent_t evs[] = {c->ev_h2d0, c->ev_h2d1, c->ev_run0, c->ev_ent1, c->ev_out1, c->ev_d2h0, c->ev_d2h1};
	for (auto ev : evs)
		if (ev)
			cudaEventDestroy(ev);
	cudaStreamSynchronize(c->dl_use);
	if (c->bounce)
		c->bounce->drain(&c->bounce_own); /* no copier thread may still hold a pointer into this context */
	if (c->pin_desc)
		cudaFreeHost(c->pin_desc);
	if (c->err_host)
		cudaFreeHost(c->err_host);
	cudaEventDestroy(c->ev_ready);
	cudaEventDestroy(c->ev_dl);
	cudaStreamDestroy(c->dl_own);
	cudaStreamDestroy(c->stream);
	delete c;
}


extern "C" int jcu_debug_guards_ok(jcu_ctx *c)
{
	if (!c || !c->guard)
		return -1;
	cudaSetDevice(c->device);
	if (cudaStreamSynchronize(c->stream) != cudaSuccess)
		return 0;
	DevBuf *bufs[] = {&c->bits, &c->images, &c->scans, &c->segs, &c->huff, &c->work, &c->coef, &c->planes, &c->out, &c->err, &c->lists,
							&c->ss_work, &c->ss_scans, &c->ss_dwork, &c->ss_dscans, &c->ss_exit_a, &c->ss_exit_b, &c->ss_entry, &c->ss_nblk, &c->ss_first, &c->ss_changed,
							&c->uslab, &c->useg_len};
	std::vector<unsigned char> h(JCU_GUARD_BYTES);
	int which = 0;
	for (auto b : bufs)
	{
		++which;
		if (!b->p)
			continue;
		if (cudaMemcpy(h.data(), (char *)b->p + b->guard_at, JCU_GUARD_BYTES, cudaMemcpyDeviceToHost) != cudaSuccess)
			return 0;
		for (size_t i = 0; i < JCU_GUARD_BYTES; ++i)
			if (h[i] != 0xA5)
			{
				snprintf(c->errstr, sizeof(c->errstr), "guard of device buffer #%d overwritten at +%zu", which, i);
				return 0;
			}
	}
	return 1;
}

extern "C" int jcu_mem_info(int device, size_t *free_bytes, size_t *total_bytes)
{
	size_t f = 0, t = 0;
	cudaError_t e = cudaSetDevice(device);
	if (e == cudaSuccess)
		e = cudaMemGetInfo(&f, &t);
	if (e != cudaSuccess)
		cudaGetLastError();
	if (free_bytes)
		*free_bytes = f;
	if (total_bytes)
		*total_bytes = t;
	return (int)e;
}

extern "C" const char *jcu_error_string(jcu_ctx *c) { return c ? c->errstr : "no context"; }

extern "C" void *jcu_host_alloc(size_t bytes)
{
	void *p = nullptr;
	if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) /* page-locked for every device: the multi-device call shares one pair of buffers */
	{
		cudaGetLastError();
		return nullptr;
	}
	return p;
}
extern "C" void jcu_host_free(void *p)
{
	if (p)
		cudaFreeHost(p);
}

extern "C" int jcu_upload(jcu_ctx *c, const jcu_batch *b)
{
	CK(cudaSetDevice(c->device));
	c->n_images = b->n_images;
	c->n_waves = b->n_waves;
	c->n_scans = b->n_scans;
	c->coef_elems = b->coef_elems;
	c->out_bytes = b->out_bytes;
	c->need_clear = b->need_coef_clear != 0;
	c->d2h_open = false;

	/* ---- per-path image lists and tile numbering ---- */
	std::vector<jcu_image> imgs(b->images, b->images + b->n_images);
	std::vector<int32_t> lists[JCU_PATH_COUNT];
	c->max_blocks_generic = 0;
	c->max_pixels_generic = 0;
	for (int p = 0; p < JCU_PATH_COUNT; ++p)
		c->tiles[p] = c->path_deq[p] = c->path_maxtiles[p] = 0;
	{
		/* band height of the 4:2:0 strip kernel: tall bands amortise the two-row chroma prologue, but the
		   grid must still fill 148 SMs x 2 CTAs several times over */
		long long strip_rows = 0;
		for (int i = 0; i < b->n_images; ++i)
			if (imgs[i].ok && !c->force_generic && imgs[i].path == JCU_PATH_FUSED_420)
				strip_rows += (long long)((imgs[i].mcu_x + S420_MS - 1) / S420_MS) * imgs[i].mcu_y;
		long long rb = strip_rows / (148 * (512 / S420_THREADS) * 8);
		c->rows_per_band = (int)std::min<long long>(std::max<long long>(rb, 2), 24);
		if (c->rows_per_band_forced) /* jcu_ctx_options.rows_per_band: tests force the band height */
			c->rows_per_band = c->rows_per_band_forced;
	}
	for (int i = 0; i < b->n_images; ++i)
	{
		jcu_image &im = imgs[i];
		if (!im.ok)
			continue;
		if (c->force_generic)
			im.path = JCU_PATH_GENERIC;
		int p = im.path;
		lists[p].push_back(i);
		c->path_n[p] = im.n_out;
		c->path_deq[p] |= im.progressive;
		if (p == JCU_PATH_GENERIC)
		{
			for (int k = 0; k < im.decode_n; ++k)
				c->max_blocks_generic = std::max(c->max_blocks_generic, im.comp[k].bw * im.comp[k].bh);
			c->max_pixels_generic = std::max(c->max_pixels_generic, (long long)im.width * im.height);
		}
		else
		{
			if (p == JCU_PATH_FUSED_420)
			{
				im.tiles_x = (im.mcu_x + S420_MS - 1) / S420_MS;						/* strips */
				im.tiles_y = (im.mcu_y + c->rows_per_band - 1) / c->rows_per_band; /* bands */
			}
			else if (p == JCU_PATH_FUSED_422)
			{
				im.tiles_x = (im.mcu_x + F2_MS - 1) / F2_MS;
				im.tiles_y = (im.mcu_y + F2_ROWS_PER_BAND - 1) / F2_ROWS_PER_BAND;
			}
			else
			{
				const int bs = p == JCU_PATH_FUSED_444 ? F1_BS(3) : F1_BS(1);
				im.tiles_x = (im.comp[0].bw + bs - 1) / bs;											 /* strips */
				im.tiles_y = ((im.height + 7) / 8 + F1_ROWS_PER_BAND - 1) / F1_ROWS_PER_BAND; /* bands */
			}
			im.tile_first = c->tiles[p];
			c->tiles[p] += im.tiles_x * im.tiles_y;
			c->path_maxtiles[p] = std::max(c->path_maxtiles[p], im.tiles_x * im.tiles_y);
		}
	}
	c->list_h.clear();
	for (int p = 0; p < JCU_PATH_COUNT; ++p)
	{
		c->list_off[p] = (int)c->list_h.size();
		c->list_n[p] = (int)lists[p].size();
		c->list_h.insert(c->list_h.end(), lists[p].begin(), lists[p].end());
	}

	/* ---- CTA work items per wave: ENT_THREADS segments of one scan each ---- */
	c->work_h.clear();
	c->ss_work_h.clear();
	c->ss_scan_h.clear();
	c->ss_dwork_h.clear();
	c->ss_dscan_h.clear();
	c->ss_nsub_total = 0;
	c->ss_max_nsub = 0;
	c->ss_dsub_total = 0;
	c->ss_max_dsub = 0;
	std::vector<char> is_ss((size_t)std::max(1, b->n_scans), 0);
	for (int s = 0; s < b->n_scans; ++s)
	{
		/* baseline scans without restart markers: one serial segment per scan, unless it is long enough
		   (or entropy_mode == 2) for the self-synchronising decoder; entropy_mode == 1 forces the serial path */
		const jcu_scan &sc = b->scans[s];
		if (!imgs[sc.image].ok || sc.kind != JCU_SCAN_BASELINE || sc.restart_interval != 0 || sc.nseg != 1 || c->entropy_mode == 1)
			continue;
		{
			/* The self-synchronising scans all run behind the last wave, whatever their own wave: a scan that shares a
			   component with another scan of its image (legal in T.81: the later scan wins, jpeg.c:1716) must keep its
			   place in the wave order, so it stays on the serial path.  Scans of one image are contiguous. */
			bool shared = false;
			for (int dir = -1; dir <= 1 && !shared; dir += 2)
				for (int o = s + dir; o >= 0 && o < b->n_scans && b->scans[o].image == sc.image && !shared; o += dir)
					for (int a = 0; a < sc.ncomp; ++a)
						for (int q = 0; q < b->scans[o].ncomp; ++q)
							shared = shared || sc.comp[a] == b->scans[o].comp[q];
			if (shared)
				continue;
		}
		const uint32_t start = b->seg_start[sc.seg_first];
		const uint32_t bytes = sc.data_end > start ? sc.data_end - start : 0u;
		if (bytes < (c->entropy_mode == 2 ? 1u : 4u * SS_SUB_BYTES))
			continue;
		SsScan e;
		e.scan = s;
		e.base = c->ss_nsub_total;
		/* the unstuffing threads cut the scan on 256-byte boundaries of the FILE (aligned 16-byte loads), so a scan that
		   starts mid-chunk needs one chunk more; the decoding subsequences (unstuffed bytes) never outnumber them */
		e.nsub = ((start & (SS_SUB_BYTES - 1u)) + bytes + SS_SUB_BYTES - 1) / SS_SUB_BYTES;
		e.pad = 0;
		c->ss_scan_h.push_back(e);
		for (uint32_t f = 0; f < e.nsub; f += SS_THREADS)
		{
			SsWork wk;
			wk.scan = s;
			wk.first = f;
			wk.count = std::min<uint32_t>(SS_THREADS, e.nsub - f);
			wk.scan_base = e.base;
			wk.scan_nsub = e.nsub;
			wk.pad = 0;
			c->ss_work_h.push_back(wk);
		}
		c->ss_nsub_total += e.nsub;
		c->ss_max_nsub = std::max(c->ss_max_nsub, e.nsub);
		is_ss[s] = 1;
	}
	if (!c->ss_scan_h.empty())
	{
		/* phase 4 writes single coefficients into the blocks that hold a subsequence boundary: k_ss_zero_shared clears
		   exactly those (the whole slab is only cleared when the parser asks for it) */
		/* Decoding subsequences: 256 unstuffed bytes when the batch fills the GPU with them; a small batch (the
		   per-image stbi_load of ONE 512 x 512 file is 409 of them) takes shorter ones, down to 32 bytes -- every phase is
		   one dependent chain per thread, so its latency is the subsequence length, and idle SMs cost nothing. */
		unsigned long long total = 0;
		c->ss_two = true;
		for (const SsScan &e : c->ss_scan_h)
		{
			const jcu_scan &sc = b->scans[e.scan];
			total += sc.data_end - b->seg_start[sc.seg_first];
			int nd = 0, na = 0;
			for (int i = 0; i < sc.ncomp; ++i)
			{
				bool newd = true, newa = true;
				for (int j = 0; j < i; ++j)
				{
					newd = newd && sc.dc_spec[j] != sc.dc_spec[i];
					newa = newa && sc.ac_spec[j] != sc.ac_spec[i];
				}
				nd += newd;
				na += newa;
			}
			if (nd > 2 || na > 2)
				c->ss_two = false; /* the kernels then hold four tables of each kind: fewer CTAs per SM */
		}
		uint32_t sub = SS_SUB_BYTES;
		while (sub > SS_MIN_SUB && total / sub < 131072ull)
			sub >>= 1;
		if (c->ss_sub_forced) /* jcu_ctx_options.selfsync_sub_bytes: tests and tuning */
			sub = c->ss_sub_forced;
		c->ss_sub_bytes = sub;
		for (const SsScan &u : c->ss_scan_h)
		{
			const jcu_scan &sc = b->scans[u.scan];
			const uint32_t bytes = sc.data_end - b->seg_start[sc.seg_first];
			SsScan e = u;
			e.base = c->ss_dsub_total;
			e.nsub = std::max(1u, (bytes + sub - 1u) / sub); /* the unstuffed stream is never longer than the raw one */
			c->ss_dscan_h.push_back(e);
			for (uint32_t f = 0; f < e.nsub; f += SS_THREADS)
			{
				SsWork wk;
				wk.scan = e.scan;
				wk.first = f;
				wk.count = std::min<uint32_t>(SS_THREADS, e.nsub - f);
				wk.scan_base = e.base;
				wk.scan_nsub = e.nsub;
				wk.pad = 0;
				c->ss_dwork_h.push_back(wk);
			}
			c->ss_dsub_total += e.nsub;
			c->ss_max_dsub = std::max(c->ss_max_dsub, e.nsub);
		}
	}
	/* which scans take the warp decoder (k_prog_coop), and which images consist of such scans only */
	std::vector<char> is_coop((size_t)std::max(1, b->n_scans), 0), img_piped((size_t)std::max(1, b->n_images), 0);
	{
		std::vector<int> n_all((size_t)std::max(1, b->n_images), 0), n_coop((size_t)std::max(1, b->n_images), 0);
		for (int s = 0; s < b->n_scans; ++s)
		{
			const jcu_scan &sc = b->scans[s];
			if (!imgs[sc.image].ok)
				continue;
			const bool ac = sc.kind == JCU_SCAN_PROG_AC_FIRST || sc.kind == JCU_SCAN_PROG_AC_REFINE;
			/* a first DC pass clears its blocks (jpeg.c:385): only one that nothing precedes may rely on the zeroed slab */
			const bool dc = sc.kind == JCU_SCAN_PROG_DC_REFINE || (sc.kind == JCU_SCAN_PROG_DC_FIRST && sc.wave == 0 && c->need_clear);
			int distinct_dc = 0; /* a first DC pass may use one table per component; the warp decoder holds PC_MAX_LUTS */
			if (sc.kind == JCU_SCAN_PROG_DC_FIRST)
				for (int i = 0; i < sc.ncomp; ++i)
				{
					bool fresh = true;
					for (int j = 0; j < i; ++j)
						fresh = fresh && sc.dc_spec[j] != sc.dc_spec[i];
					distinct_dc += fresh;
				}
			is_coop[s] = ((ac && sc.ncomp == 1) || dc) && sc.nseg == 1 && c->entropy_mode != 1 && !is_ss[s] && distinct_dc <= PC_MAX_LUTS;
			++n_all[sc.image];
			n_coop[sc.image] += is_coop[s];
		}
		for (int i = 0; i < b->n_images; ++i)
			img_piped[i] = n_all[i] > 1 && n_all[i] == n_coop[i];
	}
	c->pipe_work_h.clear();
	c->pipe_preds_h.assign((size_t)std::max(1, b->n_scans) * PC_MAX_PREDS, -1);
	std::vector<uint64_t> behind((size_t)std::max(1, b->n_scans), 0); /* scans (bit = ordinal in the image) a scan is known to stay behind */
	for (int s = 0; s < b->n_scans; ++s)
	{
		const jcu_scan &sc = b->scans[s];
		if (!imgs[sc.image].ok || !img_piped[sc.image])
			continue;
		int first = s; /* first scan of the image */
		while (first > 0 && b->scans[first - 1].image == sc.image)
			--first;
		if (s - first >= 64)
		{
			img_piped[sc.image] = 2;
			continue;
		}
		/* the predecessors of a scan: earlier scans of its image that share a component and overlap its write set -- the
		   rule that gives the dependency levels in jpeg_parse.c (one_scan).  The NEAREST ones are kept: they are behind
		   their own predecessors in turn. */
		auto lo = [](const jcu_scan &x) { return (x.kind == JCU_SCAN_BASELINE || x.kind == JCU_SCAN_PROG_DC_FIRST) ? 0 : x.ss; };
		auto hi = [](const jcu_scan &x) {
			return (x.kind == JCU_SCAN_BASELINE || x.kind == JCU_SCAN_PROG_DC_FIRST) ? 63
					 : x.kind == JCU_SCAN_PROG_AC_FIRST											  ? std::min(63, x.se + 15)
																													  : x.se;
		};
		int np = 0;
		for (int j = s - 1; j >= 0 && b->scans[j].image == sc.image; --j)
		{
			const jcu_scan &o = b->scans[j];
			bool shared = false;
			for (int a = 0; a < sc.ncomp; ++a)
				for (int q = 0; q < o.ncomp; ++q)
					shared = shared || sc.comp[a] == o.comp[q];
			if (!shared || lo(sc) > hi(o) || lo(o) > hi(sc))
				continue;
			/* a predecessor of a predecessor needs no wait of its own: the nearer one never overtakes it */
			if (behind[s] & (1ull << (j - first)))
				continue;
			behind[s] |= (1ull << (j - first)) | behind[j];
			if (np == PC_MAX_PREDS)
			{
				img_piped[sc.image] = 2; /* more predecessors than the table holds: this image keeps the level-by-level launches */
				break;
			}
			c->pipe_preds_h[(size_t)s * PC_MAX_PREDS + np++] = j;
		}
	}
	c->wave_off.assign((size_t)b->n_waves + 1, 0);
	c->wave_nbase.assign((size_t)b->n_waves, 0);
	c->wave_two.assign((size_t)b->n_waves, 1);
	c->wave_nsmall.assign((size_t)b->n_waves, 0);
	c->wave_ncoop.assign((size_t)b->n_waves, 0);
	for (int w = 0; w < b->n_waves; ++w)
	{
		c->wave_off[w] = (int)c->work_h.size();
		/* four groups per wave: baseline scans (k_entropy_baseline), progressive scans of at most
		   ENT_SMALL_THREADS segments (32-thread CTAs), the other progressive scans, and progressive scans
		   that are one long segment (warp-cooperative decoder, jpeg_entropy_prog.cuh; entropy_mode == 1 keeps
		   them on the one-thread-per-segment path) */
		for (int pass = 0; pass < 4; ++pass)
		{
			for (int s = 0; s < b->n_scans; ++s)
			{
				const jcu_scan &sc = b->scans[s];
				if (sc.wave != w || !imgs[sc.image].ok || is_ss[s])
					continue;
				const bool coop = is_coop[s] != 0;
				const int group = sc.kind == JCU_SCAN_BASELINE ? 0 : coop ? 3 : sc.nseg <= ENT_SMALL_THREADS ? 1 : 2;
				if (group != pass)
					continue;
				if (coop && img_piped[sc.image] == 1)
				{
					/* the whole chain of this image goes into the one pipelined launch, in level order */
					EntWork wk;
					wk.scan = s;
					wk.seg_first = 0;
					wk.nseg = 1;
					c->pipe_work_h.push_back(wk);
					continue;
				}
				if (pass == 0)
				{
					int nd = 0, na = 0;
					for (int i = 0; i < sc.ncomp; ++i)
					{
						bool newd = true, newa = true;
						for (int j = 0; j < i; ++j)
						{
							newd = newd && sc.dc_spec[j] != sc.dc_spec[i];
							newa = newa && sc.ac_spec[j] != sc.ac_spec[i];
						}
						nd += newd;
						na += newa;
					}
					if (nd > 2 || na > 2)
						c->wave_two[w] = 0;
				}
				const uint32_t per_cta = pass == 0 ? ENTB_THREADS : pass == 1 ? ENT_SMALL_THREADS : ENT_THREADS;
				for (uint32_t f = 0; f < sc.nseg; f += per_cta)
				{
					EntWork wk;
					wk.scan = s;
					wk.seg_first = f;
					wk.nseg = std::min<uint32_t>(per_cta, sc.nseg - f);
					c->work_h.push_back(wk);
				}
			}
			if (pass == 0)
				c->wave_nbase[w] = (int)c->work_h.size() - c->wave_off[w];
			if (pass == 1)
				c->wave_nsmall[w] = (int)c->work_h.size() - c->wave_off[w] - c->wave_nbase[w];
			if (pass == 2)
				c->wave_ncoop[w] = (int)c->work_h.size(); /* start of the last group; a count after the loop */
		}
		c->wave_ncoop[w] = (int)c->work_h.size() - c->wave_ncoop[w];
	}
	c->wave_off[b->n_waves] = (int)c->work_h.size();
	if (!c->pipe_work_h.empty())
	{
		/* Ticket order of the pipelined launch.  Level-major order over the whole batch would fill the machine with the
		   first levels of all images; instead the images go in groups whose scans are all resident at once, level-major
		   inside a group -- predecessors still come first, and every chain of a group runs as a pipeline. */
		int max_scans = 1;
		std::vector<int> per_img((size_t)b->n_images, 0), rank((size_t)b->n_images, -1);
		for (const EntWork &wk : c->pipe_work_h)
			max_scans = std::max(max_scans, ++per_img[b->scans[wk.scan].image]);
		int nrank = 0;
		for (int i = 0; i < b->n_images; ++i)
			if (per_img[i])
				rank[i] = nrank++;
		const int group = std::max(1, (c->pc_resident * 9 / 10) / max_scans);
		std::stable_sort(c->pipe_work_h.begin(), c->pipe_work_h.end(), [&](const EntWork &x, const EntWork &y) {
			const jcu_scan &sx = b->scans[x.scan], &sy = b->scans[y.scan];
			const int gx = rank[sx.image] / group, gy = rank[sy.image] / group;
			if (gx != gy)
				return gx < gy;
			if (sx.wave != sy.wave)
				return sx.wave < sy.wave;
			return x.scan < y.scan;
		});
	}

	/* ---- device buffers ---- */
	CK(ensure(c->bits, b->bitstream_bytes + 64));
	{
		bool any_base = false;
		for (int w = 0; w < b->n_waves; ++w)
			any_base = any_base || c->wave_nbase[w] > 0;
		if (any_base || !c->ss_scan_h.empty()) /* the self-synchronising path unstuffs its scans into the same slab */
		{
			CK(ensure(c->uslab, b->bitstream_bytes + 32 * ((size_t)b->n_segs + 4)));
			CK(ensure(c->useg_len, sizeof(uint32_t) * std::max<size_t>(1, b->n_segs)));
		}
	}
	CK(ensure(c->images, sizeof(jcu_image) * std::max(1, b->n_images)));
	CK(ensure(c->scans, sizeof(jcu_scan) * std::max(1, b->n_scans)));
	CK(ensure(c->segs, sizeof(uint32_t) * std::max<size_t>(1, b->n_segs)));
	CK(ensure(c->huff, sizeof(jcu_huff_spec) * std::max(1, b->n_huff)));
	CK(ensure(c->work, sizeof(EntWork) * std::max<size_t>(1, c->work_h.size())));
	CK(ensure(c->coef, sizeof(int16_t) * b->coef_elems + 256));
	CK(ensure(c->planes, b->plane_bytes + 256));
	CK(ensure(c->out, b->out_bytes + 256));
	CK(ensure(c->err, sizeof(int32_t) * std::max(1, b->n_images)));
	CK(ensure(c->lists, sizeof(int32_t) * std::max<size_t>(1, c->list_h.size())));
	if (!c->ss_scan_h.empty())
	{
		const size_t n = std::max(c->ss_nsub_total, c->ss_dsub_total);
		CK(ensure(c->ss_work, sizeof(SsWork) * c->ss_work_h.size()));
		CK(ensure(c->ss_scans, sizeof(SsScan) * c->ss_scan_h.size()));
		CK(ensure(c->ss_dwork, sizeof(SsWork) * c->ss_dwork_h.size()));
		CK(ensure(c->ss_dscans, sizeof(SsScan) * c->ss_dscan_h.size()));
		CK(ensure(c->ss_exit_a, sizeof(SsState) * n));
		CK(ensure(c->ss_exit_b, sizeof(SsState) * n));
		CK(ensure(c->ss_entry, sizeof(SsState) * n));
		CK(ensure(c->ss_nblk, sizeof(uint32_t) * n));
		CK(ensure(c->ss_first, sizeof(uint32_t) * n));
		CK(ensure(c->ss_changed, 256));
	}

	if (!c->pipe_work_h.empty())
	{
		CK(ensure(c->pipe_work, sizeof(EntWork) * c->pipe_work_h.size()));
		CK(ensure(c->pipe_preds, sizeof(int32_t) * c->pipe_preds_h.size()));
		CK(ensure(c->pipe_progress, sizeof(int32_t) * ((size_t)b->n_scans + 4)));
	}
	if ((size_t)b->n_images + 2 > c->err_host_cap)
	{
		if (c->err_host)
			cudaFreeHost(c->err_host);
		c->err_host = nullptr;
		c->err_host_cap = 0;
		const size_t want = (size_t)b->n_images + 2 + 64;
		CK(cudaHostAlloc((void **)&c->err_host, want * sizeof(int32_t), cudaHostAllocMapped));
		CK(cudaHostGetDevicePointer((void **)&c->err_host_dev, c->err_host, 0));
		c->err_host_cap = want;
	}
	/* ---- descriptors go through one pinned staging block, so that no copy below blocks the host ---- */
	struct Part
	{
		const void *src;
		size_t bytes;
		void *dst;
	};
	const Part parts[] = {
		{imgs.data(), sizeof(jcu_image) * (size_t)b->n_images, c->images.p},
		{b->scans, sizeof(jcu_scan) * (size_t)b->n_scans, c->scans.p},
		{b->seg_start, sizeof(uint32_t) * (size_t)b->n_segs, c->segs.p},
		{b->huff, sizeof(jcu_huff_spec) * (size_t)b->n_huff, c->huff.p},
		{c->work_h.data(), sizeof(EntWork) * c->work_h.size(), c->work.p},
		{c->ss_work_h.data(), sizeof(SsWork) * c->ss_work_h.size(), c->ss_work.p},
		{c->ss_scan_h.data(), sizeof(SsScan) * c->ss_scan_h.size(), c->ss_scans.p},
		{c->ss_dwork_h.data(), sizeof(SsWork) * c->ss_dwork_h.size(), c->ss_dwork.p},
		{c->ss_dscan_h.data(), sizeof(SsScan) * c->ss_dscan_h.size(), c->ss_dscans.p},
		{c->list_h.data(), sizeof(int32_t) * c->list_h.size(), c->lists.p},
		{c->pipe_work_h.data(), sizeof(EntWork) * c->pipe_work_h.size(), c->pipe_work.p},
		{c->pipe_preds_h.data(), c->pipe_work_h.empty() ? 0 : sizeof(int32_t) * c->pipe_preds_h.size(), c->pipe_preds.p},
	};
	size_t desc_bytes = 0;
	for (const Part &pt : parts)
		desc_bytes += (pt.bytes + 255) & ~(size_t)255;
	if (desc_bytes > c->pin_desc_cap)
	{
		/* the previous batch of this context has been waited for (jcu_sync) before its buffers are prepared again */
		if (c->pin_desc)
			cudaFreeHost(c->pin_desc);
		c->pin_desc = nullptr;
		c->pin_desc_cap = 0;
		const size_t want = desc_bytes + desc_bytes / 4 + 4096;
		CK(cudaHostAlloc((void **)&c->pin_desc, want, cudaHostAllocDefault));
		c->pin_desc_cap = want;
	}
	if (c->profile)
		CK(cudaEventRecord(c->ev_h2d0, c->stream));
	{
		size_t off = 0;
		for (const Part &pt : parts)
		{
			if (!pt.bytes)
				continue;
			memcpy(c->pin_desc + off, pt.src, pt.bytes);
			CK(cudaMemcpyAsync(pt.dst, c->pin_desc + off, pt.bytes, cudaMemcpyHostToDevice, c->stream));
			off += (pt.bytes + 255) & ~(size_t)255;
		}
	}
	if (b->bitstream_bytes)
		CK(cudaMemcpyAsync(c->bits.p, b->bitstream, b->bitstream_bytes, cudaMemcpyHostToDevice, c->stream));
	if (c->profile)
		CK(cudaEventRecord(c->ev_h2d1, c->stream));
	return 0;
}

/* grow the big device buffers ahead of time so that a later jcu_upload of at most this size meets no cudaMalloc */
extern "C" int jcu_reserve(jcu_ctx *c, size_t bitstream_bytes, size_t coef_elems, size_t out_bytes)
{
	CK(cudaSetDevice(c->device));
	CK(ensure(c->bits, bitstream_bytes + 64));
	CK(ensure(c->uslab, bitstream_bytes + 32 * (bitstream_bytes / 64 + 4)));
	CK(ensure(c->useg_len, sizeof(uint32_t) * (bitstream_bytes / 64 + 1)));
	CK(ensure(c->segs, sizeof(uint32_t) * (bitstream_bytes / 64 + 1)));
	CK(ensure(c->coef, sizeof(int16_t) * coef_elems + 256));
	CK(ensure(c->out, out_bytes + 256));
	CK(ensure(c->images, sizeof(jcu_image) * 4));
	CK(ensure(c->scans, sizeof(jcu_scan) * 64));
	CK(ensure(c->huff, sizeof(jcu_huff_spec) * 64));
	CK(ensure(c->work, sizeof(EntWork) * 4096));
	CK(ensure(c->err, sizeof(int32_t) * 4));
	CK(ensure(c->lists, sizeof(int32_t) * 16));
	return 0;
}

__global__ void k_publish_status(const int32_t *__restrict__ err, int n, const int32_t *__restrict__ ss_result, int32_t *__restrict__ host)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n)
		host[i] = err[i];
	if (i < 2)
		host[n + i] = ss_result ? ss_result[i] : 0;
}

extern "C" int jcu_run(jcu_ctx *c)
{
	CK(cudaSetDevice(c->device));
	c->launches = 0;
	cudaStream_t st = c->stream;
	if (c->profile)
		CK(cudaEventRecord(c->ev_run0, st)); /* start of the compute region */
	CK(cudaMemsetAsync(c->err.p, 0, sizeof(int32_t) * std::max(1, c->n_images), st));
	if (c->need_clear && c->coef_elems)
		CK(cudaMemsetAsync(c->coef.p, 0, sizeof(int16_t) * c->coef_elems, st));

	const jcu_image *d_img = (const jcu_image *)c->images.p;
	const int16_t *d_coef = (const int16_t *)c->coef.p;
	PcPipe no_pipe;
	no_pipe.preds = nullptr;
	no_pipe.progress = nullptr;
	no_pipe.ticket = nullptr;
	if (!c->pipe_work_h.empty())
	{
		/* progressive files without restart markers: every scan of an image at once, each behind its predecessors
		   (jpeg_entropy_prog.cuh, PcPipe); the word behind the progress array is the work ticket */
		PcPipe pp;
		pp.preds = (const int32_t *)c->pipe_preds.p;
		pp.progress = (int32_t *)c->pipe_progress.p;
		pp.ticket = (uint32_t *)c->pipe_progress.p + c->n_scans;
		CK(cudaMemsetAsync(c->pipe_progress.p, 0, sizeof(int32_t) * ((size_t)c->n_scans + 4), st));
		k_prog_coop<<<(unsigned)c->pipe_work_h.size(), PC_THREADS, 0, st>>>((const EntWork *)c->pipe_work.p, (const jcu_scan *)c->scans.p, d_img,
																								  (const uint32_t *)c->segs.p, (const jcu_huff_spec *)c->huff.p,
																								  (const uint8_t *)c->bits.p, (int16_t *)c->coef.p, (int32_t *)c->err.p, pp);
		c->launches++;
	}
	for (int w = 0; w < c->n_waves; ++w)
	{
		const int n = c->wave_off[w + 1] - c->wave_off[w], nb = c->wave_nbase[w];
		const EntWork *wk = (const EntWork *)c->work.p + c->wave_off[w];
		if (nb > 0)
		{
			k_unstuff<<<nb, ENTB_THREADS, 0, st>>>(wk, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p,
																(const uint8_t *)c->bits.p, (uint8_t *)c->uslab.p, (uint32_t *)c->useg_len.p);
			c->launches++;
			if (c->wave_two[w])
				k_entropy_baseline<2, 2><<<nb, ENTB_THREADS, ENTB_DYN_SMEM, st>>>(
					wk, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p, (const jcu_huff_spec *)c->huff.p,
					(const uint8_t *)c->uslab.p, (const uint32_t *)c->useg_len.p, (int16_t *)c->coef.p, (int32_t *)c->err.p,
					(const uint8_t *)c->bits.p);
			else
				k_entropy_baseline<4, 4><<<nb, ENTB_THREADS, ENTB_DYN_SMEM, st>>>(
					wk, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p, (const jcu_huff_spec *)c->huff.p,
					(const uint8_t *)c->uslab.p, (const uint32_t *)c->useg_len.p, (int16_t *)c->coef.p, (int32_t *)c->err.p,
					(const uint8_t *)c->bits.p);
			c->launches++;
		}
		const int nsm = c->wave_nsmall[w], nco = c->wave_ncoop[w];
		if (nco > 0)
		{
			const EntWork *wc = wk + (n - nco);
			k_prog_coop<<<nco, PC_THREADS, 0, st>>>(wc, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p,
																(const jcu_huff_spec *)c->huff.p, (const uint8_t *)c->bits.p, (int16_t *)c->coef.p,
																(int32_t *)c->err.p, no_pipe);
			c->launches++;
		}
		if (nsm > 0)
		{
			k_entropy_segments<ENT_SMALL_THREADS><<<nsm, ENT_SMALL_THREADS, 0, st>>>(
				wk + nb, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p, (const jcu_huff_spec *)c->huff.p,
				(const uint8_t *)c->bits.p, (int16_t *)c->coef.p, (int32_t *)c->err.p);
			c->launches++;
		}
		if (n - nb - nsm - nco > 0)
		{
			k_entropy_segments<ENT_THREADS><<<n - nb - nsm - nco, ENT_THREADS, 0, st>>>(
				wk + nb + nsm, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p, (const jcu_huff_spec *)c->huff.p,
				(const uint8_t *)c->bits.p, (int16_t *)c->coef.p, (int32_t *)c->err.p);
			c->launches++;
		}
	}
	if (!c->ss_scan_h.empty())
	{
		/* self-synchronising decode of the restart-less baseline scans (jpeg_selfsync.cuh) */
		SsBuffers sb;
		sb.exit_a = (SsState *)c->ss_exit_a.p;
		sb.exit_b = (SsState *)c->ss_exit_b.p;
		sb.entry_used = (SsState *)c->ss_entry.p;
		sb.nblk = (uint32_t *)c->ss_nblk.p;
		sb.first_blk = (uint32_t *)c->ss_first.p;
		sb.changed = (int32_t *)c->ss_changed.p;
		sb.sync = (uint32_t *)c->ss_changed.p + 4;
		sb.result = (int32_t *)c->ss_changed.p + 8;
		sb.ticket = (uint32_t *)c->ss_changed.p + 12;
		sb.sub_bytes = c->ss_sub_bytes;
		const unsigned nw = (unsigned)c->ss_work_h.size(), ns = (unsigned)c->ss_scan_h.size();
		const unsigned ndw = (unsigned)c->ss_dwork_h.size();
		const SsWork *wk = (const SsWork *)c->ss_work.p, *dwk = (const SsWork *)c->ss_dwork.p;
		const SsScan *dscans = (const SsScan *)c->ss_dscans.p;
#define SS_ARGS dwk, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p, (const jcu_huff_spec *)c->huff.p, \
					 (const uint8_t *)c->bits.p, (const uint8_t *)c->uslab.p, (const uint32_t *)c->useg_len.p, (int16_t *)c->coef.p, \
					 (int32_t *)c->err.p, sb
		/* phase U: count the bytes each 256-byte raw chunk keeps, prefix sum per scan, write the unstuffed stream */
		k_ss_unstuff<false><<<nw, SS_THREADS, 0, st>>>(wk, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p,
																	  (const uint8_t *)c->bits.p, (uint8_t *)c->uslab.p, (uint32_t *)c->useg_len.p, sb);
		k_ss_scan<<<ns, 256, 0, st>>>((const SsScan *)c->ss_scans.p, sb.nblk, sb.first_blk);
		k_ss_unstuff<true><<<nw, SS_THREADS, 0, st>>>(wk, (const jcu_scan *)c->scans.p, d_img, (const uint32_t *)c->segs.p,
																	 (const uint8_t *)c->bits.p, (uint8_t *)c->uslab.p, (uint32_t *)c->useg_len.p, sb);
		if (c->ss_two)
			k_ss_decode<SS_GUESS, EntbShared2><<<ndw, SS_THREADS, 0, st>>>(SS_ARGS);
		else
			k_ss_decode<SS_GUESS, EntbShared><<<ndw, SS_THREADS, 0, st>>>(SS_ARGS);
		c->launches += 4;
		/* the Jacobi rounds run to convergence inside ONE cooperative launch (grid barrier per round, jpeg_selfsync.cuh):
		   no flag copy, no stream synchronisation, the host never learns the round count before jcu_image_errors */
		CK(cudaMemsetAsync(c->ss_changed.p, 0, 64, st));
		{
			const SsWork *a_work = dwk;
			unsigned a_nw = ndw, a_max = c->ss_max_dsub + 2u;
			const jcu_scan *a_scans = (const jcu_scan *)c->scans.p;
			const uint32_t *a_segs = (const uint32_t *)c->segs.p;
			const jcu_huff_spec *a_huff = (const jcu_huff_spec *)c->huff.p;
			const uint8_t *a_uslab = (const uint8_t *)c->uslab.p;
			const uint32_t *a_ulen = (const uint32_t *)c->useg_len.p;
			void *args[] = {&a_work, &a_nw, &a_scans, &d_img, &a_segs, &a_huff, &a_uslab, &a_ulen, &sb, &a_max};
			const unsigned grid = std::min<unsigned>(ndw, (unsigned)std::max(1, c->ss_two ? c->ss_coop_ctas2 : c->ss_coop_ctas));
			const void *fn = c->ss_two ? (const void *)k_ss_refine_loop<EntbShared2> : (const void *)k_ss_refine_loop<EntbShared>;
			CK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(SS_THREADS), args, 0, st));
		}
		c->launches++;
		k_ss_scan<<<ns, 256, 0, st>>>(dscans, sb.nblk, sb.first_blk);
		if (!c->need_clear)
		{
			k_ss_zero_shared<<<ndw, SS_THREADS, 0, st>>>(dwk, (const jcu_scan *)c->scans.p, d_img, (int16_t *)c->coef.p, sb);
			c->launches++;
		}
		if (c->ss_two) /* entry states: side sb.result[0] */
			k_ss_decode<SS_WRITE, EntbShared2><<<ndw, SS_THREADS, 0, st>>>(SS_ARGS);
		else
			k_ss_decode<SS_WRITE, EntbShared><<<ndw, SS_THREADS, 0, st>>>(SS_ARGS);
		k_ss_dc<<<dim3(ns, 4), SS_DC_THREADS, 0, st>>>((const SsScan *)c->ss_scans.p, (const jcu_scan *)c->scans.p, d_img, (int16_t *)c->coef.p);
		c->launches += 3;
#undef SS_ARGS
	}
	CK(cudaGetLastError());
	if (c->profile)
		CK(cudaEventRecord(c->ev_ent1, st));

	/* ---- output stage ---- */
	if (c->list_n[JCU_PATH_GENERIC] > 0)
	{
		const int32_t *lst = (const int32_t *)c->lists.p + c->list_off[JCU_PATH_GENERIC];
		dim3 g1((unsigned)((c->max_blocks_generic + 127) / 128), (unsigned)c->list_n[JCU_PATH_GENERIC], 4);
		k_generic_idct<<<g1, 128, 0, st>>>(lst, d_img, d_coef, (uint8_t *)c->planes.p);
		long long want = (c->max_pixels_generic + 255) / 256;
		dim3 g2((unsigned)std::min<long long>(std::max<long long>(want, 1), 65535), (unsigned)c->list_n[JCU_PATH_GENERIC], 1);
		k_generic_color<<<g2, 256, 0, st>>>(lst, d_img, (const uint8_t *)c->planes.p, (uint8_t *)c->out.p);
		c->launches += 2;
	}
	for (int p = 1; p < JCU_PATH_COUNT; ++p)
		if (c->list_n[p] > 0)
		{
			const int32_t *lst = (const int32_t *)c->lists.p + c->list_off[p];
			cudaError_t e = p == JCU_PATH_FUSED_422
									  ? fused422_launch(c->path_maxtiles[p], lst, c->list_n[p], d_img, d_coef, (uint8_t *)c->out.p, st,
															  c->path_n[p], c->path_deq[p])
								  : p == JCU_PATH_FUSED_420
									  ? fused420_launch(c->path_maxtiles[p], lst, c->list_n[p], d_img, d_coef, (uint8_t *)c->out.p, st,
															  c->path_n[p], c->path_deq[p], c->rows_per_band)
								  : p == JCU_PATH_FUSED_444
									  ? fused1x1_launch<3>(c->path_maxtiles[p], lst, c->list_n[p], d_img, d_coef, (uint8_t *)c->out.p, st,
																  c->path_n[p], c->path_deq[p])
									  : fused1x1_launch<1>(c->path_maxtiles[p], lst, c->list_n[p], d_img, d_coef, (uint8_t *)c->out.p, st,
																  c->path_n[p], c->path_deq[p]);
			if (e != cudaSuccess)
				return set_err(c, e, "fused_launch");
			c->launches++;
		}
	CK(cudaGetLastError());
	if (c->profile)
		CK(cudaEventRecord(c->ev_out1, st));
	if (c->err_host_dev)
	{
		const int n = std::max(0, c->n_images);
		k_publish_status<<<(n + 2 + 255) / 256, 256, 0, st>>>((const int32_t *)c->err.p, n,
																				c->ss_scan_h.empty() ? nullptr : (const int32_t *)c->ss_changed.p + 8, c->err_host_dev);
		CK(cudaGetLastError());
	}
	return 0;
}

extern "C" int jcu_sync(jcu_ctx *c)
{
	CK(cudaSetDevice(c->device));
	if (c->dl_open)
	{
		if (c->profile)
			CK(cudaEventRecord(c->ev_d2h1, c->dl_use));
		CK(cudaEventRecord(c->ev_dl, c->dl_use));
	}
	CK(cudaStreamSynchronize(c->stream));
	if (c->dl_open)
	{
		CK(cudaEventSynchronize(c->ev_dl));
		c->dl_open = false;
	}
	if (c->bounce)
		CK(c->bounce->drain(&c->bounce_own)); /* pieces on their way from the pinned ring into pageable caller memory */
	return 0;
}

/* tuning: downloads of `c` are queued on `owner`'s download stream (one FIFO for a whole chunk pipeline) */
extern "C" void jcu_share_download_stream(jcu_ctx *c, jcu_ctx *owner)
{
	if (c && owner)
		c->dl_use = owner->dl_own;
}

extern "C" int jcu_download(jcu_ctx *c, size_t off, size_t bytes, void *dst)
{
	CK(cudaSetDevice(c->device));
	if (!c->dl_open)
	{
		/* first download after a run: the download stream waits for the kernels */
		CK(cudaEventRecord(c->ev_ready, c->stream));
		CK(cudaStreamWaitEvent(c->dl_use, c->ev_ready, 0));
		if (c->profile)
			CK(cudaEventRecord(c->ev_d2h0, c->dl_use));
		c->dl_open = true;
		c->d2h_open = true;
	}
	if (bytes >= ((size_t)1 << 20) && !host_ptr_is_pinned(dst))
	{
		if (!c->bounce)
		{
			cudaError_t e = cudaSuccess;
			c->bounce = bounce_for_device(c->device, c->host_threads_hint, &e);
			if (!c->bounce)
				return set_err(c, e, "jcu_download: pinned staging ring");
		}
		/* a single image (a few MB) is cut into about eight pieces so that as many copier threads share it */
		const size_t piece = std::min(Bounce::SLOT_BYTES, std::max((size_t)256 << 10, ((bytes / 8 + 4095) & ~(size_t)4095)));
		for (size_t done = 0; done < bytes; done += piece)
		{
			const size_t n = std::min(piece, bytes - done);
			const int slot = c->bounce->take_slot();
			cudaError_t e = cudaMemcpyAsync(c->bounce->buf[slot], (const uint8_t *)c->out.p + off + done, n, cudaMemcpyDeviceToHost, c->dl_use);
			if (e == cudaSuccess)
				e = cudaEventRecord(c->bounce->ev[slot], c->dl_use);
			c->bounce->submit(slot, (uint8_t *)dst + done, e == cudaSuccess ? n : 0, &c->bounce_own); /* the slot goes back through the workers either way */
			CK(e);
		}
		return 0;
	}
	CK(cudaMemcpyAsync(dst, (const uint8_t *)c->out.p + off, bytes, cudaMemcpyDeviceToHost, c->dl_use));
	return 0;
}

/* ---- widening epilogues (jpeg_widen.cuh) ---- */
extern "C" int jcu_widen(jcu_ctx *c, size_t off, size_t count, int elem_bytes, int comp, int nonalpha, const float *lut512)
{
	CK(cudaSetDevice(c->device));
	if ((elem_bytes != 2 && elem_bytes != 4) || comp < 1 || (off & 15) || off + count > c->out.cap)
		return set_err(c, cudaErrorInvalidValue, "jcu_widen: bad arguments");
	CK(ensure(c->wide, count * (size_t)elem_bytes + 256));
	if (count == 0)
		return 0;
	const unsigned grid = (unsigned)((count + (size_t)WIDEN_THREADS * WIDEN_PER_THREAD - 1) / ((size_t)WIDEN_THREADS * WIDEN_PER_THREAD));
	if (elem_bytes == 2)
		k_widen16<<<grid, WIDEN_THREADS, 0, c->stream>>>((const uint8_t *)c->out.p + off, (uint16_t *)c->wide.p, count);
	else
	{
		CK(ensure(c->wide_lut, 512 * sizeof(float)));
		CK(cudaMemcpyAsync(c->wide_lut.p, lut512, 512 * sizeof(float), cudaMemcpyHostToDevice, c->stream));
		k_widenf<<<grid, WIDEN_THREADS, 0, c->stream>>>((const uint8_t *)c->out.p + off, (float *)c->wide.p, count, comp, nonalpha,
																		(const float *)c->wide_lut.p);
	}
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

extern "C" int jcu_download_wide(jcu_ctx *c, size_t bytes, void *dst)
{
	CK(cudaSetDevice(c->device));
	CK(cudaMemcpyAsync(dst, c->wide.p, bytes, cudaMemcpyDeviceToHost, c->stream)); /* same stream as the kernel; jcu_sync waits */
	return 0;
}

extern "C" void *jcu_device_wide(jcu_ctx *c) { return c->wide.p; }

extern "C" int jcu_image_errors(jcu_ctx *c, int32_t *errs, int n)
{
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream)); /* the last kernel of jcu_run has written the mapped host words */
	if (!c->err_host || n > c->n_images)
		return set_err(c, cudaErrorInvalidValue, "jcu_image_errors: no run to report");
	memcpy(errs, c->err_host, sizeof(int32_t) * (size_t)n);
	c->ss_rounds = c->ss_scan_h.empty() ? 0 : c->err_host[c->n_images + 1];
	return 0;
}

extern "C" void *jcu_device_out(jcu_ctx *c) { return c->out.p; }

extern "C" int jcu_get_timings(jcu_ctx *c, jcu_timings *t)
{
	memset(t, 0, sizeof(*t));
	t->launches = c->launches;
	t->ss_scans = (int32_t)c->ss_scan_h.size();
	t->ss_rounds = c->ss_rounds;
	if (!c->profile)
		return 0;
	CK(cudaSetDevice(c->device));
	cudaEventElapsedTime(&t->h2d_ms, c->ev_h2d0, c->ev_h2d1);
	cudaEventElapsedTime(&t->entropy_ms, c->ev_run0, c->ev_ent1);
	cudaEventElapsedTime(&t->idct_ms, c->ev_ent1, c->ev_out1);
	cudaEventElapsedTime(&t->total_ms, c->ev_run0, c->ev_out1);
	if (c->d2h_open)
		cudaEventElapsedTime(&t->d2h_ms, c->ev_d2h0, c->ev_d2h1);
	cudaGetLastError();
	return 0;
}

/* ---- encoder (jpeg_encode.cuh) ---- */
extern "C" int jcu_encode(jcu_ctx *c, const jcu_enc_image *imgs, int n, const uint8_t *pixels, const uint8_t *const *pix_ptrs,
								  size_t pixel_bytes, int pixels_on_device, const uint8_t *headers, const uint32_t *huff, size_t out_bytes, uint32_t *out_len, float *kernel_ms)
{
	CK(cudaSetDevice(c->device));
	if (n <= 0)
		return 0;
	cudaStream_t st = c->stream;
	size_t nblocks = 0;
	uint32_t max_blocks = 0;
	for (int i = 0; i < n; ++i)
	{
		nblocks = std::max(nblocks, (size_t)imgs[i].blk_first + (size_t)imgs[i].nblocks);
		max_blocks = std::max(max_blocks, (uint32_t)imgs[i].nblocks);
	}
	const size_t nchunks = nblocks * (ENC_BLOCK_BYTES / ENC_CHUNK);
	if (!c->ev_enc0)
	{
		CK(cudaEventCreate(&c->ev_enc0));
		CK(cudaEventCreate(&c->ev_enc1));
	}
	if (!pixels_on_device)
		CK(ensure(c->enc_pix, pixel_bytes + 16));
	CK(ensure(c->enc_img, sizeof(jcu_enc_image) * (size_t)n));
	CK(ensure(c->enc_coef, nblocks * 128 + 256));
	CK(ensure(c->enc_nbits, nblocks * 4 + 16));
	CK(ensure(c->enc_bitoff, nblocks * 4 + 16));
	CK(ensure(c->enc_tot, (size_t)n * 8 + 16)); /* [0, n): bit totals, [n, 2n): 0xFF totals */
	CK(ensure(c->enc_bitbuf, nblocks * ENC_BLOCK_BYTES + 256));
	CK(ensure(c->enc_ffcnt, nchunks * 4 + 16));
	CK(ensure(c->enc_ffoff, nchunks * 4 + 16));
	CK(ensure(c->enc_out, out_bytes + 16));
	CK(ensure(c->enc_len, (size_t)n * 4 + 16));
	CK(ensure(c->enc_hdr, (size_t)n * JCU_ENC_HDR_STRIDE));
	CK(cudaMemcpyToSymbolAsync(c_enc_huff, huff, sizeof(uint32_t) * 4 * 256, 0, cudaMemcpyHostToDevice, st));
	if (!pixels_on_device)
		for (int i = 0; i < n; ++i)
			if (imgs[i].nblocks && pix_ptrs[i])
				CK(cudaMemcpyAsync((uint8_t *)c->enc_pix.p + imgs[i].pix_off, pix_ptrs[i],
										 (size_t)imgs[i].width * imgs[i].height * imgs[i].comp, cudaMemcpyHostToDevice, st));
	CK(cudaMemcpyAsync(c->enc_img.p, imgs, sizeof(jcu_enc_image) * (size_t)n, cudaMemcpyHostToDevice, st));
	CK(cudaMemcpyAsync(c->enc_hdr.p, headers, (size_t)n * JCU_ENC_HDR_STRIDE, cudaMemcpyHostToDevice, st));
	const uint8_t *d_pix = pixels_on_device ? pixels : (const uint8_t *)c->enc_pix.p;
	const jcu_enc_image *d_img = (const jcu_enc_image *)c->enc_img.p;
	uint32_t *d_tot = (uint32_t *)c->enc_tot.p;
	CK(cudaEventRecord(c->ev_enc0, st));
	CK(cudaMemsetAsync(c->enc_bitbuf.p, 0, nblocks * ENC_BLOCK_BYTES, st));
	const dim3 gb((max_blocks + ENC_THREADS - 1) / ENC_THREADS, (unsigned)n);
	const dim3 gc((max_blocks * (ENC_BLOCK_BYTES / ENC_CHUNK) + ENC_THREADS - 1) / ENC_THREADS, (unsigned)n);
	{
		uint32_t max_tiles = 0, max_generic = 0;
		for (int i = 0; i < n; ++i)
			if (imgs[i].subsample)
				max_tiles = std::max(max_tiles, (uint32_t)(((imgs[i].mcu_x + ENC_TILE_MCUS - 1) / ENC_TILE_MCUS) * imgs[i].mcu_y));
			else
				max_generic = std::max(max_generic, (uint32_t)imgs[i].nblocks);
		if (max_generic)
			k_enc_dct<<<dim3((max_generic + ENC_THREADS - 1) / ENC_THREADS, (unsigned)n), ENC_THREADS, 0, st>>>(d_img, d_pix, (int16_t *)c->enc_coef.p);
		if (max_tiles)
			k_enc_dct420<<<dim3(max_tiles, (unsigned)n), ENC_THREADS, 0, st>>>(d_img, d_pix, (int16_t *)c->enc_coef.p);
	}
	k_enc_count<<<gb, ENC_THREADS, 0, st>>>(d_img, (const int16_t *)c->enc_coef.p, (uint32_t *)c->enc_nbits.p);
	k_enc_scan<<<n, 1024, 0, st>>>(d_img, (const uint32_t *)c->enc_nbits.p, (uint32_t *)c->enc_bitoff.p, d_tot, nullptr, 0);
	k_enc_emit<<<gb, ENC_THREADS, 0, st>>>(d_img, (const int16_t *)c->enc_coef.p, (const uint32_t *)c->enc_bitoff.p, (uint8_t *)c->enc_bitbuf.p);
	k_enc_ffcount<<<gc, ENC_THREADS, 0, st>>>(d_img, (const uint8_t *)c->enc_bitbuf.p, d_tot, (uint32_t *)c->enc_ffcnt.p);
	k_enc_scan<<<n, 1024, 0, st>>>(d_img, (const uint32_t *)c->enc_ffcnt.p, (uint32_t *)c->enc_ffoff.p, d_tot + n, d_tot, 1);
	k_enc_stuff<<<gc, ENC_THREADS, 0, st>>>(d_img, (const uint8_t *)c->enc_bitbuf.p, d_tot, (const uint32_t *)c->enc_ffoff.p, d_tot + n,
														  (const uint8_t *)c->enc_hdr.p, (uint8_t *)c->enc_out.p, (uint32_t *)c->enc_len.p);
	CK(cudaGetLastError());
	CK(cudaEventRecord(c->ev_enc1, st));
	CK(cudaMemcpyAsync(out_len, c->enc_len.p, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, st));
	CK(cudaStreamSynchronize(st));
	if (kernel_ms)
		cudaEventElapsedTime(kernel_ms, c->ev_enc0, c->ev_enc1);
	c->launches = 7;
	return 0;
}

extern "C" int jcu_encode_fetch(jcu_ctx *c, size_t off, size_t bytes, void *dst)
{
	CK(cudaSetDevice(c->device));
	CK(cudaMemcpyAsync(dst, (const uint8_t *)c->enc_out.p + off, bytes, cudaMemcpyDeviceToHost, c->stream));
	return 0;
}

extern "C" int jcu_encode_sync(jcu_ctx *c)
{
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

extern "C" void *jcu_encode_device_out(jcu_ctx *c) { return c->enc_out.p; }

/* ---- stage-level entry points: the IDCT stage alone (jcu_test_idct_blocks, and its allocation-free twin below) ---- */
__global__ void k_test_idct(const int16_t *coef, const uint16_t *qt, int n, uint8_t *out)
{
	int b = blockIdx.x * blockDim.x + threadIdx.x;
	if (b >= n)
		return;
	int v[64];
	uint32_t rows[16];
	load_dequant(coef + (size_t)b * 64, qt, v);
	idct_block_regs(v, rows);
	uint32_t *o = reinterpret_cast<uint32_t *>(out + (size_t)b * 64);
	for (int i = 0; i < 16; ++i)
		o[i] = rows[i];
}
extern "C" int jcu_test_idct_blocks(jcu_ctx *c, const int16_t *coef, const uint16_t *qt, int nblocks, uint8_t *out64)
{
	CK(cudaSetDevice(c->device));
	int16_t *d_c;
	uint16_t *d_q;
	uint8_t *d_o;
	CK(cudaMalloc(&d_c, (size_t)nblocks * 128));
	CK(cudaMalloc(&d_q, 128));
	CK(cudaMalloc(&d_o, (size_t)nblocks * 64));
	{
		/* callers pass natural-order blocks: store them in the slab layout (jpeg_coef_layout.cuh) */
		std::vector<int16_t> sw((size_t)nblocks * 64);
		for (int b = 0; b < nblocks; ++b)
		{
			const unsigned key = coef_key(d_c + (size_t)b * 64);
			for (int r = 0; r < 8; ++r)
				memcpy(&sw[(size_t)b * 64 + ((r ^ (int)key) * 8)], coef + (size_t)b * 64 + r * 8, 16);
		}
		CK(cudaMemcpy(d_c, sw.data(), (size_t)nblocks * 128, cudaMemcpyHostToDevice));
	}
	CK(cudaMemcpy(d_q, qt, 128, cudaMemcpyHostToDevice));
	k_test_idct<<<(nblocks + 63) / 64, 64, 0, c->stream>>>(d_c, d_q, nblocks, d_o);
	CK(cudaStreamSynchronize(c->stream));
	CK(cudaMemcpy(out64, d_o, (size_t)nblocks * 64, cudaMemcpyDeviceToHost));
	cudaFree(d_c);
	cudaFree(d_q);
	cudaFree(d_o);
	return 0;
}

extern "C" int jcu_read_coefs(jcu_ctx *c, size_t elem_off, size_t elems, int16_t *dst)
{
	CK(cudaSetDevice(c->device));
	CK(cudaMemcpyAsync(dst, (const int16_t *)c->coef.p + elem_off, elems * sizeof(int16_t), cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	/* undo the slab's chunk permutation (jpeg_coef_layout.cuh); callers pass whole blocks */
	for (size_t b = 0; b + 64 <= elems; b += 64)
	{
		const unsigned key = coef_key((const int16_t *)c->coef.p + elem_off + b);
		int16_t tmp[64];
		memcpy(tmp, dst + b, sizeof(tmp));
		for (int r = 0; r < 8; ++r)
			memcpy(dst + b + r * 8, tmp + ((r ^ (int)key) * 8), 16);
	}
	return 0;
}

/* ---- allocation-free stage entry point (include/jpeg_cuda_abi.h: jcu_idct_blocks_noalloc) ----
 * jcu_test_idct_blocks keeps its slab-layout copy of the input blocks in a cudaMalloc'ed scratch buffer.  Callers
 * that may not allocate (CUDA graph capture, serving loops) get the same kernel over a scratch slab that lives in
 * this file-scope device array: every GPU that loads the library holds its 4 MiB until exit. */
#include <pthread.h>
static __device__ __align__(256) int16_t g_noalloc_slab[(size_t)JCU_NOALLOC_MAX_BLOCKS * 64];

#define JCU_NOALLOC_MAX_DEVICES 64
/* In host code `g_noalloc_slab` is only a handle for the runtime, not a device address: the address each device
   gave its copy of the array is looked up once per device (with that device current) and kept here. */
static int16_t *g_noalloc_ptr[JCU_NOALLOC_MAX_DEVICES];
static pthread_mutex_t g_noalloc_mu = PTHREAD_MUTEX_INITIALIZER;

static cudaError_t noalloc_slab_address(int device, int16_t **out)
{
	*out = nullptr;
	if (device < 0 || device >= JCU_NOALLOC_MAX_DEVICES)
		return cudaErrorInvalidDevice;
	cudaError_t e = cudaSuccess;
	pthread_mutex_lock(&g_noalloc_mu);
	if (!g_noalloc_ptr[device])
	{
		void *p = nullptr;
		e = cudaGetSymbolAddress(&p, g_noalloc_slab); /* the caller made `device` current */
		if (e == cudaSuccess && !p)
			e = cudaErrorInvalidSymbol;
		if (e == cudaSuccess)
			g_noalloc_ptr[device] = (int16_t *)p;
	}
	*out = g_noalloc_ptr[device];
	pthread_mutex_unlock(&g_noalloc_mu);
	return e;
}

/* natural-order blocks (caller's memory) -> slab layout (jpeg_coef_layout.cuh) in the scratch: one 16-byte chunk
   (= one coefficient row) per thread.  Every chunk of blocks [0, n) is rewritten, so nothing a previous call
   left in the scratch is ever read and nothing needs clearing. */
__global__ void k_noalloc_stage(const uint4 *__restrict__ in, int16_t *slab, int n)
{
	const int g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= n * 8)
		return;
	const int b = g >> 3, r = g & 7;
	int16_t *blk = slab + (size_t)b * 64;
	reinterpret_cast<uint4 *>(blk)[r ^ (int)coef_key(blk)] = in[g];
}

extern "C" int jcu_idct_blocks_noalloc(jcu_ctx *c, const int16_t *d_coef, const uint16_t *d_qt, int nblocks, uint8_t *d_out64,
													void *stream)
{
	if (!c)
		return (int)cudaErrorInvalidValue;
	if (nblocks < 0 || nblocks > JCU_NOALLOC_MAX_BLOCKS)
	{
		snprintf(c->errstr, sizeof(c->errstr), "jcu_idct_blocks_noalloc: %d blocks exceed the scratch of JCU_NOALLOC_MAX_BLOCKS = %d blocks",
					nblocks, JCU_NOALLOC_MAX_BLOCKS);
		return (int)cudaErrorInvalidValue;
	}
	if (nblocks == 0)
		return 0;
	if (!d_coef || !d_qt || !d_out64 || ((uintptr_t)d_coef & 15) || ((uintptr_t)d_qt & 1) || ((uintptr_t)d_out64 & 3))
	{
		snprintf(c->errstr, sizeof(c->errstr), "jcu_idct_blocks_noalloc: null or misaligned device pointer (coef 16, qt 2, out 4 bytes)");
		return (int)cudaErrorInvalidValue;
	}
	CK(cudaSetDevice(c->device));
	int16_t *slab = nullptr;
	CK(noalloc_slab_address(c->device, &slab));
	cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
	k_noalloc_stage<<<(nblocks * 8 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const uint4 *>(d_coef), slab, nblocks);
	k_test_idct<<<(nblocks + 63) / 64, 64, 0, st>>>(slab, d_qt, nblocks, d_out64);
	CK(cudaGetLastError());
	return 0;
}
