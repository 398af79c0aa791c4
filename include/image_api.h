/*
 * image_api.h -- public C API of the B200-native JPEG decode path.
 *
 * Drop-in for the JPEG leg of rpehkone/image-codecs.  The reference ships this header EMPTY
 * (SURVEY.md F2); its public names exist only as STBIDEF definitions in convert.c and
 * image_api.c.  Every per-image entry point below keeps the reference's name, argument order,
 * types, ownership rule and error convention, citing the definition it replaces:
 *
 *   stbi_load                  convert.c:188      stbi_info                  image_api.c:74
 *   stbi_load_from_file        convert.c:199      stbi_info_from_file        image_api.c:85
 *   stbi_load_from_memory      convert.c:254      stbi_info_from_memory      image_api.c:119
 *   stbi_load_from_callbacks   convert.c:261      stbi_info_from_callbacks   image_api.c:126
 *
 * Behaviour kept from the reference:
 *   - the returned pixel buffer is malloc'ed by the library, caller frees it with
 *     stbi_image_free (== free);
 *   - *x,*y = image size, *comp = components IN THE FILE (3 or 1), not req_comp (jpeg.c:2436);
 *   - output has (req_comp ? req_comp : *comp) channels, tightly packed, top-down;
 *   - failure returns NULL / 0 and stbi_failure_reason() gives the reference's short string;
 *   - only JPEG is recognised: anything else fails with "unknown image type" (image_api.c:55).
 *
 * The decode itself runs on the GPU (hand-written sm_100a CUDA, image_codecs_b200/csrc).
 * There is NO CPU fallback: without a usable CUDA device the load functions fail with
 * "no cuda device".  stbi_info* is header-only and needs no GPU (jpeg.c:2466).
 *
 * NEW beside the per-image calls (BASELINE.json north_star): a batched decoder,
 * stbi_jpeg_batch_*, at the bottom of this file.
 */
#ifndef IMAGE_API_H
#define IMAGE_API_H

#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef STBIDEF
#define STBIDEF extern
#endif

typedef unsigned char stbi_uc;
typedef unsigned short stbi_us;

enum
{
	STBI_default = 0, /* only used for req_comp */
	STBI_grey = 1,
	STBI_grey_alpha = 2,
	STBI_rgb = 3,
	STBI_rgb_alpha = 4
};

/* shape inferred from the reference's uses at common.c:12,49,79 */
typedef struct
{
	int (*read)(void *user, char *data, int size); /* fill 'data' with 'size' bytes; return bytes read */
	void (*skip)(void *user, int n);					  /* skip the next 'n' bytes */
	int (*eof)(void *user);								  /* nonzero at end of data */
} stbi_io_callbacks;

/* ---- per-image load (reference convert.c:188-266) ---- */
STBIDEF stbi_uc *stbi_load(char const *filename, int *x, int *y, int *comp, int req_comp);
STBIDEF stbi_uc *stbi_load_from_file(FILE *f, int *x, int *y, int *comp, int req_comp);
STBIDEF stbi_uc *stbi_load_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp, int req_comp);
STBIDEF stbi_uc *stbi_load_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp,
														int req_comp);

/* ---- header-only queries (reference image_api.c:74-131; JPEG leg jpeg.c:2482) ---- */
STBIDEF int stbi_info(char const *filename, int *x, int *y, int *comp);
STBIDEF int stbi_info_from_file(FILE *f, int *x, int *y, int *comp);
STBIDEF int stbi_info_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp);
STBIDEF int stbi_info_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp);

/* ---- 16-bit and float variants (reference convert.c:213-252, 286-337): the 8-bit GPU decode is widened ON THE
 *      DEVICE by a store epilogue (v * 257; a 512-entry table of the reference's pow(v / 255, gamma) * scale built by
 *      the host's pow()), so the wide samples are what crosses the bus; same values as the reference's host loop ---- */
STBIDEF stbi_us *stbi_load_16(char const *filename, int *x, int *y, int *comp, int req_comp);
STBIDEF stbi_us *stbi_load_from_file_16(FILE *f, int *x, int *y, int *comp, int req_comp);
STBIDEF stbi_us *stbi_load_16_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp, int req_comp);
STBIDEF stbi_us *stbi_load_16_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp,
															int req_comp);
STBIDEF float *stbi_loadf(char const *filename, int *x, int *y, int *comp, int req_comp);
STBIDEF float *stbi_loadf_from_file(FILE *f, int *x, int *y, int *comp, int req_comp);
STBIDEF float *stbi_loadf_from_memory(stbi_uc const *buffer, int len, int *x, int *y, int *comp, int req_comp);
STBIDEF float *stbi_loadf_from_callbacks(stbi_io_callbacks const *clbk, void *user, int *x, int *y, int *comp,
													  int req_comp);
STBIDEF void stbi_ldr_to_hdr_gamma(float gamma); /* convert.c:404-411 */
STBIDEF void stbi_ldr_to_hdr_scale(float scale);
STBIDEF void stbi_hdr_to_ldr_gamma(float gamma);
STBIDEF void stbi_hdr_to_ldr_scale(float scale);

/* JPEG is never HDR or 16-bit (reference convert.c:345-400, image_api.c:96-145 return 0 for it) */
STBIDEF int stbi_is_hdr(char const *filename);
STBIDEF int stbi_is_hdr_from_file(FILE *f);
STBIDEF int stbi_is_hdr_from_memory(stbi_uc const *buffer, int len);
STBIDEF int stbi_is_hdr_from_callbacks(stbi_io_callbacks const *clbk, void *user);
STBIDEF int stbi_is_16_bit(char const *filename);
STBIDEF int stbi_is_16_bit_from_file(FILE *f);
STBIDEF int stbi_is_16_bit_from_memory(stbi_uc const *buffer, int len);
STBIDEF int stbi_is_16_bit_from_callbacks(stbi_io_callbacks const *clbk, void *user);

/* ---- result ownership / errors / flags (names used by the reference, defined in its missing header) ---- */
STBIDEF void stbi_image_free(void *retval_from_stbi_load);
STBIDEF const char *stbi_failure_reason(void); /* thread-local */
STBIDEF void stbi_set_flip_vertically_on_load(int flag_true_if_should_flip); /* convert.c:97-101 */
/* PNG-only switches of the reference, accepted and ignored (a JPEG has neither alpha nor a CgBI chunk) */
STBIDEF void stbi_set_unpremultiply_on_load(int flag_true_if_should_unpremultiply);
STBIDEF void stbi_convert_iphone_png_to_rgb(int flag_true_if_should_convert);

/* ---- JPEG writer (reference codec/jpeg_write.c:368-390): same byte stream, encoded on the GPU ---- */
typedef void stbi_write_func(void *context, void *data, int size);
STBIDEF int stbi_write_jpg_to_func(stbi_write_func *func, void *context, int x, int y, int comp, const void *data, int quality);
STBIDEF int stbi_write_jpg(char const *filename, int x, int y, int comp, const void *data, int quality);
STBIDEF void stbi_flip_vertically_on_write(int flag);
/* batched (new): files are malloc'ed, free each with stbi_image_free; returns the number produced */
STBIDEF int stbi_jpeg_encode_batch(int device, const stbi_uc *const *pixels, const int *xs, const int *ys, const int *comps,
											  int count, int quality, stbi_uc **outs, int *out_lens, float *kernel_ms);
STBIDEF int stbi_jpeg_encode_batch_device(int device, const stbi_uc *const *pixels, const int *xs, const int *ys, const int *comps,
														int count, int quality, stbi_uc **outs, int *out_lens, float *kernel_ms);

/* ===================================================================================== */
/* Batched decode (new; semantics per image identical to stbi_load_from_memory)          */
/* ===================================================================================== */

typedef struct stbi_jpeg_batch stbi_jpeg_batch; /* opaque: owns one device's streams, pinned rings, scratch */

typedef struct
{
	int device;				 /* CUDA device ordinal */
	int host_threads;		 /* marker-parser worker threads; 0 = number of online cores (max 64) */
	int chunk_images;		 /* images per pipeline chunk for the host-buffer call; 0 = auto */
	int profile;			 /* 1 = record CUDA events around each stage (stbi_jpeg_batch_timings) */
	int entropy_mode;		 /* 0 = auto, 1 = one thread per restart segment, 2 = self-synchronising */
	/* tuning (0 = default) */
	int pipeline_depth;	 /* chunk objects in flight in stbi_jpeg_decode_batch: 0 = 3, 1..4 */
	int device_mem_mb;	 /* device memory ONE prepared chunk may ask for; 0 = auto (a share of the device).  A chunk
									 that would need more is cut and continued; an image that needs more by itself fails
									 alone with "outofmem" */
	int no_download_merge; /* bit 0: one device->host copy per image instead of merged copies of up to 64 MB; bit 1: all
									 pipeline objects queue their copies on ONE stream (FIFO) instead of one stream each */
	int trace;				 /* 1 = host timeline of the chunk pipeline on stderr */
	/* test switches (0 in production) */
	int force_generic;	 /* every image takes the generic two-pass output path */
	int rows_per_band;	 /* > 0: MCU rows per band of the 4:2:0 strip kernel */
	int debug_guards;		 /* canaries behind every device buffer, see stbi_jpeg_batch_debug_guards */
	int selfsync_sub_bytes; /* 32 / 64 / 128 / 256: bytes per subsequence of the self-synchronising decoder (0 = by batch size) */
} stbi_jpeg_batch_params;

/* per-image status values written by the batch calls */
enum
{
	STBI_JPEG_OK = 1,
	STBI_JPEG_FAILED = 0 /* reason: stbi_jpeg_batch_reason(batch, i) */
};

typedef struct
{
	float parse_ms;	/* host marker parse (wall) */
	float h2d_ms;		/* bitstream + descriptor upload */
	float entropy_ms; /* Huffman decode kernels (all scans) */
	float idct_ms;		/* fused dequant+IDCT+upsample+colour kernel(s) */
	float d2h_ms;		/* pixel download */
	float total_ms;	/* first to last event of the last run */
	double entropy_bits; /* entropy-coded bits consumed by the last run (incl. stuffed bytes, excl. markers) */
	double coef_bytes;	/* algorithmic bytes read by the fused kernel: 128 B x coefficient blocks */
	double out_bytes;		/* algorithmic bytes written: n x width x height summed */
	int launches;			/* kernels launched by the last run */
	int selfsync_scans;	/* scans decoded by the self-synchronising path (streams without restart markers) */
	int selfsync_rounds; /* refine rounds that path needed to converge */
	int reserved[3];
} stbi_jpeg_batch_timings;

STBIDEF stbi_jpeg_batch *stbi_jpeg_batch_create(const stbi_jpeg_batch_params *params);
STBIDEF void stbi_jpeg_batch_destroy(stbi_jpeg_batch *b);

/*
 * One-call batched decode with HOST buffers (the end-to-end path).
 *   bufs[i], lens[i]   compressed streams (any host memory; pinned is faster)
 *   req_comp           0..4, as stbi_load_from_memory
 *   outs[i]            caller buffers of out_caps[i] bytes (pinned is faster).  An image whose
 *                      x*y*n exceeds out_caps[i] fails with "output too small".
 *                      outs may be NULL: then only xs/ys/comps/status are produced (sizing pass).
 *   xs, ys, comps      per-image size and components-in-file (any may be NULL)
 *   status             per-image STBI_JPEG_OK / STBI_JPEG_FAILED
 * Returns the number of images decoded successfully, or -1 if the batch could not run at all
 * (reason in stbi_failure_reason()).  A corrupt image fails alone; the batch continues.
 */
STBIDEF int stbi_jpeg_decode_batch(stbi_jpeg_batch *b, const stbi_uc *const *bufs, const int *lens, int count,
											  int req_comp, stbi_uc *const *outs, const size_t *out_caps, int *xs, int *ys,
											  int *comps, int *status);

/*
 * Staged form, for callers that keep data resident on the device:
 *   prepare   parse markers on the host, upload bitstreams + descriptors, size device buffers
 *   run       launch entropy decode + fused IDCT/colour on the batch's stream (asynchronous)
 *   sync      wait for the stream
 *   fetch     copy image i's pixels to host memory
 * device_pixels gives the device address / byte offsets of the packed outputs.
 */
STBIDEF int stbi_jpeg_batch_prepare(stbi_jpeg_batch *b, const stbi_uc *const *bufs, const int *lens, int count,
												int req_comp, int *xs, int *ys, int *comps, int *status);
STBIDEF int stbi_jpeg_batch_run(stbi_jpeg_batch *b);
STBIDEF int stbi_jpeg_batch_sync(stbi_jpeg_batch *b);
STBIDEF int stbi_jpeg_batch_fetch(stbi_jpeg_batch *b, int image, stbi_uc *out, size_t out_cap);
STBIDEF int stbi_jpeg_batch_fetch_all(stbi_jpeg_batch *b, stbi_uc *const *outs, const size_t *out_caps);
STBIDEF void *stbi_jpeg_batch_device_pixels(stbi_jpeg_batch *b, int image, size_t *nbytes);
STBIDEF int stbi_jpeg_batch_status(stbi_jpeg_batch *b, int image);
STBIDEF const char *stbi_jpeg_batch_reason(stbi_jpeg_batch *b, int image);
STBIDEF int stbi_jpeg_batch_timings_get(stbi_jpeg_batch *b, stbi_jpeg_batch_timings *t);
/* test hook: a batch created with params.debug_guards = 1 carries a canary behind the logical size of every device
   buffer; 1 = all intact, 0 = one was overwritten, -1 = guard mode off */
STBIDEF int stbi_jpeg_batch_debug_guards(stbi_jpeg_batch *b);

/*
 * All GPUs of the box behind one call (BASELINE.json north_star: "independent image batches are sharded across the 8
 * GPUs of one box with no collective").  The object owns one stbi_jpeg_batch per device; stbi_jpeg_decode_batch_multi
 * splits the images into one contiguous range per device, balanced by compressed bytes (entropy-decode time follows
 * the bits), and runs stbi_jpeg_decode_batch on every range at once, each on its own host worker thread with its own
 * pinned staging, streams and parser threads (params.host_threads is per device; 0 = online cores / devices).
 * Arguments, results and per-image semantics are those of stbi_jpeg_decode_batch.  devices == NULL: devices
 * 0 .. n_devices-1 (n_devices <= 0: every visible device).
 */
typedef struct stbi_jpeg_multi stbi_jpeg_multi;
STBIDEF stbi_jpeg_multi *stbi_jpeg_multi_create(const int *devices, int n_devices, const stbi_jpeg_batch_params *params);
STBIDEF void stbi_jpeg_multi_destroy(stbi_jpeg_multi *m);
STBIDEF int stbi_jpeg_multi_device_count(const stbi_jpeg_multi *m);
STBIDEF int stbi_jpeg_decode_batch_multi(stbi_jpeg_multi *m, const stbi_uc *const *bufs, const int *lens, int count, int req_comp,
													  stbi_uc *const *outs, const size_t *out_caps, int *xs, int *ys, int *comps, int *status);
/* images [first[d], first[d+1]) went to device d in the last call (n_devices + 1 entries); returns n_devices */
STBIDEF int stbi_jpeg_multi_last_split(const stbi_jpeg_multi *m, int *first, int cap);

/* Device used by the per-image entry points (stbi_load*, stbi_write_jpg*); default 0.  Takes effect for contexts
   created afterwards. */
STBIDEF void stbi_jpeg_set_default_device(int device);

/*
 * Host-only full marker parse of one stream (no GPU): what the batch loader would hand to the
 * device.  Returns 1 if the stream parses (0 + stbi_failure_reason() otherwise) and fills
 * info[16] = { width, height, comps_in_file, progressive, n_scans, n_restart_segments,
 *              output_path (JCU_PATH_*), colour_mode (JCU_CM_*), n_out, decode_n, hmax, vmax,
 *              mcu_x, mcu_y, coefficient_blocks, entropy_coded_bytes }.
 */
STBIDEF int stbi_jpeg_inspect(stbi_uc const *buffer, int len, int req_comp, long long info[16]);

/*
 * Host-only: the decode schedule of one stream.  Fills waves[i] (i < cap) with the dependency level of scan
 * i -- scans of one level touch disjoint (component, coefficient band) sets and are entropy decoded by the
 * same launch; the reference runs them one after the other in file order (jpeg.c:1716).  Returns the number
 * of scans, 0 if the stream does not parse.
 */
STBIDEF int stbi_jpeg_inspect_scans(stbi_uc const *buffer, int len, int *waves, int cap);

/* number of CUDA devices visible (0 when there is none); never falls back to a CPU decode */
STBIDEF int stbi_jpeg_cuda_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* IMAGE_API_H */
