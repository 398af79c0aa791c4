/*
 * jpeg_cuda_abi.h -- the thin C-ABI between the host C code (marker parser, stbi_* entry points,
 * batch loader) and the hand-written sm_100a CUDA kernels.  extern "C", plain pointers and sizes,
 * no C++ or torch types.  This is the boundary that replaces, inside the reference,
 *
 *   stbi__parse_entropy_coded_data   codec/jpeg.c:1155   (entropy decode of one scan)
 *   stbi__jpeg_finish                codec/jpeg.c:1326   (progressive dequant + IDCT)
 *   the row loop of load_jpeg_image  codec/jpeg.c:2301   (upsample + colour + pack)
 *
 * i.e. everything below stbi__decode_jpeg_image's marker loop.  The host keeps the marker
 * parser (jpeg.c:1349-1756) and hands the device: file bytes, raw DHT/DQT contents, per-scan
 * parameters and the byte offsets of restart segments.
 */
#ifndef JPEG_CUDA_ABI_H
#define JPEG_CUDA_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* colour/pack modes of the output stage (cases of load_jpeg_image, jpeg.c:2320-2431) */
enum
{
	JCU_CM_Y = 0,			/* replicate / copy component 0            jpeg.c:2373-2378, 2418-2430 */
	JCU_CM_YCBCR = 1,		/* YCbCr -> RGB                            jpeg.c:2338, 2369 */
	JCU_CM_RGB = 2,		/* components are R,G,B: pass through      jpeg.c:2325-2335 */
	JCU_CM_CMYK = 3,		/* CMYK -> RGB via blinn(c, k)             jpeg.c:2343-2353 */
	JCU_CM_YCCK = 4,		/* YCCK -> RGB -> blinn(255-c, k)          jpeg.c:2355-2366 */
	JCU_CM_RGB_LUMA = 5, /* compute_y(r,g,b)                        jpeg.c:2382-2394 */
	JCU_CM_CMYK_LUMA = 6, /* compute_y of CMYK->RGB                 jpeg.c:2396-2408 */
	JCU_CM_YCCK_LUMA = 7	 /* blinn(255 - Y, K)                      jpeg.c:2409-2417 */
};

/* scan kinds (the four block decoders of jpeg.c:308-558) */
enum
{
	JCU_SCAN_BASELINE = 0,	  /* stbi__jpeg_decode_block            jpeg.c:308 */
	JCU_SCAN_PROG_DC_FIRST = 1, /* ..._prog_dc, succ_high == 0         jpeg.c:382 */
	JCU_SCAN_PROG_DC_REFINE = 2, /* ..._prog_dc, succ_high != 0        jpeg.c:395 */
	JCU_SCAN_PROG_AC_FIRST = 3, /* ..._prog_ac, succ_high == 0         jpeg.c:412 */
	JCU_SCAN_PROG_AC_REFINE = 4 /* ..._prog_ac, succ_high != 0         jpeg.c:468 */
};

/* output-stage kernel selection (decided on the host from the frame header) */
enum
{
	JCU_PATH_GENERIC = 0, /* two-pass: IDCT to planes, then resample+colour (any H,V, any mode) */
	JCU_PATH_FUSED_420 = 1, /* fused tile kernel, 3 comps 2x2/1x1/1x1, YCbCr -> RGB/RGBA */
	JCU_PATH_FUSED_444 = 2, /* fused tile kernel, 3 comps 1x1, YCbCr -> RGB/RGBA */
	JCU_PATH_FUSED_GRAY = 3, /* fused tile kernel, component 0 only -> 1..4 channels */
	JCU_PATH_FUSED_422 = 4,	 /* fused strip kernel, 3 comps 2x1/1x1/1x1 (h_2 upsampling, jpeg.c:1784), YCbCr -> RGB/RGBA */
	JCU_PATH_COUNT = 5
};

/* raw DHT contents (jpeg.c:1381-1416): the device builds its own lookup tables from these */
typedef struct
{
	uint8_t bits[16];	 /* number of codes of length 1..16 */
	uint8_t vals[256]; /* symbols in code order */
} jcu_huff_spec;

typedef struct
{
	int32_t h, v;		 /* sampling factors */
	int32_t hs, vs;	 /* expansion factors hmax/h, vmax/v (jpeg.c:2273) */
	int32_t cx, cy;	 /* effective size in samples (jpeg.c:1627) */
	int32_t bw, bh;	 /* blocks incl. MCU padding = w2/8, h2/8 (jpeg.c:1636) */
	int32_t cbw, cbh;	 /* blocks covering cx, cy (non-interleaved scans, jpeg.c:1169) */
	int32_t w_lores;	 /* (img_x + hs - 1) / hs (jpeg.c:2276) */
	int32_t pad;
	uint64_t coef_off; /* first coefficient (int16 index) of this component in the coefficient slab */
	uint64_t plane_off; /* byte offset of this component's sample plane (generic path only) */
} jcu_comp;

typedef struct
{
	int32_t width, height;
	int32_t ncomp;		  /* components in the frame */
	int32_t decode_n;	  /* components the output stage needs (jpeg.c:2246) */
	int32_t n_out;		  /* output channels */
	int32_t color_mode; /* JCU_CM_* */
	int32_t path;		  /* JCU_PATH_* */
	int32_t hmax, vmax, mcu_x, mcu_y;
	int32_t progressive;
	int32_t ok;			  /* host verdict: 0 = do not decode (parse failed) */
	int32_t tile_first; /* first tile index of this image in its path's launch */
	int32_t tiles_x, tiles_y;
	jcu_comp comp[4];
	uint16_t qt[4][64]; /* per COMPONENT, natural order (jpeg.c:1376): applied by the entropy stage for baseline */
	uint16_t qt_out[4][64]; /* table the OUTPUT stage applies: = qt for progressive (jpeg.c:1341), all ones for baseline */
	uint64_t out_off;	  /* byte offset of the packed output in the pixel slab */
	uint64_t data_off;  /* byte offset of the file in the bitstream slab (16-byte aligned) */
	uint32_t data_len;
	uint32_t flip; /* 1 = store the rows bottom-up (stbi_set_flip_vertically_on_load, convert.c:97-101) */
} jcu_image;

typedef struct
{
	int32_t image;
	int32_t kind; /* JCU_SCAN_* */
	int32_t ncomp;
	int32_t comp[4];		/* frame component index, in SOS order (jpeg.c:1495) */
	int32_t dc_spec[4];	/* index into the jcu_huff_spec pool, -1 = undefined */
	int32_t ac_spec[4];
	int32_t ss, se, ah, al;
	int32_t restart_interval; /* in units; 0 = none */
	int32_t total_units;		  /* MCUs (interleaved) or blocks (single component) */
	int32_t units_per_row;
	int32_t wave;				  /* ordinal of the scan within its image; waves run in order */
	uint32_t seg_first;		  /* first entry in the segment-offset array */
	uint32_t nseg;
	uint32_t data_end;		  /* offset (in the file) of the marker that terminates the scan */
	uint32_t end_rst;			  /* 1 = that marker is an RSTn behind the last expected interval (the reference's restart check consumes it) */
} jcu_scan;

typedef struct
{
	int32_t n_images;
	const jcu_image *images;
	int32_t n_scans;
	const jcu_scan *scans;
	uint32_t n_segs;
	const uint32_t *seg_start; /* per segment: offset in the file of its first entropy-coded byte */
	int32_t n_huff;
	const jcu_huff_spec *huff;
	const uint8_t *bitstream; /* host slab: files concatenated at 16-byte aligned offsets */
	size_t bitstream_bytes;
	size_t coef_elems;  /* int16 elements in the coefficient slab */
	size_t plane_bytes; /* generic-path sample planes */
	size_t out_bytes;	  /* pixel slab */
	int32_t n_waves;
	int32_t need_coef_clear; /* 1 = some image is progressive / multi-scan: zero the coefficient slab */
} jcu_batch;

typedef struct
{
	float h2d_ms, entropy_ms, idct_ms, d2h_ms, total_ms;
	int32_t launches;
	int32_t ss_scans;	 /* scans decoded by the self-synchronising path in the last run */
	int32_t ss_rounds; /* refine rounds it needed */
} jcu_timings;

typedef struct jcu_ctx jcu_ctx;

#ifndef JCU_API
#define JCU_API __attribute__((visibility("default")))
#endif

JCU_API int jcu_device_count(void);
/* all functions below return 0 on success, else a CUDA error code; text via jcu_error_string */
JCU_API int jcu_ctx_create(int device, int profile, int entropy_mode, jcu_ctx **out);
/* same, with the tuning / test switches (all zero = the defaults of jcu_ctx_create) */
typedef struct
{
	int32_t device, profile, entropy_mode;
	int32_t force_generic; /* 1 = every image takes the generic two-pass output path (parity tests of that path) */
	int32_t rows_per_band; /* > 0 = MCU rows per band of the 4:2:0 strip kernel instead of the batch-size heuristic */
	int32_t guard;			  /* 1 = every device buffer carries a canary behind its logical size (jcu_debug_guards_ok) */
	int32_t selfsync_sub_bytes; /* 32 / 64 / 128 / 256 = bytes per subsequence of the self-synchronising decoder instead of the batch-size heuristic */
	int32_t host_threads; /* hint: host threads the caller gives this context (its parser pool); the copier threads of the pageable download ring are sized by it (0 = by the machine) */
} jcu_ctx_options;
JCU_API int jcu_ctx_create_opts(const jcu_ctx_options *o, jcu_ctx **out);
JCU_API void jcu_ctx_destroy(jcu_ctx *c);
JCU_API const char *jcu_error_string(jcu_ctx *c);

JCU_API int jcu_mem_info(int device, size_t *free_bytes, size_t *total_bytes); /* cudaMemGetInfo: the batch loader budgets its chunks with it */
JCU_API void *jcu_host_alloc(size_t bytes); /* pinned host memory */
JCU_API void jcu_host_free(void *p);

JCU_API int jcu_reserve(jcu_ctx *c, size_t bitstream_bytes, size_t coef_elems, size_t out_bytes); /* pre-size the device buffers */
JCU_API int jcu_upload(jcu_ctx *c, const jcu_batch *b);								  /* H2D, asynchronous on the ctx stream */
JCU_API int jcu_run(jcu_ctx *c);															  /* entropy waves + output stage, asynchronous */
JCU_API int jcu_sync(jcu_ctx *c);
/* Asynchronous D2H of out[off, off + bytes) on the context's download stream; complete after jcu_sync.  dst may be
   page-locked (the copy engine writes it directly) or pageable: 1 MB and more into pageable memory goes through the
   device's ring of pinned slots and copier threads (csrc/jpeg_cuda.cu, struct Bounce), which jcu_sync also drains.
   No counterpart in the reference (its pixels are produced in host memory, jpeg.c:2293). */
JCU_API int jcu_download(jcu_ctx *c, size_t off, size_t bytes, void *dst);
JCU_API void jcu_share_download_stream(jcu_ctx *c, jcu_ctx *owner);				  /* queue c's downloads on owner's download stream (FIFO over a chunk pipeline) */
/* per-image failure word: 0 = decoded; else JCU_ERR_CODE(word) says why (the earliest event in file order wins) */
#define JCU_ERR_BAD_HUFFMAN 1u	 /* "bad huffman code", jpeg.c:317, 351, 446, 505, 549 */
#define JCU_ERR_UNKNOWN_MARKER 2u /* a segment was decoded before its bytes were used up: the reference takes what follows
												for a marker segment, jpeg.c:1186, 1722-1735 -> "unknown marker" (:1469) */
#define JCU_ERR_CODE(word) ((~(uint32_t)(word)) & 3u)
JCU_API int jcu_image_errors(jcu_ctx *c, int32_t *errs, int n);					  /* per-image device error flags (after sync) */
JCU_API void *jcu_device_out(jcu_ctx *c);
JCU_API int jcu_get_timings(jcu_ctx *c, jcu_timings *t);

/* Widening epilogues: `count` samples of the pixel slab starting at byte `off` (16-byte aligned) become
   16-bit (elem_bytes 2: v * 257, stbi__convert_8_to_16 codec/convert.c:18-35) or float (elem_bytes 4: lut512[v] for
   the first `nonalpha` of every `comp` interleaved channels, lut512[256 + v] for the alpha channel, stbi__ldr_to_hdr
   codec/common.c:391-424) in the context's wide buffer; asynchronous on the ctx stream.  The pageable host table is
   read before the call returns. */
JCU_API int jcu_widen(jcu_ctx *c, size_t off, size_t count, int elem_bytes, int comp, int nonalpha, const float *lut512);
JCU_API int jcu_download_wide(jcu_ctx *c, size_t bytes, void *dst); /* async D2H of the wide buffer; jcu_sync waits */
JCU_API void *jcu_device_wide(jcu_ctx *c);

/* stage-level entry point: the IDCT stage alone (host pointers in, host pointers out).  The colour-conversion test hook
   lives in the separate test library (csrc/test_hooks.cu -> libimage_codecs_b200_testhooks.so), not in the product. */
JCU_API int jcu_test_idct_blocks(jcu_ctx *c, const int16_t *coef, const uint16_t *qt, int nblocks, uint8_t *out64);
/* test hook: a context created with jcu_ctx_options.guard carries a 4 KB canary behind the logical size of every
   device buffer; returns 1 if all canaries are intact, 0 if one was overwritten (jcu_error_string says which),
   -1 if guard mode is off */
JCU_API int jcu_debug_guards_ok(jcu_ctx *c);
JCU_API int jcu_read_coefs(jcu_ctx *c, size_t elem_off, size_t elems, int16_t *dst); /* coefficient slab after run */

/* Allocation-free variant of jcu_test_idct_blocks.  It has NO counterpart in the reference: it was added
   for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
   serving loops, harnesses that forbid it).  Same kernel and arithmetic (stbi__idct_block, codec/jpeg.c:615), but
   every pointer is a DEVICE pointer supplied by the caller -- d_coef: nblocks x 64 int16 in natural order, 16-byte
   aligned; d_qt: 64 uint16 multipliers, natural order; d_out64: nblocks x 64 samples, 4-byte aligned -- and the
   slab-layout copy of the blocks the kernel reads lives in a file-scope `static __device__` array of
   JCU_NOALLOC_MAX_BLOCKS blocks (4 MiB that every GPU which loads the library holds until exit) instead of a
   cudaMalloc'ed buffer.  The call makes c's device current, enqueues two kernels on `stream` (a cudaStream_t; NULL =
   the context's own stream) and returns without synchronising; it calls no cudaMalloc, cudaFree or allocating
   library routine.  More than JCU_NOALLOC_MAX_BLOCKS blocks -> cudaErrorInvalidValue, reason in jcu_error_string.
   All calls on one device share the scratch: calls that overlap on the GPU (different streams without an ordering
   between them) race; calls on one stream are safe, and each call rewrites every scratch byte it later reads, so
   nothing of the previous call survives into its result.  The first call on a device looks the scratch address up
   (cudaGetSymbolAddress) and should happen outside graph capture. */
#define JCU_NOALLOC_MAX_BLOCKS 32768
JCU_API int jcu_idct_blocks_noalloc(jcu_ctx *c, const int16_t *d_coef, const uint16_t *d_qt, int nblocks, uint8_t *d_out64,
												void *stream);

/* ---- encoder (replaces stbi_write_jpg_core, codec/jpeg_write.c:172-366) ---- */
#define JCU_ENC_HDR_STRIDE 1024 /* bytes reserved per image in the header staging array */
typedef struct
{
	int32_t width, height, comp; /* source pixels: comp interleaved bytes per pixel (1..4), rows top-down */
	int32_t subsample;			  /* 1 = 4:2:0 (16x16 MCUs, quality <= 90 in the reference), 0 = 4:4:4 */
	int32_t flip;					  /* read the rows bottom-up (stbi__flip_vertically_on_write) */
	int32_t mcu_x, mcu_y, nblocks;
	uint64_t pix_off;	  /* bytes into the pixel slab */
	uint64_t blk_first; /* index of the image's first block among all blocks of the batch */
	uint64_t out_off;	  /* bytes into the output slab */
	uint32_t out_cap;	  /* room at out_off */
	uint32_t hdr_len;	  /* bytes of this image's header in the staging array */
	float fdtbl_y[64], fdtbl_uv[64]; /* 1 / (quantiser * AAN scale), natural order (jpeg_write.c:227-234) */
} jcu_enc_image;

/* Encodes n images: pixels -> complete JPEG files in the encoder's device output slab.  `pixels` is a device
   pointer used in place (pixels_on_device; pix_off relative to it), else NULL and pix_ptrs[i] are the host
   pointers of the images, each copied to pix_off of the encoder's own pixel slab.  huff = code | length << 16 per
   [Y DC, UV DC, Y AC, UV AC][symbol].  out_len (host, n entries) receives each file's length, 0 = out_cap too
   small.  kernel_ms (optional) = device time of the kernels alone. */
JCU_API int jcu_encode(jcu_ctx *c, const jcu_enc_image *imgs, int n, const uint8_t *pixels, const uint8_t *const *pix_ptrs,
							  size_t pixel_bytes, int pixels_on_device, const uint8_t *headers, const uint32_t *huff, size_t out_bytes, uint32_t *out_len, float *kernel_ms);
JCU_API int jcu_encode_fetch(jcu_ctx *c, size_t off, size_t bytes, void *dst); /* asynchronous D2H from the output slab */
JCU_API int jcu_encode_sync(jcu_ctx *c);
JCU_API void *jcu_encode_device_out(jcu_ctx *c);

#ifdef __cplusplus
}
#endif
#endif
