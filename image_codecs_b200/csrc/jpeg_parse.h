/*
 * jpeg_parse.h -- host-side JPEG marker/header parser (C).
 *
 * Replaces the host half of the reference decoder: stbi__decode_jpeg_header (jpeg.c:1670),
 * stbi__process_marker (:1349), stbi__process_frame_header (:1549), stbi__process_scan_header
 * (:1471) and the marker loop of stbi__decode_jpeg_image (:1702).  It performs NO entropy
 * decoding: for each scan it walks the entropy-coded bytes only to find marker positions and
 * emits the byte offsets of restart segments for the GPU (BASELINE.json north_star: "the
 * jpeg.c marker/header parser stays on the host, emitting segment offsets").
 */
#ifndef JPEG_PARSE_H
#define JPEG_PARSE_H

#include "../../include/jpeg_cuda_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

enum
{
	JP_MODE_LOAD = 0,	 /* full parse (STBI__SCAN_load) */
	JP_MODE_TYPE = 1,	 /* SOI only (STBI__SCAN_type; stbi__jpeg_test, jpeg.c:2454) */
	JP_MODE_HEADER = 2 /* up to and including SOF (STBI__SCAN_header; stbi__jpeg_info_raw, jpeg.c:2466) */
};

typedef struct
{
	int ok;				  /* 1 = parsed, image can be decoded */
	const char *reason; /* reference's short failure string when !ok (may be NULL: reference left it stale) */

	jcu_image img; /* geometry, colour mode, quant tables; slab offsets are filled by the batch builder */
	int comps_in_file; /* what *comp reports: img_n >= 3 ? 3 : 1 (jpeg.c:2437) */

	jcu_scan *scans; /* dc_spec/ac_spec index THIS result's huff[] until rebased by the builder */
	int nscans, cap_scans;
	uint32_t *segs; /* segment start offsets (relative to the file) for all scans, scan.seg_first indexes here */
	int nsegs, cap_segs;
	jcu_huff_spec *huff;
	int nhuff, cap_huff;

	size_t consumed; /* bytes of the file the reference has read when it returns: up to and including the EOI marker */
	double entropy_bytes; /* entropy-coded bytes over all scans (incl. stuffing, excl. markers) */
	long long coef_blocks; /* blocks the output stage reads (MCU-padded, decode_n components) */
	int needs_clear; /* some block of the image is written by no scan (lost RSTn, component without a scan): zero the slab first */
} jp_result;

/* Parse one JPEG held in memory.  Returns r->ok.  r must be zero-initialised or previously freed. */
int jp_parse(const unsigned char *buf, size_t len, int req_comp, int mode, jp_result *r);
void jp_result_free(jp_result *r);

#ifdef __cplusplus
}
#endif
#endif
