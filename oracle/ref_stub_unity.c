/*
 * oracle/ref_stub_unity.c -- TEST INFRASTRUCTURE ONLY: the reference-side binding of INTEGRATION.md section 2, compiled.
 *
 * The reference's own public layer and dispatch (common.c, image_api.c, convert.c, #included UNMODIFIED from
 * REF_ROOT) are built around a codec/jpeg.c whose stbi__jpeg_load (jpeg.c:2442-2452) is replaced by the stub a
 * maintainer would add: stbi_load_from_memory (convert.c:254) -> stbi__load_and_postprocess_8bit (convert.c:78)
 * -> stbi__load_main (image_api.c:3-13) -> stbi__jpeg_test (jpeg.c:2454, the reference's) -> the stub below ->
 * stbi_jpeg_decode_batch of libimage_codecs_b200.so.  The reference's CPU decoder is still compiled (its
 * stbi__jpeg_test / stbi__jpeg_info are used) but its load entry is renamed away and never called.
 *
 * Public types come from THIS repository's include/image_api.h -- the header the reference ships empty -- so the
 * stub is compiled exactly as INTEGRATION.md prints it; ref_prelude.h then supplies only the internals.
 */
#define STBI_NO_PNG
#define STBI_NO_BMP
#define STBI_NO_GIF
#define STBI_NO_PSD
#define STBI_NO_PIC
#define STBI_NO_PNM
#define STBI_NO_HDR
#define STBI_NO_TGA
#define STBI_NO_LINEAR
#define STBI_NO_STDIO
#define STBI_NO_ZLIB
#define STBI_WRITE_NO_STDIO

#define ORACLE_PUBLIC_TYPES_FROM_PRODUCT_HEADER
#include "../include/image_api.h" /* stbi_uc, stbi_io_callbacks, STBI_rgb..., stbi_jpeg_batch_* */
#include "ref_prelude.h"

#define ORACLE_STR2(x) #x
#define ORACLE_STR(x) ORACLE_STR2(x)
#define ORACLE_REF(file) ORACLE_STR(REF_ROOT/file)

#include ORACLE_REF(common.c)

/* the reference's CPU entry keeps its body under another name; the dispatch below reaches the stub instead */
#define stbi__jpeg_load stbi__jpeg_load_reference_cpu
#include ORACLE_REF(codec/jpeg.c)
#undef stbi__jpeg_load

/* ---- BEGIN the stub of INTEGRATION.md section 2, verbatim ---- */
static stbi_jpeg_batch *stbi__b200;           /* one per process; create once */

static void *stbi__jpeg_load(stbi__context *s, int *x, int *y, int *comp, int req_comp, stbi__result_info *ri)
{
   /* the GPU path needs the whole stream: memory contexts already have it, callback/file contexts are
      drained first (what stbi_load_from_file does here: csrc/image_api.c drain_file / drain_callbacks) */
   const stbi_uc *buf = s->img_buffer_original;
   int len = (int)(s->img_buffer_original_end - s->img_buffer_original);
   int status = 0, n;
   stbi_uc *out;
   size_t cap;
   STBI_NOTUSED(ri);
   if (!stbi__b200) {
      stbi_jpeg_batch_params p = {0};          /* device 0, auto threads */
      stbi__b200 = stbi_jpeg_batch_create(&p);
      if (!stbi__b200) return stbi__errpuc("no cuda device", "B200 path unavailable");
   }
   /* sizing pass (header parse only), then decode into a malloc'ed buffer the caller will free */
   if (stbi_jpeg_decode_batch(stbi__b200, &buf, &len, 1, req_comp, NULL, NULL, x, y, comp, &status) < 0 || !status)
      return stbi__errpuc(stbi_jpeg_batch_reason(stbi__b200, 0), "Corrupt JPEG");
   n = req_comp ? req_comp : *comp;
   cap = (size_t)n * *x * *y;
   out = (stbi_uc *)stbi__malloc(cap + 1);    /* +1: the reference's slack byte, jpeg.c:2293 */
   if (!out) return stbi__errpuc("outofmem", "Out of memory");
   if (stbi_jpeg_decode_batch(stbi__b200, &buf, &len, 1, req_comp, &out, &cap, x, y, comp, &status) != 1) {
      STBI_FREE(out);
      return stbi__errpuc(stbi_jpeg_batch_reason(stbi__b200, 0), "Corrupt JPEG");
   }
   return out;
}
/* ---- END of the stub ---- */

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

/* what the tests call: the REFERENCE's stbi_load_from_memory (convert.c:254) and friends of this library, under names
   that cannot be confused with the product's exports of the same name */
#define STUB_API __attribute__((visibility("default")))
STUB_API unsigned char *refstub_load_from_memory(const unsigned char *buf, int len, int *x, int *y, int *comp, int req_comp)
{
	return stbi_load_from_memory(buf, len, x, y, comp, req_comp);
}
STUB_API int refstub_info_from_memory(const unsigned char *buf, int len, int *x, int *y, int *comp)
{
	return stbi_info_from_memory(buf, len, x, y, comp);
}
STUB_API const char *refstub_failure_reason(void) { return stbi__g_failure_reason; }
STUB_API void refstub_set_flip(int flag) { stbi__vertically_flip_on_load = flag; }
STUB_API void refstub_free(void *p) { free(p); }
