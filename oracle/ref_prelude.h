/*
 * oracle/ref_prelude.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * The reference's image_api.h is a 0-byte file, so none of its .c files
 * compile as shipped.  This header supplies exactly the names those files use but never
 * define, so that the reference sources can be #included UNMODIFIED
 * from /root/reference by oracle/ref_unity.c.  It contains no JPEG arithmetic: every
 * operation on the decode path comes from the reference files themselves.
 *
 * Each definition below is derived from how the reference uses the name:
 *   stbi__context fields      common.c:10-28, convert.c:208, jpeg.c:1559-1572
 *   stbi_io_callbacks         common.c:12,49,79
 *   stbi__err / errpuc        jpeg.c:110, image_api.c:55, convert.c:193
 *   stbi__malloc_mad*         jpeg.c:1641,1651,2293
 *   stbi__start_mem/...       convert.c:203,243,257
 *   writer context            write_bmp.c:62-65, jpeg_write.c:249,292,370
 */
#ifndef ORACLE_REF_PRELUDE_H
#define ORACLE_REF_PRELUDE_H

#include <assert.h>
#include <limits.h>
#include <math.h>
#include <stdarg.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef ORACLE_USE_SSE2
#include <emmintrin.h>
#define STBI_SSE2
#define STBI_SIMD_ALIGN(type, name) type name __attribute__((aligned(16)))
static int stbi__sse2_available(void) { return 1; }
#else
#define STBI_SIMD_ALIGN(type, name) type name
#endif

#ifndef ORACLE_PUBLIC_TYPES_FROM_PRODUCT_HEADER /* ref_stub_unity.c takes the public types from include/image_api.h */
typedef unsigned char stbi_uc;
typedef unsigned short stbi_us;
#endif
typedef uint16_t stbi__uint16;
typedef int16_t stbi__int16;
typedef uint32_t stbi__uint32;
typedef int32_t stbi__int32;

#ifdef ORACLE_PUBLIC_TYPES_FROM_PRODUCT_HEADER
#undef STBIDEF
#define STBIDEF /* the reference's own definitions of stbi_load* stay inside the library (-fvisibility=hidden) */
#else
#define STBIDEF extern __attribute__((visibility("default")))
#endif
#define STBIWDEF extern __attribute__((visibility("default")))
#define stbi_inline inline
#define STBI_ASSERT(x) assert(x)
#define STBIW_ASSERT(x) assert(x)
#define STBI_FREE(p) free(p)
#define STBI_MALLOC(sz) malloc(sz)
#define STBI_REALLOC_SIZED(p, oldsz, newsz) realloc(p, newsz)
#define STBI_NOTUSED(v) (void)sizeof(v)
#define STBI_MAX_DIMENSIONS (1 << 24)
#define stbi_lrot(x, y) (((x) << (y)) | ((x) >> (-(y)&31)))
#define STBIW_UCHAR(x) (unsigned char)((x)&0xff)

#ifndef ORACLE_PUBLIC_TYPES_FROM_PRODUCT_HEADER
enum { STBI_default = 0, STBI_grey = 1, STBI_grey_alpha = 2, STBI_rgb = 3, STBI_rgb_alpha = 4 };
#endif
enum { STBI_ORDER_RGB, STBI_ORDER_BGR };

#ifndef ORACLE_PUBLIC_TYPES_FROM_PRODUCT_HEADER
typedef struct
{
	int (*read)(void *user, char *data, int size);
	void (*skip)(void *user, int n);
	int (*eof)(void *user);
} stbi_io_callbacks;
#endif

typedef struct
{
	stbi__uint32 img_x, img_y;
	int img_n, img_out_n;

	stbi_io_callbacks io;
	void *io_user_data;

	int read_from_callbacks;
	int buflen;
	stbi_uc buffer_start[128];
	int callback_already_read;

	stbi_uc *img_buffer, *img_buffer_end;
	stbi_uc *img_buffer_original, *img_buffer_original_end;
} stbi__context;

typedef struct
{
	int bits_per_channel;
	int num_channels;
	int channel_order;
} stbi__result_info;

/* ---- failure reason (short strings, as upstream's default) ---- */
static __thread const char *stbi__g_failure_reason;
STBIDEF const char *stbi_failure_reason(void) { return stbi__g_failure_reason; }
static int stbi__err_set(const char *str)
{
	stbi__g_failure_reason = str;
	return 0;
}
#define stbi__err(x, y) stbi__err_set(x)
#define stbi__errpf(x, y) ((float *)(size_t)(stbi__err(x, y) ? NULL : NULL))
#define stbi__errpuc(x, y) ((unsigned char *)(size_t)(stbi__err(x, y) ? NULL : NULL))

STBIDEF void stbi_image_free(void *p) { free(p); }

/* ---- overflow-checked allocation helpers ---- */
static void *stbi__malloc(size_t size) { return malloc(size); }
static int stbi__addsizes_valid(int a, int b)
{
	if (b < 0)
		return 0;
	return a <= INT_MAX - b;
}
static int stbi__mul2sizes_valid(int a, int b)
{
	if (a < 0 || b < 0)
		return 0;
	if (b == 0)
		return 1;
	return a <= INT_MAX / b;
}
static int stbi__mad2sizes_valid(int a, int b, int add)
{
	return stbi__mul2sizes_valid(a, b) && stbi__addsizes_valid(a * b, add);
}
static int stbi__mad3sizes_valid(int a, int b, int c, int add)
{
	return stbi__mul2sizes_valid(a, b) && stbi__mul2sizes_valid(a * b, c) && stbi__addsizes_valid(a * b * c, add);
}
static int stbi__mad4sizes_valid(int a, int b, int c, int d, int add)
{
	return stbi__mul2sizes_valid(a, b) && stbi__mul2sizes_valid(a * b, c) && stbi__mul2sizes_valid(a * b * c, d) &&
			 stbi__addsizes_valid(a * b * c * d, add);
}
static void *stbi__malloc_mad2(int a, int b, int add)
{
	if (!stbi__mad2sizes_valid(a, b, add))
		return NULL;
	return stbi__malloc(a * b + add);
}
static void *stbi__malloc_mad3(int a, int b, int c, int add)
{
	if (!stbi__mad3sizes_valid(a, b, c, add))
		return NULL;
	return stbi__malloc(a * b * c + add);
}
static void *stbi__malloc_mad4(int a, int b, int c, int d, int add)
{
	if (!stbi__mad4sizes_valid(a, b, c, d, add))
		return NULL;
	return stbi__malloc(a * b * c * d + add);
}

/* ---- context start / rewind ---- */
static void stbi__refill_buffer(stbi__context *s); /* defined in the reference's common.c */

static void stbi__start_mem(stbi__context *s, stbi_uc const *buffer, int len)
{
	s->io.read = NULL;
	s->read_from_callbacks = 0;
	s->callback_already_read = 0;
	s->img_buffer = s->img_buffer_original = (stbi_uc *)buffer;
	s->img_buffer_end = s->img_buffer_original_end = (stbi_uc *)buffer + len;
}
static void stbi__start_callbacks(stbi__context *s, stbi_io_callbacks *c, void *user)
{
	s->io = *c;
	s->io_user_data = user;
	s->buflen = sizeof(s->buffer_start);
	s->read_from_callbacks = 1;
	s->callback_already_read = 0;
	s->img_buffer = s->img_buffer_original = s->buffer_start;
	stbi__refill_buffer(s);
	s->img_buffer_original_end = s->img_buffer_end;
}
static void stbi__rewind(stbi__context *s)
{
	s->img_buffer = s->img_buffer_original;
	s->img_buffer_end = s->img_buffer_original_end;
}

static int stbi__vertically_flip_on_load = 0;

/* ---- writer side (only what jpeg_write.c + stbiw__putc need) ---- */
typedef void stbi_write_func(void *context, void *data, int size);
typedef struct
{
	stbi_write_func *func;
	void *context;
	unsigned char buffer[64];
	int buf_used;
} stbi__write_context;
static void stbi__start_write_callbacks(stbi__write_context *s, stbi_write_func *c, void *context)
{
	s->func = c;
	s->context = context;
}
static int stbi__flip_vertically_on_write = 0;


/* LDR->HDR widening (common.c:391, only with ORACLE_LINEAR): the reference defines these in convert.c:402 AFTER
   common.c uses them, so its missing header must have carried tentative definitions like these */
#ifndef STBI_NO_LINEAR
static float stbi__l2h_gamma, stbi__l2h_scale;
#endif
#endif
