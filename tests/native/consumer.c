/*
 * tests/native/consumer.c -- a plain-C application of include/image_api.h, linked against libimage_codecs_b200.so the way
 * INTEGRATION.md section 1 describes: the same calls a user of the reference makes (stbi_info_from_memory
 * image_api.c:119, stbi_load_from_memory convert.c:254, stbi_failure_reason, stbi_image_free), then the batched and
 * the multi-device entry points on the same bytes.  Prints one line per call with an FNV-1a 64 hash of the pixels;
 * the GPU test compares the hashes with the oracle's decode.
 *
 *   consumer FILE.jpg [req_comp]     decode on the GPU
 *   consumer --info FILE.jpg         header only (no GPU)
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/image_api.h"

static unsigned long long fnv1a64(const unsigned char *p, size_t n)
{
	unsigned long long h = 0xcbf29ce484222325ull;
	size_t i;
	for (i = 0; i < n; ++i)
	{
		h ^= p[i];
		h *= 0x100000001b3ull;
	}
	return h;
}

static unsigned char *slurp(const char *path, int *len)
{
	FILE *f = fopen(path, "rb");
	unsigned char *buf;
	long n;
	if (!f)
		return NULL;
	fseek(f, 0, SEEK_END);
	n = ftell(f);
	fseek(f, 0, SEEK_SET);
	buf = (unsigned char *)malloc((size_t)n + 1);
	if (buf && fread(buf, 1, (size_t)n, f) != (size_t)n)
	{
		free(buf);
		buf = NULL;
	}
	fclose(f);
	*len = (int)n;
	return buf;
}

int main(int argc, char **argv)
{
	int info_only = argc > 1 && !strcmp(argv[1], "--info");
	const char *path = argc > 1 + info_only ? argv[1 + info_only] : NULL;
	int req_comp = argc > 2 && !info_only ? atoi(argv[2]) : 3;
	int len = 0, x = 0, y = 0, comp = 0, n;
	unsigned char *file, *px;
	if (!path)
	{
		fprintf(stderr, "usage: consumer [--info] FILE.jpg [req_comp]\n");
		return 2;
	}
	file = slurp(path, &len);
	if (!file)
	{
		fprintf(stderr, "cannot read %s\n", path);
		return 2;
	}
	if (!stbi_info_from_memory(file, len, &x, &y, &comp))
	{
		printf("info failed: %s\n", stbi_failure_reason());
		return 1;
	}
	printf("info %d %d %d\n", x, y, comp);
	if (info_only)
		return 0;
	n = req_comp ? req_comp : comp;

	/* the reference's per-image call */
	px = stbi_load_from_memory(file, len, &x, &y, &comp, req_comp);
	if (!px)
	{
		printf("load failed: %s\n", stbi_failure_reason());
		return 1;
	}
	printf("load %d %d %d %016llx\n", x, y, comp, fnv1a64(px, (size_t)x * y * n));
	stbi_image_free(px);

	/* the batched call: the same stream three times, caller-owned output buffers */
	{
		stbi_jpeg_batch_params p;
		stbi_jpeg_batch *b;
		const stbi_uc *bufs[3];
		int lens[3], xs[3], ys[3], cs[3], st[3], i, ok;
		stbi_uc *outs[3];
		size_t caps[3];
		memset(&p, 0, sizeof(p));
		b = stbi_jpeg_batch_create(&p);
		if (!b)
		{
			printf("batch_create failed: %s\n", stbi_failure_reason());
			return 1;
		}
		for (i = 0; i < 3; ++i)
		{
			bufs[i] = file;
			lens[i] = len;
			caps[i] = (size_t)x * y * n;
			outs[i] = (stbi_uc *)malloc(caps[i]);
		}
		ok = stbi_jpeg_decode_batch(b, bufs, lens, 3, req_comp, outs, caps, xs, ys, cs, st);
		printf("batch %d %016llx %016llx %016llx\n", ok, fnv1a64(outs[0], caps[0]), fnv1a64(outs[1], caps[1]), fnv1a64(outs[2], caps[2]));
		stbi_jpeg_batch_destroy(b);

		/* every visible device behind one call */
		{
			stbi_jpeg_multi *m = stbi_jpeg_multi_create(NULL, 0, NULL);
			if (!m)
			{
				printf("multi_create failed: %s\n", stbi_failure_reason());
				return 1;
			}
			for (i = 0; i < 3; ++i)
				memset(outs[i], 0, caps[i]);
			ok = stbi_jpeg_decode_batch_multi(m, bufs, lens, 3, req_comp, outs, caps, xs, ys, cs, st);
			printf("multi %d devices %d %016llx %016llx %016llx\n", stbi_jpeg_multi_device_count(m), ok, fnv1a64(outs[0], caps[0]),
					 fnv1a64(outs[1], caps[1]), fnv1a64(outs[2], caps[2]));
			stbi_jpeg_multi_destroy(m);
		}
		for (i = 0; i < 3; ++i)
			free(outs[i]);
	}
	free(file);
	return 0;
}
