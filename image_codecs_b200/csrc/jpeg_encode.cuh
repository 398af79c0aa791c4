/*
 * jpeg_encode.cuh -- baseline JPEG encoder on the device (sm_100a), byte-identical to the reference's
 * stbi_write_jpg_core (codec/jpeg_write.c:172-366).  SURVEY 8(f) row N3: the step on the other side of the
 * wire format, so that batches of test/benchmark inputs can be produced on the device.
 *
 * The reference encodes strictly serially: one bit buffer for the whole image, DC predictors carried from
 * block to block (jpeg_write.c:303-357).  Here:
 *
 *   k_enc_dct    one thread = one 8x8 block.  RGB -> YCbCr in float (jpeg_write.c:321-323), 2x2 chroma average
 *                for subsampled files (:337-338), the float AAN DCT (stbiw__jpg_DCT, :24-74; rows then
 *                columns, :95-105), quantise/round/zig-zag (:107-118).  Every float operation is a separate
 *                IEEE round-to-nearest instruction (__fmul_rn / __fadd_rn / __fsub_rn): the reference is
 *                compiled without FMA contraction on x86-64, and a fused multiply-add would round differently.
 *                Output: 64 int16 in zig-zag order per block.
 *   k_enc_count  one thread = one block: DC difference against the previous block of the same component in
 *                encode order, then the exact number of bits stbiw__jpg_processDU (:120-168) emits.
 *   k_enc_scan   per image: exclusive prefix sum of the bit counts (and, later, of the 0xFF counts).
 *   k_enc_emit   one thread = one block: the same symbol walk again, writing the bits at the block's bit
 *                offset into a zeroed buffer; words wholly inside the block are plain stores, the two boundary
 *                words are OR-ed atomically.
 *   k_enc_ffcount / k_enc_stuff   byte stuffing (stbiw__jpg_writeBits, :4-22: 0xFF is followed by 0x00) as a
 *                count / prefix sum / scatter over 64-byte chunks, including the final 7 one-bits of padding
 *                (:360) and the EOI marker.
 *
 * Headers (quantisation tables, SOF0, the four standard Huffman tables, SOS) are assembled on the host
 * (image_api.c) and uploaded with the image descriptors.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"

#define ENC_THREADS 128
#define ENC_BLOCK_BYTES 256 /* room for the bits of one block: 64 x (16 + 12) bits = 224 bytes at most */
#define ENC_CHUNK 64			 /* bytes per stuffing chunk */

/* code | length << 16, indexed [table][symbol]; tables: 0 = Y DC, 1 = UV DC, 2 = Y AC, 3 = UV AC */
__constant__ uint32_t c_enc_huff[4][256];
__constant__ uint8_t c_enc_zigzag[64] = {0,	1,	 5,  6,	14, 15, 27, 28, 2,  4,	7,	 13, 16, 26, 29, 42, 3,	 8,  12, 17, 25, 30,
													  41, 43, 9,  11, 18, 24, 31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38,
													  46, 51, 55, 60, 21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63};

#define EF_A(a, b) __fadd_rn((a), (b))
#define EF_S(a, b) __fsub_rn((a), (b))
#define EF_M(a, b) __fmul_rn((a), (b))

/* stbiw__jpg_DCT (jpeg_write.c:24-74), operation for operation */
__device__ __forceinline__ void enc_dct8(float &d0, float &d1, float &d2, float &d3, float &d4, float &d5, float &d6, float &d7)
{
	const float tmp0 = EF_A(d0, d7), tmp7 = EF_S(d0, d7), tmp1 = EF_A(d1, d6), tmp6 = EF_S(d1, d6);
	const float tmp2 = EF_A(d2, d5), tmp5 = EF_S(d2, d5), tmp3 = EF_A(d3, d4), tmp4 = EF_S(d3, d4);
	float tmp10 = EF_A(tmp0, tmp3);
	const float tmp13 = EF_S(tmp0, tmp3);
	float tmp11 = EF_A(tmp1, tmp2), tmp12 = EF_S(tmp1, tmp2);
	const float e0 = EF_A(tmp10, tmp11), e4 = EF_S(tmp10, tmp11);
	const float z1 = EF_M(EF_A(tmp12, tmp13), 0.707106781f);
	const float e2 = EF_A(tmp13, z1), e6 = EF_S(tmp13, z1);
	tmp10 = EF_A(tmp4, tmp5);
	tmp11 = EF_A(tmp5, tmp6);
	tmp12 = EF_A(tmp6, tmp7);
	const float z5 = EF_M(EF_S(tmp10, tmp12), 0.382683433f);
	const float z2 = EF_A(EF_M(tmp10, 0.541196100f), z5);
	const float z4 = EF_A(EF_M(tmp12, 1.306562965f), z5);
	const float z3 = EF_M(tmp11, 0.707106781f);
	const float z11 = EF_A(tmp7, z3), z13 = EF_S(tmp7, z3);
	d5 = EF_A(z13, z2);
	d3 = EF_S(z13, z2);
	d1 = EF_A(z11, z4);
	d7 = EF_S(z11, z4);
	d0 = e0;
	d2 = e2;
	d4 = e4;
	d6 = e6;
}

/* component value of one source pixel (jpeg_write.c:318-323); which: 0 = Y, 1 = U, 2 = V */
__device__ __forceinline__ float enc_sample(const jcu_enc_image &im, const uint8_t *__restrict__ pix, int row, int col, int which)
{
	const int cr = row < im.height ? row : im.height - 1, cc = col < im.width ? col : im.width - 1;
	const size_t p = ((size_t)(im.flip ? im.height - 1 - cr : cr) * im.width + cc) * im.comp;
	const int ofs_g = im.comp > 2 ? 1 : 0, ofs_b = im.comp > 2 ? 2 : 0;
	const float r = pix[p], g = pix[p + ofs_g], b = pix[p + ofs_b];
	if (which == 0)
		return EF_S(EF_A(EF_A(EF_M(0.29900f, r), EF_M(0.58700f, g)), EF_M(0.11400f, b)), 128.0f);
	if (which == 1)
		return EF_A(EF_S(EF_M(-0.16874f, r), EF_M(0.33126f, g)), EF_M(0.50000f, b));
	return EF_S(EF_S(EF_M(0.50000f, r), EF_M(0.41869f, g)), EF_M(0.08131f, b));
}

/* block b of an image, in encode order: component (0 Y, 1 U, 2 V), pixel origin, and whether it is a 2x2-averaged block */
__device__ __forceinline__ void enc_locate(const jcu_enc_image &im, uint32_t b, int &which, int &x0, int &y0, bool &sub)
{
	const uint32_t per = im.subsample ? 6u : 3u;
	const uint32_t m = b / per, j = b - m * per;
	const int mx = (int)(m % (uint32_t)im.mcu_x), my = (int)(m / (uint32_t)im.mcu_x);
	sub = false;
	if (im.subsample)
	{
		if (j < 4u)
		{
			which = 0;
			x0 = mx * 16 + (int)(j & 1u) * 8;
			y0 = my * 16 + (int)(j >> 1) * 8;
		}
		else
		{
			which = (int)j - 3;
			x0 = mx * 16;
			y0 = my * 16;
			sub = true;
		}
	}
	else
	{
		which = (int)j;
		x0 = mx * 8;
		y0 = my * 8;
	}
}

/* grid = (ceil(max blocks / ENC_THREADS), images) */
__global__ void __launch_bounds__(ENC_THREADS)
k_enc_dct(const jcu_enc_image *__restrict__ images, const uint8_t *__restrict__ pixels, int16_t *__restrict__ coef)
{
	const jcu_enc_image &im = images[blockIdx.y];
	const uint32_t b = blockIdx.x * ENC_THREADS + threadIdx.x;
	if (b >= (uint32_t)im.nblocks || im.subsample)
		return; /* 4:2:0 images go to k_enc_dct420 */
	int which, x0, y0;
	bool sub;
	enc_locate(im, b, which, x0, y0, sub);
	const uint8_t *pix = pixels + im.pix_off;
	float v[64];
#pragma unroll
	for (int y = 0; y < 8; ++y)
#pragma unroll
		for (int x = 0; x < 8; ++x)
		{
			if (sub)
			{
				/* (U[j] + U[j+1] + U[j+16] + U[j+17]) * 0.25f (jpeg_write.c:337) */
				const float a = enc_sample(im, pix, y0 + 2 * y, x0 + 2 * x, which), c = enc_sample(im, pix, y0 + 2 * y, x0 + 2 * x + 1, which);
				const float d = enc_sample(im, pix, y0 + 2 * y + 1, x0 + 2 * x, which), e = enc_sample(im, pix, y0 + 2 * y + 1, x0 + 2 * x + 1, which);
				v[y * 8 + x] = EF_M(EF_A(EF_A(EF_A(a, c), d), e), 0.25f);
			}
			else
				v[y * 8 + x] = enc_sample(im, pix, y0 + y, x0 + x, which);
		}
#pragma unroll
	for (int y = 0; y < 8; ++y)
		enc_dct8(v[y * 8], v[y * 8 + 1], v[y * 8 + 2], v[y * 8 + 3], v[y * 8 + 4], v[y * 8 + 5], v[y * 8 + 6], v[y * 8 + 7]);
#pragma unroll
	for (int x = 0; x < 8; ++x)
		enc_dct8(v[x], v[x + 8], v[x + 16], v[x + 24], v[x + 32], v[x + 40], v[x + 48], v[x + 56]);
	const float *fd = which == 0 ? im.fdtbl_y : im.fdtbl_uv;
	int16_t *out = coef + ((size_t)im.blk_first + b) * 64;
#pragma unroll
	for (int j = 0; j < 64; ++j)
	{
		const float q = EF_M(v[j], fd[j]);
		out[c_enc_zigzag[j]] = (int16_t)(int)(q < 0 ? EF_S(q, 0.5f) : EF_A(q, 0.5f)); /* jpeg_write.c:117 */
	}
}

/*
 * 4:2:0 files (the reference's choice at quality <= 90): the same arithmetic on a shared-memory tile.  A CTA takes
 * ENC_TILE_MCUS consecutive MCUs of one MCU row (16 x 256 pixels).  Phase 1: one 2x2 pixel quad per thread step --
 * neighbouring threads read neighbouring pixels -- converts each source pixel ONCE into Y (four samples) and the
 * averaged U and V sample (jpeg_write.c:321-323, 337-338).  Phase 2: one thread = one block, in encode order,
 * DCT + quantise from the tile.  The generic kernel reads every pixel three times with strided byte loads.
 */
#define ENC_TILE_MCUS 16
#define ENC_TILE_W (16 * ENC_TILE_MCUS)
__global__ void __launch_bounds__(ENC_THREADS)
k_enc_dct420(const jcu_enc_image *__restrict__ images, const uint8_t *__restrict__ pixels, int16_t *__restrict__ coef)
{
	__shared__ float sY[16][ENC_TILE_W + 4];
	__shared__ float sU[8][ENC_TILE_W / 2 + 2], sV[8][ENC_TILE_W / 2 + 2];
	const jcu_enc_image &im = images[blockIdx.y];
	if (!im.subsample)
		return;
	const int tiles_x = (im.mcu_x + ENC_TILE_MCUS - 1) / ENC_TILE_MCUS;
	if ((int)blockIdx.x >= tiles_x * im.mcu_y)
		return;
	const int ty = blockIdx.x / tiles_x, tx = blockIdx.x - ty * tiles_x;
	const int m0 = tx * ENC_TILE_MCUS, nm = min(ENC_TILE_MCUS, im.mcu_x - m0);
	const int x0 = m0 * 16, y0 = ty * 16;
	const uint8_t *pix = pixels + im.pix_off;
	const int ofs_g = im.comp > 2 ? 1 : 0, ofs_b = im.comp > 2 ? 2 : 0;
	for (int q = threadIdx.x; q < 8 * nm * 8; q += ENC_THREADS)
	{
		const int qy = q / (nm * 8), qx = q - qy * (nm * 8);
		float u[4], v[4];
#pragma unroll
		for (int i = 0; i < 4; ++i)
		{
			const int row = y0 + 2 * qy + (i >> 1), col = x0 + 2 * qx + (i & 1);
			const int cr = row < im.height ? row : im.height - 1, cc = col < im.width ? col : im.width - 1;
			const size_t p = ((size_t)(im.flip ? im.height - 1 - cr : cr) * im.width + cc) * im.comp;
			const float r = pix[p], g = pix[p + ofs_g], b = pix[p + ofs_b];
			sY[2 * qy + (i >> 1)][2 * qx + (i & 1)] = EF_S(EF_A(EF_A(EF_M(0.29900f, r), EF_M(0.58700f, g)), EF_M(0.11400f, b)), 128.0f);
			u[i] = EF_A(EF_S(EF_M(-0.16874f, r), EF_M(0.33126f, g)), EF_M(0.50000f, b));
			v[i] = EF_S(EF_S(EF_M(0.50000f, r), EF_M(0.41869f, g)), EF_M(0.08131f, b));
		}
		sU[qy][qx] = EF_M(EF_A(EF_A(EF_A(u[0], u[1]), u[2]), u[3]), 0.25f);
		sV[qy][qx] = EF_M(EF_A(EF_A(EF_A(v[0], v[1]), v[2]), v[3]), 0.25f);
	}
	__syncthreads();
	const int t = threadIdx.x;
	if (t >= nm * 6)
		return;
	const int m = t / 6, j = t - m * 6;
	float d[64];
	if (j < 4)
	{
		const int r0 = (j >> 1) * 8, c0 = m * 16 + (j & 1) * 8;
#pragma unroll
		for (int y = 0; y < 8; ++y)
#pragma unroll
			for (int x = 0; x < 8; ++x)
				d[y * 8 + x] = sY[r0 + y][c0 + x];
	}
	else
	{
		const float(*sC)[ENC_TILE_W / 2 + 2] = j == 4 ? sU : sV;
#pragma unroll
		for (int y = 0; y < 8; ++y)
#pragma unroll
			for (int x = 0; x < 8; ++x)
				d[y * 8 + x] = sC[y][m * 8 + x];
	}
#pragma unroll
	for (int y = 0; y < 8; ++y)
		enc_dct8(d[y * 8], d[y * 8 + 1], d[y * 8 + 2], d[y * 8 + 3], d[y * 8 + 4], d[y * 8 + 5], d[y * 8 + 6], d[y * 8 + 7]);
#pragma unroll
	for (int x = 0; x < 8; ++x)
		enc_dct8(d[x], d[x + 8], d[x + 16], d[x + 24], d[x + 32], d[x + 40], d[x + 48], d[x + 56]);
	const float *fd = j < 4 ? im.fdtbl_y : im.fdtbl_uv;
	const uint32_t b = ((uint32_t)ty * (uint32_t)im.mcu_x + (uint32_t)(m0 + m)) * 6u + (uint32_t)j;
	int16_t *out = coef + ((size_t)im.blk_first + b) * 64;
#pragma unroll
	for (int i = 0; i < 64; ++i)
	{
		const float q = EF_M(d[i], fd[i]);
		out[c_enc_zigzag[i]] = (int16_t)(int)(q < 0 ? EF_S(q, 0.5f) : EF_A(q, 0.5f));
	}
}

/* previous block of the same component in encode order, or -1 */
__device__ __forceinline__ int64_t enc_prev_block(const jcu_enc_image &im, uint32_t b)
{
	const uint32_t per = im.subsample ? 6u : 3u;
	const uint32_t m = b / per, j = b - m * per;
	if (im.subsample && j >= 1u && j <= 3u)
		return (int64_t)b - 1;
	if (m == 0u)
		return -1;
	if (im.subsample && j == 0u)
		return (int64_t)(b - per + 3u); /* Y3 of the previous MCU */
	return (int64_t)(b - per);
}

/* stbiw__jpg_calcBits (jpeg_write.c:76-86): magnitude category and the bits that follow the code */
__device__ __forceinline__ void enc_calc_bits(int val, uint32_t &bits, int &len)
{
	const int a = val < 0 ? -val : val;
	const int v = val < 0 ? val - 1 : val;
	len = a > 1 ? 32 - __clz(a) : 1;
	bits = (uint32_t)v & ((1u << len) - 1u);
}

/* the symbol walk of stbiw__jpg_processDU (jpeg_write.c:120-168); PUT(code, length) is called for every emitted field */
template <typename PUT>
__device__ __forceinline__ void enc_walk(const int16_t *__restrict__ du, int dc_prev, int is_chroma, PUT put)
{
	const uint32_t *hdc = c_enc_huff[is_chroma ? 1 : 0], *hac = c_enc_huff[is_chroma ? 3 : 2];
	const int diff = (int)du[0] - dc_prev;
	if (diff == 0)
		put(hdc[0] & 0xffffu, (int)(hdc[0] >> 16));
	else
	{
		uint32_t bits;
		int len;
		enc_calc_bits(diff, bits, len);
		put(hdc[len & 255] & 0xffffu, (int)(hdc[len & 255] >> 16));
		put(bits, len);
	}
	int end0 = 63;
	while (end0 > 0 && du[end0] == 0)
		--end0;
	if (end0 == 0)
	{
		put(hac[0] & 0xffffu, (int)(hac[0] >> 16));
		return;
	}
	for (int i = 1; i <= end0; ++i)
	{
		const int start = i;
		while (du[i] == 0 && i <= end0)
			++i;
		int nz = i - start;
		if (nz >= 16)
		{
			for (int k = 0; k < (nz >> 4); ++k)
				put(hac[0xF0] & 0xffffu, (int)(hac[0xF0] >> 16));
			nz &= 15;
		}
		uint32_t bits;
		int len;
		enc_calc_bits((int)du[i], bits, len);
		const uint32_t h = hac[((nz << 4) + len) & 255];
		put(h & 0xffffu, (int)(h >> 16));
		put(bits, len);
	}
	if (end0 != 63)
		put(hac[0] & 0xffffu, (int)(hac[0] >> 16));
}

__device__ __forceinline__ int enc_dc_prev(const jcu_enc_image &im, const int16_t *__restrict__ coef, uint32_t b)
{
	const int64_t p = enc_prev_block(im, b);
	return p < 0 ? 0 : (int)coef[((size_t)im.blk_first + (size_t)p) * 64];
}

__global__ void __launch_bounds__(ENC_THREADS)
k_enc_count(const jcu_enc_image *__restrict__ images, const int16_t *__restrict__ coef, uint32_t *__restrict__ nbits)
{
	const jcu_enc_image &im = images[blockIdx.y];
	const uint32_t b = blockIdx.x * ENC_THREADS + threadIdx.x;
	if (b >= (uint32_t)im.nblocks)
		return;
	const uint32_t per = im.subsample ? 6u : 3u, j = b % per;
	const int chroma = im.subsample ? (j >= 4u) : (j >= 1u);
	uint32_t total = 0;
	enc_walk(coef + ((size_t)im.blk_first + b) * 64, enc_dc_prev(im, coef, b), chroma, [&](uint32_t, int len) { total += (uint32_t)len; });
	nbits[im.blk_first + b] = total;
}

/*
 * Per image: exclusive prefix sum of vals[first .. first + n) into offs, total into totals[image].
 * grid = images, 1024 threads.  `mode` 0: blocks (first = blk_first, n = nblocks); 1: stuffing chunks.
 */
__global__ void __launch_bounds__(1024)
k_enc_scan(const jcu_enc_image *__restrict__ images, const uint32_t *__restrict__ vals, uint32_t *__restrict__ offs,
			  uint32_t *__restrict__ totals, const uint32_t *__restrict__ bit_totals, int mode)
{
	__shared__ uint32_t s_part[1024];
	const jcu_enc_image &im = images[blockIdx.x];
	const int tid = threadIdx.x;
	size_t first;
	uint32_t n;
	if (mode == 0)
	{
		first = im.blk_first;
		n = (uint32_t)im.nblocks;
	}
	else
	{
		first = im.blk_first * (ENC_BLOCK_BYTES / ENC_CHUNK);
		n = ((bit_totals[blockIdx.x] + 7u) / 8u + ENC_CHUNK - 1u) / ENC_CHUNK;
	}
	const uint32_t chunk = (n + 1023u) / 1024u;
	const uint32_t lo = min(n, (uint32_t)tid * chunk), hi = min(n, lo + chunk);
	uint32_t sum = 0;
	for (uint32_t i = lo; i < hi; ++i)
		sum += vals[first + i];
	s_part[tid] = sum;
	__syncthreads();
	if (tid == 0)
	{
		uint32_t run = 0;
		for (int i = 0; i < 1024; ++i)
		{
			const uint32_t v = s_part[i];
			s_part[i] = run;
			run += v;
		}
		totals[blockIdx.x] = run;
	}
	__syncthreads();
	uint32_t run = s_part[tid];
	for (uint32_t i = lo; i < hi; ++i)
	{
		offs[first + i] = run;
		run += vals[first + i];
	}
}

/* bits of one block at its bit offset; the buffer is zeroed, big-endian bit order */
__global__ void __launch_bounds__(ENC_THREADS)
k_enc_emit(const jcu_enc_image *__restrict__ images, const int16_t *__restrict__ coef, const uint32_t *__restrict__ bitoff,
			  uint8_t *__restrict__ bitbuf)
{
	const jcu_enc_image &im = images[blockIdx.y];
	const uint32_t b = blockIdx.x * ENC_THREADS + threadIdx.x;
	if (b >= (uint32_t)im.nblocks)
		return;
	const uint32_t per = im.subsample ? 6u : 3u, j = b % per;
	const int chroma = im.subsample ? (j >= 4u) : (j >= 1u);
	uint32_t *words = reinterpret_cast<uint32_t *>(bitbuf + im.blk_first * ENC_BLOCK_BYTES);
	const uint32_t off = bitoff[im.blk_first + b];
	uint32_t widx = off >> 5;
	int cnt = (int)(off & 31u); /* leading bits of the first word belong to the previous block */
	uint64_t acc = 0;
	bool first = true;
	enc_walk(coef + ((size_t)im.blk_first + b) * 64, enc_dc_prev(im, coef, b), chroma, [&](uint32_t code, int len) {
		acc = (acc << len) | code;
		cnt += len;
		if (cnt >= 32)
		{
			const uint32_t w = __byte_perm((uint32_t)(acc >> (cnt - 32)), 0, 0x0123);
			if (first)
				atomicOr(&words[widx], w);
			else
				words[widx] = w;
			first = false;
			++widx;
			cnt -= 32;
			acc &= (1ull << cnt) - 1ull;
		}
	});
	if (cnt > 0)
		atomicOr(&words[widx], __byte_perm((uint32_t)(acc << (32 - cnt)), 0, 0x0123));
}

/* byte i of the unstuffed stream of an image, with the final 7 one-bits of padding (jpeg_write.c:360) */
__device__ __forceinline__ uint32_t enc_byte(const uint8_t *__restrict__ src, uint32_t i, uint32_t total_bits)
{
	uint32_t v = src[i];
	if (i == (total_bits >> 3) && (total_bits & 7u))
		v |= 0xffu >> (total_bits & 7u);
	return v;
}

/* grid = (ceil(max chunks / ENC_THREADS), images): 0xFF bytes per ENC_CHUNK-byte chunk */
__global__ void __launch_bounds__(ENC_THREADS)
k_enc_ffcount(const jcu_enc_image *__restrict__ images, const uint8_t *__restrict__ bitbuf, const uint32_t *__restrict__ bit_totals,
				  uint32_t *__restrict__ ffcount)
{
	const jcu_enc_image &im = images[blockIdx.y];
	const uint32_t tb = bit_totals[blockIdx.y], nbytes = (tb + 7u) / 8u, nchunks = (nbytes + ENC_CHUNK - 1u) / ENC_CHUNK;
	const uint32_t c = blockIdx.x * ENC_THREADS + threadIdx.x;
	if (c >= nchunks)
		return;
	const uint8_t *src = bitbuf + im.blk_first * ENC_BLOCK_BYTES;
	uint32_t n = 0;
	for (uint32_t i = c * ENC_CHUNK; i < min(nbytes, (c + 1u) * ENC_CHUNK); ++i)
		n += enc_byte(src, i, tb) == 0xffu;
	ffcount[im.blk_first * (ENC_BLOCK_BYTES / ENC_CHUNK) + c] = n;
}

/* stuffed copy behind the header, then EOI; out_len[image] = total file length (0 if it does not fit out_cap) */
__global__ void __launch_bounds__(ENC_THREADS)
k_enc_stuff(const jcu_enc_image *__restrict__ images, const uint8_t *__restrict__ bitbuf, const uint32_t *__restrict__ bit_totals,
				const uint32_t *__restrict__ ffoff, const uint32_t *__restrict__ fftotal, const uint8_t *__restrict__ headers,
				uint8_t *__restrict__ out, uint32_t *__restrict__ out_len)
{
	const jcu_enc_image &im = images[blockIdx.y];
	const uint32_t tb = bit_totals[blockIdx.y], nbytes = (tb + 7u) / 8u, nchunks = (nbytes + ENC_CHUNK - 1u) / ENC_CHUNK;
	const uint32_t total = im.hdr_len + nbytes + fftotal[blockIdx.y] + 2u;
	const bool fits = total <= im.out_cap;
	uint8_t *dst = out + im.out_off;
	if (blockIdx.x == 0 && threadIdx.x == 0)
	{
		out_len[blockIdx.y] = fits ? total : 0u;
		if (fits)
		{
			dst[total - 2u] = 0xff;
			dst[total - 1u] = 0xd9;
		}
	}
	if (blockIdx.x == 0 && fits)
		for (uint32_t i = threadIdx.x; i < im.hdr_len; i += ENC_THREADS)
			dst[i] = headers[(size_t)blockIdx.y * JCU_ENC_HDR_STRIDE + i];
	const uint32_t c = blockIdx.x * ENC_THREADS + threadIdx.x;
	if (c >= nchunks || !fits)
		return;
	const uint8_t *src = bitbuf + im.blk_first * ENC_BLOCK_BYTES;
	uint32_t o = im.hdr_len + c * ENC_CHUNK + ffoff[im.blk_first * (ENC_BLOCK_BYTES / ENC_CHUNK) + c];
	for (uint32_t i = c * ENC_CHUNK; i < min(nbytes, (c + 1u) * ENC_CHUNK); ++i)
	{
		const uint32_t v = enc_byte(src, i, tb);
		dst[o++] = (uint8_t)v;
		if (v == 0xffu)
			dst[o++] = 0;
	}
}
