/*
 * Widening store epilogues: the reference widens the 8-bit JPEG result on the host,
 *   16 bit: stbi__convert_8_to_16        codec/convert.c:18-35     out = (v << 8) + v
 *   float : stbi__ldr_to_hdr              codec/common.c:391-424    out = (float)(pow(v / 255.0f, gamma) * scale),
 *                                                                   alpha (last channel of 2 / 4): v / 255.0f
 * Both depend on the byte (and its channel) alone, so the device applies them as a table look-up behind the
 * output kernels and the wide samples travel to the host once.  The 512-entry float table ([0,256) colour,
 * [256,512) alpha) is built by the caller with the host's pow(), which keeps the result bit-identical to the
 * reference whatever the device's libm rounds to.  HBM-bound: 1 byte read, 2 / 4 bytes written per sample.
 */
#pragma once
#include <cstdint>

#define WIDEN_THREADS 256
#define WIDEN_PER_THREAD 16 /* samples per thread: one 16-byte load */

__global__ void __launch_bounds__(WIDEN_THREADS) k_widen16(const uint8_t *__restrict__ in, uint16_t *__restrict__ out, size_t n)
{
	const size_t i0 = ((size_t)blockIdx.x * WIDEN_THREADS + threadIdx.x) * WIDEN_PER_THREAD;
	if (i0 >= n)
		return;
	if (i0 + WIDEN_PER_THREAD <= n)
	{
		const uint4 w = *reinterpret_cast<const uint4 *>(in + i0);
		const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
		uint32_t o[8];
#pragma unroll
		for (int j = 0; j < 4; ++j)
		{
			/* bytes b0..b3 -> halves b*257: __byte_perm duplicates each byte into both bytes of its half */
			o[2 * j] = __byte_perm(ws[j], 0, 0x1100);
			o[2 * j + 1] = __byte_perm(ws[j], 0, 0x3322);
		}
		uint4 *d = reinterpret_cast<uint4 *>(out + i0);
		d[0] = make_uint4(o[0], o[1], o[2], o[3]);
		d[1] = make_uint4(o[4], o[5], o[6], o[7]);
	}
	else
		for (size_t i = i0; i < n; ++i)
			out[i] = (uint16_t)(in[i] * 257u);
}

__global__ void __launch_bounds__(WIDEN_THREADS)
k_widenf(const uint8_t *__restrict__ in, float *__restrict__ out, size_t n, int comp, int nonalpha, const float *__restrict__ lut512)
{
	__shared__ float lut[512];
	for (int i = threadIdx.x; i < 512; i += WIDEN_THREADS)
		lut[i] = lut512[i];
	__syncthreads();
	const size_t i0 = ((size_t)blockIdx.x * WIDEN_THREADS + threadIdx.x) * WIDEN_PER_THREAD;
	if (i0 >= n)
		return;
	int ch = (int)(i0 % (size_t)comp); /* channel of the first sample; pixels are comp interleaved bytes */
	if (i0 + WIDEN_PER_THREAD <= n)
	{
		const uint4 w = *reinterpret_cast<const uint4 *>(in + i0);
		const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
		float4 *d = reinterpret_cast<float4 *>(out + i0);
#pragma unroll
		for (int j = 0; j < 4; ++j)
		{
			float f[4];
#pragma unroll
			for (int k = 0; k < 4; ++k)
			{
				f[k] = lut[((ws[j] >> (8 * k)) & 255u) + (ch >= nonalpha ? 256u : 0u)];
				ch = (ch + 1 == comp) ? 0 : ch + 1;
			}
			d[j] = make_float4(f[0], f[1], f[2], f[3]);
		}
	}
	else
		for (size_t i = i0; i < n; ++i)
		{
			out[i] = lut[in[i] + (ch >= nonalpha ? 256u : 0u)];
			ch = (ch + 1 == comp) ? 0 : ch + 1;
		}
}
