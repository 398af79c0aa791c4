/*
 * test_hooks.cu -- TEST BUILD ONLY: libimage_codecs_b200_testhooks.so, never linked into the product library.
 * Exposes device arithmetic that has no entry point of its own in the product so that the parity tests can check it in
 * isolation: the colour conversion of jpeg_output.cuh (stbi__YCbCr_to_RGB_row, jpeg.c:1976-2018).
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include "jpeg_output.cuh"

__global__ void k_test_ycbcr(const uint8_t *y, const uint8_t *cb, const uint8_t *cr, int n, int step, uint8_t *out)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	int r, g, b;
	ycbcr_to_rgb(y[i], cb[i], cr[i], r, g, b);
	out[(size_t)i * step] = (uint8_t)r;
	out[(size_t)i * step + 1] = (uint8_t)g;
	out[(size_t)i * step + 2] = (uint8_t)b;
	if (step == 4)
		out[(size_t)i * step + 3] = 255;
}

#define HOOK_CK(call)           \
	do                           \
	{                            \
		cudaError_t e__ = (call); \
		if (e__ != cudaSuccess)   \
			return (int)e__;       \
	} while (0)

extern "C" __attribute__((visibility("default"))) int jcu_test_ycbcr(int device, const uint8_t *y, const uint8_t *cb, const uint8_t *cr, int n,
																							int step, uint8_t *out)
{
	uint8_t *d_in, *d_o;
	HOOK_CK(cudaSetDevice(device));
	HOOK_CK(cudaMalloc(&d_in, (size_t)n * 3));
	HOOK_CK(cudaMalloc(&d_o, (size_t)n * step));
	HOOK_CK(cudaMemcpy(d_in, y, n, cudaMemcpyHostToDevice));
	HOOK_CK(cudaMemcpy(d_in + n, cb, n, cudaMemcpyHostToDevice));
	HOOK_CK(cudaMemcpy(d_in + 2 * (size_t)n, cr, n, cudaMemcpyHostToDevice));
	k_test_ycbcr<<<(n + 255) / 256, 256>>>(d_in, d_in + n, d_in + 2 * (size_t)n, n, step, d_o);
	HOOK_CK(cudaDeviceSynchronize());
	HOOK_CK(cudaMemcpy(out, d_o, (size_t)n * step, cudaMemcpyDeviceToHost));
	cudaFree(d_in);
	cudaFree(d_o);
	return 0;
}
