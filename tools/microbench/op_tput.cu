// Instruction-throughput microbenchmark (sm_100a): lane-ops per clock per SM for the integer ops
// the JPEG kernels are built from.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 op_tput.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

#define ITER 256
#define UNR 8

template <int OP> __device__ __forceinline__ void step(uint32_t (&r)[UNR], uint32_t a, uint32_t b)
{
#pragma unroll
	for (int i = 0; i < UNR; ++i)
	{
		uint32_t x = r[i];
		if (OP == 0) x = x * a + b;                                   // IMAD
		else if (OP == 1) x = x + a + b;                              // IADD3
		else if (OP == 2) x = (x & a) ^ b;                            // LOP3
		else if (OP == 3) x = __funnelshift_r(x, a, 7);               // SHF
		else if (OP == 4) x = __byte_perm(x, a, 0x2143);              // PRMT
		else if (OP == 5) x = (uint32_t)max((int)x, (int)a);          // VIMNMX
		else if (OP == 6) x = (uint32_t)__dp2a_lo((int)a, (int)x, (int)b);       // IDP.2A
		else if (OP == 7) x = (uint32_t)__dp4a((int)a, (int)x, (int)b);          // IDP.4A
		else if (OP == 8) { uint32_t d; asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(a), "r"(b)); x = d; } // I2IP
		else if (OP == 9) x = (uint32_t)__viaddmax_s32_relu((int)x, (int)a, (int)b); // VIADDMNMX
		else if (OP == 10) x = (uint32_t)(((int)x) >> 5);             // SHF.R.S32 / arithmetic shift
		else if (OP == 11) x = __float_as_uint(__uint_as_float(x) * __uint_as_float(a) + __uint_as_float(b)); // FFMA
		else if (OP == 12) x = (uint32_t)__float2int_rd(__uint_as_float(x));     // F2I
		else if (OP == 13) x = __float_as_uint((float)(int)x);        // I2F
		else if (OP == 14) x = __shfl_xor_sync(0xffffffffu, x, 1);    // SHFL
		else if (OP == 15) { int v = (int)x; x = (uint32_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); } // clamp (2 VIMNMX or 1 w/ relu)
		else if (OP == 16) x = (uint32_t)__vimin_s32_relu((int)x, (int)a);       // VIMNMX.RELU
		else if (OP == 17) x = (x << 3) + a;                          // LEA
		else if (OP == 18) x = (uint32_t)(int)(short)x;               // sext16 (PRMT?)
		else if (OP == 19) x = (uint32_t)__mulhi((int)x, (int)a);     // IMAD.HI
		r[i] = x;
	}
}

template <int OP> __global__ void k(uint32_t *out, uint32_t a, uint32_t b, long long *cyc)
{
	uint32_t r[UNR];
	for (int i = 0; i < UNR; ++i) r[i] = threadIdx.x * 2654435761u + i * 40503u + a;
	__syncthreads();
	long long t0 = clock64();
	for (int it = 0; it < ITER; ++it) step<OP>(r, a, b);
	long long t1 = clock64();
	uint32_t s = 0;
	for (int i = 0; i < UNR; ++i) s ^= r[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP> void run(const char *name, uint32_t *out, long long *cyc)
{
	const int threads = 1024, blocks = 148;
	k<OP><<<blocks, threads>>>(out, 3, 5, cyc);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	cudaEventRecord(e0);
	k<OP><<<blocks, threads>>>(out, 3, 5, cyc);
	cudaEventRecord(e1); cudaEventSynchronize(e1);
	long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
	double c = 0; for (int i = 0; i < blocks; ++i) c += (double)h[i]; c /= blocks;
	double ops = (double)threads * ITER * UNR;
	printf("%-12s %8.1f lane-ops/clk/SM   (%.0f cycles)\n", name, ops / c, c);
}

int main()
{
	uint32_t *out; long long *cyc;
	cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
	run<0>("IMAD", out, cyc); run<1>("IADD3", out, cyc); run<2>("LOP3", out, cyc); run<3>("SHF.funnel", out, cyc);
	run<4>("PRMT", out, cyc); run<5>("VIMNMX", out, cyc); run<6>("IDP.2A", out, cyc); run<7>("IDP.4A", out, cyc);
	run<8>("I2IP.sat", out, cyc); run<9>("VIADDMNMX", out, cyc); run<10>("SHR.S32", out, cyc); run<11>("FFMA", out, cyc);
	run<12>("F2I", out, cyc); run<13>("I2F", out, cyc); run<14>("SHFL", out, cyc); run<15>("clamp0-255", out, cyc);
	run<16>("VIMNMX.RELU", out, cyc); run<17>("LEA", out, cyc); run<18>("sext16", out, cyc); run<19>("IMAD.HI", out, cyc);
	printf("%s\n", cudaGetErrorString(cudaGetLastError()));
	return 0;
}
