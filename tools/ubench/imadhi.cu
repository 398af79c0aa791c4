// Throughput probe: mad.lo vs mul.hi/mad.hi vs shift on sm_100a (instructions per clock per SM).
#include <cstdio>
#include <cuda_runtime.h>
#define CHAINS 8
#define ITERS 16384
template <int MODE>
__global__ void k(int *out, int seed, long long *clk)
{
	int x[CHAINS];
#pragma unroll
	for (int i = 0; i < CHAINS; ++i)
		x[i] = seed + threadIdx.x * 7 + i;
	long long t0 = clock64();
#pragma unroll 1
	for (int it = 0; it < ITERS; ++it)
	{
#pragma unroll
		for (int i = 0; i < CHAINS; ++i)
		{
			if (MODE == 0)
				asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(seed), "r"(it));
			else if (MODE == 1)
				asm volatile("mad.hi.s32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(seed), "r"(it));
			else if (MODE == 2)
				asm volatile("shr.s32 %0, %0, %1;" : "+r"(x[i]) : "r"(seed & 1));
			else if (MODE == 3)
			{ /* mad.lo + shr + add (today's colour channel) */
				int t;
				asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(x[i]), "r"(seed), "r"(it));
				asm volatile("shr.s32 %0, %0, 12;" : "+r"(t));
				asm volatile("add.s32 %0, %0, %1;" : "+r"(x[i]) : "r"(t));
			}
			else if (MODE == 4)
			{ /* mad.lo + mad.hi (proposed) */
				int t;
				asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(x[i]), "r"(seed), "r"(it));
				asm volatile("mad.hi.s32 %0, %1, 1048576, %0;" : "+r"(x[i]) : "r"(t));
			}
			else if (MODE == 5)
				asm volatile("mul.hi.s32 %0, %0, 1048576;" : "+r"(x[i]));
		}
	}
	long long t1 = clock64();
	int s = 0;
#pragma unroll
	for (int i = 0; i < CHAINS; ++i)
		s ^= x[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0 && blockIdx.x == 0)
		*clk = t1 - t0;
}
template <int MODE>
void run(const char *name, int per_iter)
{
	int *out;
	long long *clk;
	cudaMalloc(&out, 148 * 4 * 256 * 4);
	cudaMalloc(&clk, 8);
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	k<MODE><<<148 * 4, 256>>>(out, 3, clk);
	cudaEventRecord(e0);
	k<MODE><<<148 * 4, 256>>>(out, 3, clk);
	cudaEventRecord(e1);
	cudaDeviceSynchronize();
	float ms;
	cudaEventElapsedTime(&ms, e0, e1);
	/* 4 CTAs x 8 warps per SM = 32 warps per SM, 8 per SMSP; 1.965 GHz */
	double inst = (double)ITERS * CHAINS * per_iter * 8; /* warp-instructions per SMSP */
	printf("%-28s %.3f ms -> %.3f warp-inst/clk/SMSP (at 1965 MHz)\n", name, ms, inst / (ms * 1e-3 * 1.965e9));
}
int main()
{
	run<0>("mad.lo", 1);
	run<1>("mad.hi", 1);
	run<5>("mul.hi imm", 1);
	run<2>("shr", 1);
	run<3>("mad.lo+shr+add", 3);
	run<4>("mad.lo+mad.hi", 2);
	return 0;
}
