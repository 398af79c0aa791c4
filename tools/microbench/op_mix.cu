// Pipe-sharing microbenchmark (sm_100a): single ops and pairs via opaque inline PTX.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
#define ITER 256
#define UNR 8
enum { IMAD, IADD, LOP, SHFN, PRMT, VMAX, SHRS, SEXT, FFMA, DP2A, I2IP, VMINRELU, LEA, SHL, MULHI, MADHI, FADD, FMUL, I2F, F2I, NOPS };
template <int OP> __device__ __forceinline__ uint32_t op(uint32_t x, uint32_t a, uint32_t b)
{
	uint32_t d;
	if (OP == IMAD) asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(a), "r"(b));
	else if (OP == IADD) asm volatile("add.s32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == LOP) asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(x), "r"(a), "r"(b));
	else if (OP == SHFN) asm volatile("shf.r.wrap.b32 %0, %1, %2, 7;" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == PRMT) asm volatile("prmt.b32 %0, %1, %2, 0x2143;" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == VMAX) asm volatile("max.s32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == SHRS) asm volatile("shr.s32 %0, %1, 5;" : "=r"(d) : "r"(x));
	else if (OP == SEXT) asm volatile("{.reg .s16 t; cvt.s16.s32 t, %1; cvt.s32.s16 %0, t;}" : "=r"(d) : "r"(x));
	else if (OP == FFMA) asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(a), "r"(b));
	else if (OP == DP2A) asm volatile("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(x), "r"(b));
	else if (OP == I2IP) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(a), "r"(b));
	else if (OP == VMINRELU) asm volatile("min.s32.relu %0, %1, %2;" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == LEA) asm volatile("{.reg .u32 t; shl.b32 t, %1, 3; add.u32 %0, t, %2;}" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == SHL) asm volatile("shl.b32 %0, %1, 3;" : "=r"(d) : "r"(x));
	else if (OP == MULHI) asm volatile("mul.hi.s32 %0, %1, 4194304;" : "=r"(d) : "r"(x));
	else if (OP == MADHI) asm volatile("mad.hi.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(a), "r"(b));
	else if (OP == FADD) asm volatile("add.rn.f32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == FMUL) asm volatile("mul.rn.f32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(a));
	else if (OP == I2F) asm volatile("cvt.rn.f32.s32 %0, %1;" : "=r"(d) : "r"(x));
	else if (OP == F2I) asm volatile("cvt.rni.s32.f32 %0, %1;" : "=r"(d) : "r"(x));
	return d;
}
template <int A, int B, int RA, int RB> __global__ void k(uint32_t *out, uint32_t a, uint32_t b, long long *cyc)
{
	uint32_t r[UNR];
	for (int i = 0; i < UNR; ++i) r[i] = threadIdx.x * 2654435761u + i * 40503u + a;
	__syncthreads();
	long long t0 = clock64();
	for (int it = 0; it < ITER; ++it)
	{
#pragma unroll
		for (int i = 0; i < UNR; ++i)
		{
#pragma unroll
			for (int q = 0; q < RA; ++q) r[i] = op<A>(r[i], a, b);
#pragma unroll
			for (int q = 0; q < RB; ++q) r[i] = op<B>(r[i], a, b);
		}
	}
	long long t1 = clock64();
	uint32_t s = 0;
	for (int i = 0; i < UNR; ++i) s ^= r[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int A, int B, int RA, int RB> void run(const char *name, uint32_t *out, long long *cyc)
{
	const int threads = 1024, blocks = 148;
	k<A, B, RA, RB><<<blocks, threads>>>(out, 3, 5, cyc);
	k<A, B, RA, RB><<<blocks, threads>>>(out, 3, 5, cyc);
	cudaDeviceSynchronize();
	long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
	double c = 0; for (int i = 0; i < blocks; ++i) c += (double)h[i]; c /= blocks;
	printf("%-22s %7.1f lane-ops/clk/SM\n", name, (double)threads * ITER * UNR * (RA + RB) / c);
}
#define R1(X) run<X, X, 1, 0>(#X, out, cyc)
#define R2(X, Y) run<X, Y, 1, 1>(#X "+" #Y, out, cyc)
int main()
{
	uint32_t *out; long long *cyc;
	cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
	R1(IMAD); R1(IADD); R1(LOP); R1(SHFN); R1(PRMT); R1(VMAX); R1(SHRS); R1(SEXT); R1(FFMA); R1(DP2A); R1(I2IP); R1(VMINRELU); R1(LEA); R1(SHL);
	R2(IMAD, IADD); R2(IMAD, LOP); R2(IMAD, SHFN); R2(IMAD, PRMT); R2(IMAD, VMAX); R2(IMAD, SHRS); R2(IMAD, DP2A); R2(IMAD, I2IP); R2(IMAD, FFMA);
	R2(PRMT, SHFN); R2(PRMT, IADD); R2(SHRS, IADD); R2(SHRS, PRMT); R2(DP2A, I2IP); R2(DP2A, IADD); R2(I2IP, IADD); R2(IADD, LOP); R2(VMAX, IADD); R2(DP2A, PRMT); R2(I2IP, PRMT);
	R1(MULHI); R1(MADHI); R1(FADD); R1(FMUL); R1(I2F); R1(F2I);
	R2(IMAD, MULHI); R2(IADD, MULHI); R2(FFMA, IADD); R2(FFMA, SHRS); R2(FFMA, MULHI); R2(FFMA, I2IP); R2(FFMA, PRMT); R2(FFMA, LEA); R2(IMAD, LEA); R2(IADD, LEA);
	run<FFMA, IMAD, 2, 1>("2xFFMA+IMAD", out, cyc); run<FFMA, IADD, 2, 1>("2xFFMA+IADD", out, cyc); run<FFMA, IMAD, 3, 1>("3xFFMA+IMAD", out, cyc);
	run<IMAD, IADD, 1, 2>("IMAD+2xIADD", out, cyc); run<IMAD, IADD, 1, 3>("IMAD+3xIADD", out, cyc);
	printf("%s\n", cudaGetErrorString(cudaGetLastError()));
	return 0;
}
