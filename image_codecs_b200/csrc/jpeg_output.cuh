/*
 * jpeg_output.cuh -- dequantisation, integer 8x8 IDCT, chroma upsampling, colour conversion.
 *
 * Arithmetic contracts reproduced bit-exactly:
 *   dequant   (short)(coef * q)                          jpeg.c:325,345,365 and :1323
 *   IDCT      stbi__idct_block, int32, 12-bit constants  jpeg.c:578-679
 *   upsample  stbi__resample_row_{v_2,h_2,hv_2,generic}  jpeg.c:1774-1840,1962
 *   colour    stbi__YCbCr_to_RGB_row (20-bit fixed)      jpeg.c:1976-2018
 *             stbi__blinn_8x8 :2218, stbi__compute_y common.c:173
 * The generic C kernels are canonical, not the SSE2/NEON variants.
 *
 * This header holds the shared device arithmetic and the GENERIC two-pass path (any H,V in
 * 1..4, any colour mode).  The fused single-pass strip kernels for the headline formats are in
 * jpeg_fused420.cuh (4:2:0) and jpeg_fused1x1.cuh (4:4:4, single component) and use the same arithmetic.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/jpeg_cuda_abi.h"
#include "jpeg_coef_layout.cuh"

/* ---- 1-D inverse DCT butterfly, 12-bit constants f2f(x) = (int)(x*4096+0.5) (jpeg.c:574-613).
 * `bias` is added to the four even-part sums (x0..x3 in the reference): 512 in the column pass,
 * 65536 + (128<<17) in the row pass.  Outputs are NOT yet shifted.  int32 wrap-around addition is
 * associative, so the regrouping below gives the reference's exact 32-bit values. ---- */
/* FMA_A: the two shift-adds a0, a1 as multiply-adds whose multiplier comes from constant memory (ptxas cannot turn
   them back into LEA): the fused kernels are bound by the ALU pipe, the FMA pipe has room */
__constant__ int c_idct_4096 = 4096;
__device__ __forceinline__ int idct_mad(int a, int b, int c) /* pinned multiply-add: one IMAD on the FMA pipe */
{
	int d;
	asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
	return d;
}
/* MODE selects how the additions are split between the two math pipes (the fused kernels are bound by the ALU
   pipe: IADD3 / SHF / PRMT / LEA / I2IP; IMAD runs on the FMA pipe).  All modes give the same 32-bit values.
     0  even part factored (3 adds + 5 multiply-adds), outputs as eight IADD3
     1  every addition of the even part rides on a multiply-add (8 IMAD, no ALU op)
     2  MODE 1 and the output butterfly folded into the odd part's chains: o[i] = x_i + d_i is the chain's addend,
        o[7-i] = 2 x_i - o[i] (26 IMAD + 6 ALU instead of 21 + 11) */
template <bool FMA_A = false, int MODE = 0>
__device__ __forceinline__ void idct_1d(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7, int bias,
													 int (&o)[8])
{
	if (MODE == 2)
	{
		const int zb = idct_mad(s2, 2217, idct_mad(s6, 2217, bias));	/* z + bias */
		const int zn = idct_mad(s2, -2217, idct_mad(s6, -2217, bias)); /* -z + bias */
		const int e3b = idct_mad(s2, 3135, zb), e3n = idct_mad(s2, -3135, zn);
		const int e2b = idct_mad(s6, -7567, zb), e2n = idct_mad(s6, 7567, zn);
		const int p = s0 + s4, m = s0 - s4;
		const int x0 = idct_mad(p, 4096, e3b), x3 = idct_mad(p, 4096, e3n);
		const int x1 = idct_mad(m, 4096, e2b), x2 = idct_mad(m, 4096, e2n);
		o[0] = idct_mad(s1, 5683, idct_mad(s3, 4816, idct_mad(s5, 3219, idct_mad(s7, 1131, x0))));
		o[1] = idct_mad(s1, 4816, idct_mad(s3, -1129, idct_mad(s5, -5681, idct_mad(s7, -3218, x1))));
		o[2] = idct_mad(s1, 3219, idct_mad(s3, -5681, idct_mad(s5, 1132, idct_mad(s7, 4816, x2))));
		o[3] = idct_mad(s1, 1131, idct_mad(s3, -3218, idct_mad(s5, 4816, idct_mad(s7, -5680, x3))));
		o[7] = x0 + x0 - o[0];
		o[6] = x1 + x1 - o[1];
		o[5] = x2 + x2 - o[2];
		o[4] = x3 + x3 - o[3];
		return;
	}
	/* even part (jpeg.c:580-591): p1 = (s2+s6)*f2f(0.5411961); t2 = p1 - s6*f2f(1.847759065); t3 = p1 + s2*f2f(0.765366865) */
	const int z = MODE == 1 ? idct_mad(s2, 2217, s6 * 2217) : (s2 + s6) * 2217;
	const int e2 = MODE == 1 ? idct_mad(s6, -7567, z) : z + s6 * -7567;
	const int e3 = MODE == 1 ? idct_mad(s2, 3135, z) : z + s2 * 3135;
	const int a0 = MODE == 1	? idct_mad(s0, 4096, idct_mad(s4, 4096, bias))
						: FMA_A	? (s0 + s4) * c_idct_4096 + bias
									: (s0 + s4) * 4096 + bias;
	const int a1 = MODE == 1	? idct_mad(s0, 4096, idct_mad(s4, -4096, bias))
						: FMA_A	? (s0 - s4) * c_idct_4096 + bias
									: (s0 - s4) * 4096 + bias;
	/* odd part (jpeg.c:592-612) multiplied out: in the ring Z/2^32 the reference's factored butterfly
	   (p5 = (p3+p4)*f2f(1.175875602), t0..t3 *= ..., t3 += p1+p4, ...) equals this 4x4 integer matrix,
	   e.g. d3 = s1*(6149+4816-3685-1597) + s3*4816 + s5*(4816-1597) + s7*(4816-3685); 16 multiply-adds, no adds */
	const int d3 = s1 * 5683 + s3 * 4816 + s5 * 3219 + s7 * 1131;
	const int d2 = s1 * 4816 + s3 * -1129 + s5 * -5681 + s7 * -3218;
	const int d1 = s1 * 3219 + s3 * -5681 + s5 * 1132 + s7 * 4816;
	const int d0 = s1 * 1131 + s3 * -3218 + s5 * 4816 + s7 * -5680;
	/* outputs as three-input sums (IADD3): x0 = a0+e3, x3 = a0-e3, x1 = a1+e2, x2 = a1-e2 are never materialised */
	o[0] = (a0 + d3) + e3;
	o[7] = (a0 - d3) + e3;
	o[1] = (a1 + d2) + e2;
	o[6] = (a1 - d2) + e2;
	o[2] = (a1 + d1) - e2;
	o[5] = (a1 - d1) - e2;
	o[3] = (a0 + d0) - e3;
	o[4] = (a0 - d0) - e3;
}

__device__ __forceinline__ int clamp_u8(int v) { return min(max(v, 0), 255); }

/*
 * Full 8x8 block: v[] holds 64 dequantised coefficients (already truncated to int16 range,
 * sign-extended) in natural order on entry and the column-pass results on exit of pass 1.
 * rows[r] receives the 8 clamped samples of row r packed little-endian into two 32-bit words.
 */
__device__ __forceinline__ void idct_block_regs(int (&v)[64], uint32_t (&rows)[16])
{
	int o[8];
#pragma unroll
	for (int c = 0; c < 8; ++c)
	{
		idct_1d(v[c], v[c + 8], v[c + 16], v[c + 24], v[c + 32], v[c + 40], v[c + 48], v[c + 56], 512, o);
#pragma unroll
		for (int r = 0; r < 8; ++r)
			v[c + 8 * r] = o[r] >> 10;
	}
#pragma unroll
	for (int r = 0; r < 8; ++r)
	{
		idct_1d(v[8 * r], v[8 * r + 1], v[8 * r + 2], v[8 * r + 3], v[8 * r + 4], v[8 * r + 5], v[8 * r + 6],
				  v[8 * r + 7], 65536 + (128 << 17), o);
		uint32_t lo = 0, hi = 0;
#pragma unroll
		for (int c = 0; c < 4; ++c)
		{
			lo |= (uint32_t)clamp_u8(o[c] >> 17) << (8 * c);
			hi |= (uint32_t)clamp_u8(o[c + 4] >> 17) << (8 * c);
		}
		rows[2 * r] = lo;
		rows[2 * r + 1] = hi;
	}
}

/* load 64 raw int16 coefficients (128 contiguous bytes) and dequantise: (short)(c*q) */
__device__ __forceinline__ void load_dequant(const int16_t *__restrict__ blk, const uint16_t *__restrict__ qt,
															int (&v)[64])
{
	const uint4 *p = reinterpret_cast<const uint4 *>(blk);
	const unsigned key = coef_key(blk);
#pragma unroll
	for (int q = 0; q < 8; ++q)
	{
		uint4 w = p[q ^ key]; /* chunk permutation of the slab layout (jpeg_coef_layout.cuh) */
		uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
		for (int j = 0; j < 4; ++j)
		{
			int c0 = (int)(int16_t)(ws[j] & 0xffffu);
			int c1 = (int)(int16_t)(ws[j] >> 16);
			v[q * 8 + 2 * j] = (int)(int16_t)(c0 * (int)qt[q * 8 + 2 * j]);
			v[q * 8 + 2 * j + 1] = (int)(int16_t)(c1 * (int)qt[q * 8 + 2 * j + 1]);
		}
	}
}

/* ---- colour ---- */
/* stbi__YCbCr_to_RGB_row (jpeg.c:1981-2011): constants are float2fixed(x) = ((int)(x*4096+.5))<<8 */
__device__ __forceinline__ void ycbcr_to_rgb(int y, int cb, int cr, int &r, int &g, int &b)
{
	int yf = (y << 20) + (1 << 19);
	cr -= 128;
	cb -= 128;
	r = yf + cr * (5743 << 8);
	g = yf + cr * -(2925 << 8) + (int)((uint32_t)(cb * -(1410 << 8)) & 0xffff0000u);
	b = yf + cb * (7258 << 8);
	r = clamp_u8(r >> 20);
	g = clamp_u8(g >> 20);
	b = clamp_u8(b >> 20);
}
/* stbi__blinn_8x8 (jpeg.c:2218) */
__device__ __forceinline__ int blinn8(int x, int y)
{
	unsigned t = (unsigned)(x * y + 128);
	return (int)(((t + (t >> 8)) >> 8) & 255u);
}
/* stbi__compute_y (common.c:173) */
__device__ __forceinline__ int luma8(int r, int g, int b) { return ((r * 77) + (g * 150) + (29 * b)) >> 8 & 255; }

/* ===================================================================================== */
/* GENERIC path, pass 1: IDCT every block of the components the output needs into planes  */
/* ===================================================================================== */
struct GenericItem
{
	int32_t image;
};

/* grid: (ceil(max blocks/128), n_generic_images, 4 components); one thread per block */
__global__ void __launch_bounds__(128)
k_generic_idct(const int32_t *__restrict__ list, const jcu_image *__restrict__ images,
					const int16_t *__restrict__ coef, uint8_t *__restrict__ planes)
{
	const jcu_image *im = &images[list[blockIdx.y]];
	const int ci = blockIdx.z;
	if (ci >= im->decode_n)
		return;
	const jcu_comp *cp = &im->comp[ci];
	const int nblk = cp->bw * cp->bh;
	const int b = blockIdx.x * blockDim.x + threadIdx.x;
	if (b >= nblk)
		return;
	int v[64];
	uint32_t rows[16];
	load_dequant(coef + cp->coef_off + (size_t)b * 64, im->qt_out[ci], v);
	idct_block_regs(v, rows);
	const int bx = b % cp->bw, by = b / cp->bw;
	uint8_t *dst = planes + cp->plane_off + ((size_t)by * 8) * ((size_t)cp->bw * 8) + (size_t)bx * 8;
#pragma unroll
	for (int r = 0; r < 8; ++r)
		*reinterpret_cast<uint2 *>(dst + (size_t)r * cp->bw * 8) = make_uint2(rows[2 * r], rows[2 * r + 1]);
}

/* one upsampled sample of component cp at output pixel (x, y): row selection jpeg.c:2301-2318 */
__device__ __forceinline__ int generic_sample(const jcu_comp *cp, const uint8_t *__restrict__ plane, int x, int y)
{
	const int stride = cp->bw * 8;
	const int hs = cp->hs, vs = cp->vs, cy = cp->cy, w = cp->w_lores;
	if (hs == 1 && vs == 1)
		return plane[(size_t)min(y, cy - 1) * stride + x];
	if (hs == 1 && vs == 2)
	{
		int k = y >> 1;
		int nr = min(k, cy - 1);
		int fr = (y & 1) ? min(k + 1, cy - 1) : min(max(k - 1, 0), cy - 1);
		return (3 * plane[(size_t)nr * stride + x] + plane[(size_t)fr * stride + x] + 2) >> 2;
	}
	if (hs == 2 && vs == 1)
	{
		const uint8_t *in = plane + (size_t)min(y, cy - 1) * stride;
		int i = x >> 1;
		if (w == 1)
			return in[0];
		if (x == 0)
			return in[0];
		if (x == 2 * w - 1)
			return in[w - 1];
		if (x == 2 * w - 2)
			return (in[w - 2] * 3 + in[w - 1] + 2) >> 2; /* as written at jpeg.c:1805 */
		if (x & 1)
			return (3 * in[i] + in[i + 1] + 2) >> 2;
		return (3 * in[i] + in[i - 1] + 2) >> 2;
	}
	if (hs == 2 && vs == 2)
	{
		int k = y >> 1;
		int nr = min(k, cy - 1);
		int fr = (y & 1) ? min(k + 1, cy - 1) : min(max(k - 1, 0), cy - 1);
		const uint8_t *nrow = plane + (size_t)nr * stride;
		const uint8_t *frow = plane + (size_t)fr * stride;
		int i = x >> 1;
		int j = (x & 1) ? min(i + 1, w - 1) : max(i - 1, 0);
		int ti = 3 * nrow[i] + frow[i];
		int tj = 3 * nrow[j] + frow[j];
		return (3 * ti + tj + 8) >> 4;
	}
	/* nearest neighbour (jpeg.c:1962); vertical by row stepping only */
	return plane[(size_t)min(y / vs, cy - 1) * stride + x / hs];
}

/* grid: (ceil(width*height/256) capped by max, n_generic_images); one thread per output pixel */
__global__ void __launch_bounds__(256)
k_generic_color(const int32_t *__restrict__ list, const jcu_image *__restrict__ images,
					 const uint8_t *__restrict__ planes, uint8_t *__restrict__ out)
{
	const jcu_image *im = &images[list[blockIdx.y]];
	const int W = im->width, H = im->height, n = im->n_out, mode = im->color_mode;
	const long long total = (long long)W * H;
	for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x)
	{
		const int x = (int)(p % W), y = (int)(p / W);
		int s[4] = {0, 0, 0, 0};
		for (int k = 0; k < im->decode_n; ++k)
		{
			/* modes that only use component 0 (and K for YCCK luma) skip the rest */
			if (mode == JCU_CM_Y && k > 0)
				break;
			s[k] = generic_sample(&im->comp[k], planes + im->comp[k].plane_off, x, y);
		}
		uint8_t *o = out + im->out_off + ((size_t)(im->flip ? H - 1 - y : y) * W + x) * n; /* flip on load: convert.c:37 */
		int r, g, b;
		switch (mode)
		{
		case JCU_CM_YCBCR: ycbcr_to_rgb(s[0], s[1], s[2], r, g, b); break;
		case JCU_CM_RGB: r = s[0]; g = s[1]; b = s[2]; break;
		case JCU_CM_CMYK:
			r = blinn8(s[0], s[3]);
			g = blinn8(s[1], s[3]);
			b = blinn8(s[2], s[3]);
			break;
		case JCU_CM_YCCK:
			ycbcr_to_rgb(s[0], s[1], s[2], r, g, b);
			r = blinn8(255 - r, s[3]);
			g = blinn8(255 - g, s[3]);
			b = blinn8(255 - b, s[3]);
			break;
		case JCU_CM_RGB_LUMA: r = g = b = luma8(s[0], s[1], s[2]); break;
		case JCU_CM_CMYK_LUMA:
			r = g = b = luma8(blinn8(s[0], s[3]), blinn8(s[1], s[3]), blinn8(s[2], s[3]));
			break;
		case JCU_CM_YCCK_LUMA: r = g = b = blinn8(255 - s[0], s[3]); break;
		default: r = g = b = s[0]; break;
		}
		if (n >= 3)
		{
			o[0] = (uint8_t)r;
			o[1] = (uint8_t)g;
			o[2] = (uint8_t)b;
			if (n == 4)
				o[3] = 255;
		}
		else
		{
			o[0] = (uint8_t)r;
			if (n == 2)
				o[1] = 255;
		}
	}
}
