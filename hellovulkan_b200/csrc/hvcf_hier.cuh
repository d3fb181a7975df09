// hvcf_hier.cuh -- device helpers shared by the kernels of HVCF_MODE_HIERARCHICAL (hvcf_hier.cu, hvcf_bin.cu).
#pragma once

#include "hvcf_kernels.cuh"

namespace hvcf
{

constexpr uint32_t kSkip = 0xffffffffu;
constexpr uint32_t kCand = 4;       // level-2 candidates in flight per trip (6 measured: A65k 0.0509 -> 0.0531 ms, profiles/r01zn)

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p)
{
	unsigned long long v;
	asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v)
{
	asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// order-preserving float <-> int map, for redux.sync min/max over a warp
__device__ __forceinline__ int float_to_ordered(float f)
{
	const int i = __float_as_int(f);
	return i ^ ((i >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float ordered_to_float(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7fffffff)); }
__device__ __forceinline__ float warp_min_f(float v) { return ordered_to_float(__reduce_min_sync(0xffffffffu, float_to_ordered(v))); }
__device__ __forceinline__ float warp_max_f(float v) { return ordered_to_float(__reduce_max_sync(0xffffffffu, float_to_ordered(v))); }

// LightCulling.comp:26-47,56 on one light and one box; see cull_kernel for why max3(lo-v, v-hi, 0) is exact.
__device__ __forceinline__ bool sphere_box_hit(float x, float y, float z, float r2, const float lo[3], const float hi[3])
{
	const float dx = fmaxf(fmaxf(__fsub_rn(lo[0], x), __fsub_rn(x, hi[0])), 0.0f);
	const float dy = fmaxf(fmaxf(__fsub_rn(lo[1], y), __fsub_rn(y, hi[1])), 0.0f);
	const float dz = fmaxf(fmaxf(__fsub_rn(lo[2], z), __fsub_rn(z, hi[2])), 0.0f);
	const float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
	return s < r2;
}

// the same test on the lane's light PAIR of the packed culling layout (FADD2 / FMUL2; sums stay scalar so
// that ptxas cannot contract them into FFMA2)
__device__ __forceinline__ void sphere_box_hit2(const float4& a, const float4& b, const float lo[3], const float hi[3], bool& h0,
	bool& h1)
{
	const f32x2 X = pack2(a.x, a.y), Y = pack2(a.z, a.w), Z = pack2(b.x, b.y);
	float l0, l1, u0, u1;
	unpack2(sub2(pack2(lo[0], lo[0]), X), l0, l1);
	unpack2(sub2(X, pack2(hi[0], hi[0])), u0, u1);
	const f32x2 dx = pack2(fmaxf(fmaxf(l0, u0), 0.0f), fmaxf(fmaxf(l1, u1), 0.0f));
	unpack2(sub2(pack2(lo[1], lo[1]), Y), l0, l1);
	unpack2(sub2(Y, pack2(hi[1], hi[1])), u0, u1);
	const f32x2 dy = pack2(fmaxf(fmaxf(l0, u0), 0.0f), fmaxf(fmaxf(l1, u1), 0.0f));
	unpack2(sub2(pack2(lo[2], lo[2]), Z), l0, l1);
	unpack2(sub2(Z, pack2(hi[2], hi[2])), u0, u1);
	const f32x2 dz = pack2(fmaxf(fmaxf(l0, u0), 0.0f), fmaxf(fmaxf(l1, u1), 0.0f));
	float sx0, sx1, sy0, sy1, sz0, sz1;
	unpack2(mul2(dx, dx), sx0, sx1);
	unpack2(mul2(dy, dy), sy0, sy1);
	unpack2(mul2(dz, dz), sz0, sz1);
	h0 = __fadd_rn(__fadd_rn(sx0, sy0), sz0) < b.z;
	h1 = __fadd_rn(__fadd_rn(sx1, sy1), sz1) < b.w;
}

// hits of the chunks before `chunk` for one cluster.  cum is laid out [cluster][kHMaxChunks] so that the whole row is two
// 128-bit loads (it was [chunk][cluster] with one predicated 32-bit load per chunk: the hottest line of the count pass,
// 10 % of its instructions and the most stall samples, profiles/r01b).  ld.global.cg: other SMs add to it while we read;
// any value read is a lower bound of the true count, which is all the early exit needs (hvcf_hier.cu header).
static_assert(kHMaxChunks == 8, "hier_prior_hits reads the row as two uint4");
__device__ __forceinline__ uint32_t hier_prior_hits(const uint32_t* cum, uint32_t cluster, uint32_t chunk)
{
	if (chunk == 0)
		return 0u;
	const uint4* row = reinterpret_cast<const uint4*>(cum + static_cast<size_t>(cluster) * kHMaxChunks);
	const uint4 a = __ldcg(row);
	uint32_t prior = a.x + (chunk > 1 ? a.y : 0u) + (chunk > 2 ? a.z : 0u) + (chunk > 3 ? a.w : 0u);
	if (chunk > 4)
	{
		const uint4 b = __ldcg(row + 1);
		prior += b.x + (chunk > 5 ? b.y : 0u) + (chunk > 6 ? b.z : 0u) + (chunk > 7 ? b.w : 0u);
	}
	return prior;
}

// segment -> chunk: chunk 0 = segments [0,4), chunk c >= 1 = segments [2^(c+1), 2^(c+2))
// (the last chunk, kHMaxChunks - 1, takes every segment from 256 on: any monotone map keeps the lower bounds valid)
__device__ __forceinline__ uint32_t hier_chunk_of(uint32_t seg) { return seg < 4 ? 0u : min(30u - __clz(seg), kHMaxChunks - 1u); }
} // namespace hvcf
