// hvcf_device.cuh -- device-side helpers for libhvcf (sm_100a only).
//
// Arithmetic rules: IEEE binary32, round-to-nearest-even,
// NO contraction.  Every float op on the parity path goes through an explicit _rn intrinsic or an
// explicit .rn PTX instruction, so the result does not depend on --fmad.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/hvcf_types.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libhvcf is written for sm_100a (B200) only"
#endif

namespace hvcf
{

// ---- culling layout of the per-frame view-space light cache --------------------------------------
// Lights are stored in groups of 64.  A group is 1 KiB: 32 x float4 "A" records followed by
// 32 x float4 "B" records; lane l of a warp owns lights (64g + l) and (64g + 32 + l):
//   A[l] = { x_l, x_{l+32}, y_l, y_{l+32} }      B[l] = { z_l, z_{l+32}, r2_l, r2_{l+32} }
// so that one LDS.128 per record is bank-conflict free and every coordinate of the lane's light
// PAIR already sits in an aligned 64-bit register pair -- the operand shape of the sm_100
// packed-FP32 instructions (FADD2 / FMUL2).  r2 = radius*radius; pad lights have r2 = -1 (never hit).
constexpr uint32_t kGroupLights = 64;
constexpr uint32_t kGroupFloats = 256;
constexpr uint32_t kGroupBytes = 1024;
constexpr uint32_t kTileGroups = 16;                          // one TMA bulk copy = 16 KiB
constexpr uint32_t kTileLights = kTileGroups * kGroupLights;   // 1024 lights
constexpr uint32_t kTileBytes = kTileGroups * kGroupBytes;
constexpr uint32_t kStages = 3;                                // smem ring depth

__host__ __device__ inline uint32_t round_up_u32(uint32_t v, uint32_t m) { return (v + m - 1) / m * m; }

// ---- exact float helpers ---------------------------------------------------------------------------
// GLSL/glm min/max: "y if y < x else x" -- AABBGenerator.comp:85-86, glm/detail/func_common.inl
__device__ __forceinline__ float glsl_min(float x, float y) { return (y < x) ? y : x; }
__device__ __forceinline__ float glsl_max(float x, float y) { return (x < y) ? y : x; }

// mat4 * vec4 in glm's association order, one row r: (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
__device__ __forceinline__ float mat_row_dot(const float* __restrict__ m, int r, float v0, float v1, float v2, float v3)
{
	const float add0 = __fadd_rn(__fmul_rn(m[0 + r], v0), __fmul_rn(m[4 + r], v1));
	const float add1 = __fadd_rn(__fmul_rn(m[8 + r], v2), __fmul_rn(m[12 + r], v3));
	return __fadd_rn(add0, add1);
}

// ---- packed FP32x2 (sm_100: FADD2 / FMUL2) -----------------------------------------------------------
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
	f32x2 r;
	asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
	return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi)
{
	asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b)
{
	f32x2 r;
	asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
	f32x2 r;
	asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
	f32x2 r;
	asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}

// ---- mbarrier + TMA bulk copy (cp.async.bulk, SASS UBLKCP) --------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"HVCF_WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra HVCF_DONE_%=;\n"
		"bra HVCF_WAIT_%=;\n"
		"HVCF_DONE_%=:\n"
		"}\n" ::"r"(smem_u32(bar)),
		"r"(parity)
		: "memory");
}
// global -> shared bulk copy, completion signalled on `bar` (bytes and both addresses 16 B aligned)
__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
					 smem_u32(smem_dst)),
				 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
				 : "memory");
}

// ---- programmatic dependent launch (PDL) --------------------------------------------------------------
// grid_dep_launch: lets the next kernel of the stream (launched with the programmatic-serialization
// attribute) start as soon as every CTA of this grid has been scheduled; grid_dep_wait: blocks until the
// previous grid has COMPLETED and its writes are visible.  Both are no-ops under a plain launch.
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ uint32_t lanemask_lt()
{
	uint32_t m;
	asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
	return m;
}

// device-resident work counters of one hvcf_cull (hvcf_stats is assembled from these on request)
struct DeviceStats
{
	unsigned long long executed_tests;
	unsigned long long prefilter_tests;
	uint32_t total_indices;
	uint32_t overflow_clusters;
	unsigned long long bin_tests;    // level-0 union tests (hier_bin_kernel / hier_bin2_kernel), also inside prefilter_tests
	unsigned long long write_tests;  // union + exact tests the write pass repeats (not in the two counters above)
	// serial of the frames culled on this frame slot, bumped ON THE DEVICE by the frame's first kernel (never cleared): the tag
	// of hier_finalize_kernel's look-back state words.  It lives on the device so that a frame captured into a CUDA graph
	// can be replayed: no kernel argument changes from one frame to the next.
	uint32_t epoch;
	uint32_t pad_;
};
__device__ __forceinline__ void stats_next_epoch(DeviceStats* st) { st->epoch = st->epoch >= 0x3fffffffu ? 1u : st->epoch + 1u; }

// CUDA loads kernels lazily, and loading one can wait for the device to go idle.  A frame of the sharded grid has kernels
// that SPIN until another shard has made progress (hvcf_peer.cu); if that other shard lives in the same process and its next
// kernel has never been launched before, the load waits for the spinning kernel and the spinning kernel waits for the load.
// So every kernel a frame can launch is loaded up front, when peers are attached (cudaFuncGetAttributes forces the load).
#define HVCF_PRELOAD(k)                                                      \
	do                                                                       \
	{                                                                        \
		cudaFuncAttributes a_;                                               \
		cudaError_t e_ = cudaFuncGetAttributes(&a_, reinterpret_cast<const void*>(k)); \
		if (e_ != cudaSuccess)                                               \
			return e_;                                                       \
	} while (0)

} // namespace hvcf
