// hvcf_kernels.cu -- hand-written sm_100a kernels of the clustered-forward light-culling path.
//
//   aabb_generate_kernel    AABBGenerator.comp:39-90            one thread per cluster
//   light_transform_kernel  LightCulling.comp:51-53 hoisted      AoS LightData -> view-space culling layout
//   cull_kernel             LightCulling.comp:26-47,56,59-77     TMA-staged light tiles, warp-ballot compaction
//   count/scan/scatter      LightCulling.comp:79-87              deterministic offsets (no global atomic append)
//
// (paths relative to /root/reference/HelloVulkan/Shaders/ClusteredForward)
#include "hvcf_kernels.cuh"

namespace hvcf
{

// ======================================================================================================
// AABB generation.  One thread per cluster, two 128-bit stores.  The Z+1 slice-plane depths
// (-near * pow(far/near, z/Z), AABBGenerator.comp:72-75) come from a host-built table: only Z+1
// distinct values exist and GLSL pow() has no exact definition, so the table (host libm powf, the
// same call the oracle makes) is what makes this kernel bit-reproducible.
// ======================================================================================================
__device__ __forceinline__ void screen_to_view(const float* __restrict__ inv, float sx, float sy, float out[3])
{
	// AABBGenerator.comp:18-32 with screen.z = 0, screen.w = 1
	const float cx = __fadd_rn(__fmul_rn(sx, 2.0f), -1.0f);
	const float cy = __fadd_rn(__fmul_rn(sy, 2.0f), -1.0f);
	const float vx = mat_row_dot(inv, 0, cx, cy, 0.0f, 1.0f);
	const float vy = mat_row_dot(inv, 1, cx, cy, 0.0f, 1.0f);
	const float vz = mat_row_dot(inv, 2, cx, cy, 0.0f, 1.0f);
	const float vw = mat_row_dot(inv, 3, cx, cy, 0.0f, 1.0f);
	out[0] = __fdiv_rn(vx, vw);
	out[1] = __fdiv_rn(vy, vw);
	out[2] = __fdiv_rn(vz, vw);
}

__global__ void __launch_bounds__(128) aabb_generate_kernel(const hvcf_cluster_forward_ubo* __restrict__ ubo,
	const float* __restrict__ zplanes, hvcf_aabb* __restrict__ out, uint32_t cluster_count)
{
	const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
	if (tile >= cluster_count)
		return;
	const uint32_t X = ubo->slice_count_x, Y = ubo->slice_count_y;
	const uint32_t x = tile % X, y = (tile / X) % Y, z = tile / (X * Y);
	const float* inv = ubo->camera_inverse_projection;

	const float normX = __fdiv_rn(1.0f, static_cast<float>(X));
	const float normY = __fdiv_rn(1.0f, static_cast<float>(Y));
	float maxVS[3], minVS[3];
	screen_to_view(inv, __fmul_rn(__fadd_rn(static_cast<float>(x), 1.0f), normX),
		__fmul_rn(__fadd_rn(static_cast<float>(y), 1.0f), normY), maxVS);
	screen_to_view(inv, __fmul_rn(static_cast<float>(x), normX), __fmul_rn(static_cast<float>(y), normY), minVS);

	const float tileNear = zplanes[z];
	const float tileFar = zplanes[z + 1];

	float lo[3], hi[3];
#pragma unroll
	for (int i = 0; i < 3; ++i)
	{
		// RayIntersectZPlane: incident * zOffset / incident.z, left to right (AABBGenerator.comp:34-37)
		const float minNear = __fdiv_rn(__fmul_rn(minVS[i], tileNear), minVS[2]);
		const float minFar = __fdiv_rn(__fmul_rn(minVS[i], tileFar), minVS[2]);
		const float maxNear = __fdiv_rn(__fmul_rn(maxVS[i], tileNear), maxVS[2]);
		const float maxFar = __fdiv_rn(__fmul_rn(maxVS[i], tileFar), maxVS[2]);
		lo[i] = glsl_min(glsl_min(minNear, minFar), glsl_min(maxNear, maxFar));
		hi[i] = glsl_max(glsl_max(minNear, minFar), glsl_max(maxNear, maxFar));
	}
	float4* o = reinterpret_cast<float4*>(out + tile);
	o[0] = make_float4(lo[0], lo[1], lo[2], 0.0f);
	o[1] = make_float4(hi[0], hi[1], hi[2], 0.0f);
}

// ======================================================================================================
// Light transform: posView = cameraView * position (glm association order), r2 = radius * radius.
// One thread per lane-pair of the culling layout (hvcf_device.cuh): two 48 B records in, two float4 out
// (fully coalesced 128-bit stores).  HBM-bound: 64 B per light.
// ======================================================================================================
template <bool ANIM>
__global__ void __launch_bounds__(256) light_transform_kernel(const hvcf_cluster_forward_ubo* __restrict__ ubo,
	const hvcf_light_data* lights, uint32_t num_lights, uint32_t num_pairs, float4* __restrict__ out,
	uint32_t* __restrict__ zero, uint32_t zero_words, DeviceStats* __restrict__ stats, const LightAnim anim)
{
	grid_dep_launch();
	const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	// per-frame scratch of the hierarchical mode (early-exit counters) and the work counters are cleared here
	// instead of by memset nodes, so that transform -> pass 1 -> finalize -> pass 2 is an unbroken PDL chain
	for (uint32_t i = p; i < zero_words; i += gridDim.x * blockDim.x)
		zero[i] = 0u;
	if (p == 0 && stats)
	{
		stats->executed_tests = 0ull;
		stats->prefilter_tests = 0ull;
		stats->total_indices = 0u;
		stats->overflow_clusters = 0u;
		stats->bin_tests = 0ull;
		stats->write_tests = 0ull;
		stats_next_epoch(stats);
	}
	if (p >= num_pairs)
		return;
	const uint32_t g = p >> 5, l = p & 31;
	const float* view = ubo->camera_view;
	float x[2], y[2], z[2], r2[2];
#pragma unroll
	for (int h = 0; h < 2; ++h)
	{
		const uint32_t i = g * kGroupLights + h * 32 + l;
		if (i < num_lights)
		{
			float4 pw = *reinterpret_cast<const float4*>(lights[i].position);
			const float radius = lights[i].radius;
			if (ANIM)
			{
				pw.y = __fadd_rn(__ldg(anim.base_y + i), __fmul_rn(anim.amplitude, __ldg(anim.table + ((anim.step + i) & anim.mask))));
				anim.lights_rw[i].position[1] = pw.y;
			}
			x[h] = mat_row_dot(view, 0, pw.x, pw.y, pw.z, pw.w);
			y[h] = mat_row_dot(view, 1, pw.x, pw.y, pw.z, pw.w);
			z[h] = mat_row_dot(view, 2, pw.x, pw.y, pw.z, pw.w);
			r2[h] = __fmul_rn(radius, radius);
		}
		else
		{
			x[h] = y[h] = z[h] = 0.0f;
			r2[h] = -1.0f; // sqDist >= 0 is never < -1: pad lights cannot hit
		}
	}
	float4* grp = out + static_cast<size_t>(g) * (kGroupFloats / 4);
	grp[l] = make_float4(x[0], x[1], y[0], y[1]);
	grp[32 + l] = make_float4(z[0], z[1], r2[0], r2[1]);
}

// ======================================================================================================
// Light culling.
//
// grid.x = cluster groups (WARPS * K consecutive clusters), grid.y = light segments.  Every warp owns K
// clusters (boxes live in registers, warp-uniform) and walks the CTA's light tiles in ascending light
// order; lane l tests the light pair (64g + l, 64g + 32 + l) of each 64-light group against each box.
// Tiles (16 KiB) are staged global->shared by TMA bulk copies into a 3-deep ring, completion on
// mbarriers; ring slots are recycled by the per-tile __syncthreads_and that also carries the
// early-exit vote (a CTA stops when every one of its clusters already holds `cap` hits, which is the
// reference's `lightIter < MAX_LIGHTS_PER_CLUSTER` loop bound, LightCulling.comp:71).
//
// The test (LightCulling.comp:26-47,56) per axis is d = max3(lo - v, v - hi, 0); sq += d*d.  With lo <= hi
// this is bit-identical to the shader's branches: v < lo -> d = lo - v (the shader's diff); v > hi ->
// d = v - hi = -(hi - v), and the sign vanishes in d*d; otherwise d = 0 and sq + 0*0 == sq exactly.
// PACKED runs the subtractions and squares on FADD2/FMUL2 (two lights per instruction).
//
// Hits are compacted with __ballot_sync/__popc: positions are ascending in light index by
// construction, and only the first `cap` are stored.
// ======================================================================================================
template <int K>
struct Boxes
{
	float lo[K][3];
	float hi[K][3];
};

template <int K, bool PACKED, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) cull_kernel(const CullParams p)
{
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ __align__(8) uint64_t full_bar[kStages];

	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t seg = blockIdx.y;
	const uint32_t t_begin = seg * p.tiles_per_seg;
	const uint32_t t_end = min(t_begin + p.tiles_per_seg, p.tiles_total);
	const uint32_t ntiles = t_end - t_begin;
	const unsigned char* src = reinterpret_cast<const unsigned char*>(p.lights_view) + static_cast<size_t>(t_begin) * kTileBytes;

	if (threadIdx.x == 0)
	{
#pragma unroll
		for (uint32_t s = 0; s < kStages; ++s)
			mbar_init(&full_bar[s], 1);
		mbar_fence_init();
	}
	__syncthreads();
	if (threadIdx.x == 0)
	{
		for (uint32_t s = 0; s < kStages - 1 && s < ntiles; ++s)
		{
			mbar_arrive_expect_tx(&full_bar[s], kTileBytes);
			bulk_copy_g2s(smem + s * kTileBytes, src + static_cast<size_t>(s) * kTileBytes, kTileBytes, &full_bar[s]);
		}
	}

	// this warp's K boxes
	const uint32_t c0 = (blockIdx.x * WARPS + warp) * K;
	Boxes<K> bx;
	uint32_t cnt[K];
	uint32_t nvalid = 0;
#pragma unroll
	for (int k = 0; k < K; ++k)
	{
		cnt[k] = 0;
		if (c0 + k < p.cluster_count)
		{
			const float4* a = reinterpret_cast<const float4*>(p.aabbs + p.cluster_begin + c0 + k);
			const float4 mn = __ldg(a), mx = __ldg(a + 1);
			bx.lo[k][0] = mn.x, bx.lo[k][1] = mn.y, bx.lo[k][2] = mn.z;
			bx.hi[k][0] = mx.x, bx.hi[k][1] = mx.y, bx.hi[k][2] = mx.z;
			++nvalid;
		}
		else
		{
			// inverted infinite box: d = +inf on every axis, never a hit; results are discarded anyway
			bx.lo[k][0] = bx.lo[k][1] = bx.lo[k][2] = __int_as_float(0x7f800000);
			bx.hi[k][0] = bx.hi[k][1] = bx.hi[k][2] = __int_as_float(0xff800000);
			cnt[k] = p.cap;
		}
	}
	f32x2 plo[PACKED ? K : 1][3], phi[PACKED ? K : 1][3];
	if (PACKED)
	{
#pragma unroll
		for (int k = 0; k < K; ++k)
#pragma unroll
			for (int i = 0; i < 3; ++i)
			{
				plo[k][i] = pack2(bx.lo[k][i], bx.lo[k][i]);
				phi[k][i] = pack2(bx.hi[k][i], bx.hi[k][i]);
			}
	}
	uint32_t* list0 = p.seg_lists + (static_cast<size_t>(seg) * p.cluster_count + c0) * p.cap;
	const uint32_t lt = lanemask_lt();
	const uint32_t cap = p.cap;
	bool warp_done = (nvalid == 0);
	uint32_t groups_done = 0;

	uint32_t it = 0;
	for (; it < ntiles; ++it)
	{
		// refill the slot released by the barrier at the end of the previous iteration
		if (threadIdx.x == 0 && it + kStages - 1 < ntiles)
		{
			const uint32_t j = it + kStages - 1, s = j % kStages;
			mbar_arrive_expect_tx(&full_bar[s], kTileBytes);
			bulk_copy_g2s(smem + s * kTileBytes, src + static_cast<size_t>(j) * kTileBytes, kTileBytes, &full_bar[s]);
		}
		const uint32_t buf = it % kStages;
		mbar_wait(&full_bar[buf], (it / kStages) & 1u);

		if (!warp_done)
		{
			const float4* tile = reinterpret_cast<const float4*>(smem + buf * kTileBytes);
			const uint32_t tile_base = (t_begin + it) * kTileLights;
#pragma unroll 2
			for (uint32_t g = 0; g < kTileGroups; ++g)
			{
				const float4 a = tile[g * 64 + lane];      // x0 x1 y0 y1
				const float4 b = tile[g * 64 + 32 + lane]; // z0 z1 r0 r1
				const uint32_t base = tile_base + g * kGroupLights;
				f32x2 X = 0, Y = 0, Z = 0;
				if (PACKED)
				{
					X = pack2(a.x, a.y);
					Y = pack2(a.z, a.w);
					Z = pack2(b.x, b.y);
				}
#pragma unroll
				for (int k = 0; k < K; ++k)
				{
					bool h0, h1;
					if (PACKED)
					{
						float l0, l1, u0, u1, d0, d1;
						unpack2(sub2(plo[k][0], X), l0, l1);
						unpack2(sub2(X, phi[k][0]), u0, u1);
						d0 = fmaxf(fmaxf(l0, u0), 0.0f), d1 = fmaxf(fmaxf(l1, u1), 0.0f);
						const f32x2 dx = pack2(d0, d1);
						unpack2(sub2(plo[k][1], Y), l0, l1);
						unpack2(sub2(Y, phi[k][1]), u0, u1);
						d0 = fmaxf(fmaxf(l0, u0), 0.0f), d1 = fmaxf(fmaxf(l1, u1), 0.0f);
						const f32x2 dy = pack2(d0, d1);
						unpack2(sub2(plo[k][2], Z), l0, l1);
						unpack2(sub2(Z, phi[k][2]), u0, u1);
						d0 = fmaxf(fmaxf(l0, u0), 0.0f), d1 = fmaxf(fmaxf(l1, u1), 0.0f);
						const f32x2 dz = pack2(d0, d1);
						// products packed (FMUL2), sums scalar: ptxas contracts mul.rn.f32x2 + add.rn.f32x2
						// into FFMA2 even under --fmad=false, which would break bit-exactness; a scalar
						// add.rn.f32 is never contracted (tests/test_host.py::test_cull_kernels_have_no_contracted_fma checks the SASS for FFMA)
						float sx0, sx1, sy0, sy1, sz0, sz1;
						unpack2(mul2(dx, dx), sx0, sx1);
						unpack2(mul2(dy, dy), sy0, sy1);
						unpack2(mul2(dz, dz), sz0, sz1);
						const float s0 = __fadd_rn(__fadd_rn(sx0, sy0), sz0);
						const float s1 = __fadd_rn(__fadd_rn(sx1, sy1), sz1);
						h0 = s0 < b.z;
						h1 = s1 < b.w;
					}
					else
					{
						const float dx0 = fmaxf(fmaxf(__fsub_rn(bx.lo[k][0], a.x), __fsub_rn(a.x, bx.hi[k][0])), 0.0f);
						const float dx1 = fmaxf(fmaxf(__fsub_rn(bx.lo[k][0], a.y), __fsub_rn(a.y, bx.hi[k][0])), 0.0f);
						const float dy0 = fmaxf(fmaxf(__fsub_rn(bx.lo[k][1], a.z), __fsub_rn(a.z, bx.hi[k][1])), 0.0f);
						const float dy1 = fmaxf(fmaxf(__fsub_rn(bx.lo[k][1], a.w), __fsub_rn(a.w, bx.hi[k][1])), 0.0f);
						const float dz0 = fmaxf(fmaxf(__fsub_rn(bx.lo[k][2], b.x), __fsub_rn(b.x, bx.hi[k][2])), 0.0f);
						const float dz1 = fmaxf(fmaxf(__fsub_rn(bx.lo[k][2], b.y), __fsub_rn(b.y, bx.hi[k][2])), 0.0f);
						const float s0 = __fadd_rn(__fadd_rn(__fmul_rn(dx0, dx0), __fmul_rn(dy0, dy0)), __fmul_rn(dz0, dz0));
						const float s1 = __fadd_rn(__fadd_rn(__fmul_rn(dx1, dx1), __fmul_rn(dy1, dy1)), __fmul_rn(dz1, dz1));
						h0 = s0 < b.z;
						h1 = s1 < b.w;
					}
					if (__any_sync(0xffffffffu, h0 || h1))
					{
						const uint32_t b0 = __ballot_sync(0xffffffffu, h0);
						const uint32_t b1 = __ballot_sync(0xffffffffu, h1);
						const uint32_t n0 = __popc(b0);
						uint32_t* list = list0 + static_cast<size_t>(k) * cap;
						if (h0)
						{
							const uint32_t pos = cnt[k] + __popc(b0 & lt);
							if (pos < cap)
								list[pos] = base + lane;
						}
						if (h1)
						{
							const uint32_t pos = cnt[k] + n0 + __popc(b1 & lt);
							if (pos < cap)
								list[pos] = base + 32 + lane;
						}
						cnt[k] += n0 + __popc(b1);
					}
				}
			}
			groups_done += kTileGroups;
			bool done = true;
#pragma unroll
			for (int k = 0; k < K; ++k)
				done = done && (cnt[k] >= cap);
			warp_done = done;
		}
		// releases ring slot `buf` for the refill above and votes on the CTA-wide early exit
		if (__syncthreads_and(warp_done))
			break;
	}
	// bulk copies still in flight must land before this CTA's shared memory is handed on
	if (threadIdx.x == 0)
	{
		const uint32_t issued = min(it + kStages, ntiles); // tiles [0, issued) were requested
		for (uint32_t j = it + 1; j < issued; ++j)
			mbar_wait(&full_bar[j % kStages], (j / kStages) & 1u);
	}
	if (lane == 0 && nvalid)
	{
#pragma unroll
		for (int k = 0; k < K; ++k)
			if (c0 + k < p.cluster_count)
				p.seg_counts[static_cast<size_t>(seg) * p.cluster_count + c0 + k] = min(cnt[k], cap);
		atomicAdd(&p.stats->executed_tests, static_cast<unsigned long long>(groups_done) * kGroupLights * nvalid);
	}
}

// ======================================================================================================
// Offsets: count -> block sums -> scan -> scatter.  Replaces the reference's atomicAdd append
// (LightCulling.comp:79-84) with an exclusive scan in cluster order, so offsets are deterministic.
// ======================================================================================================
__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v, uint32_t lane)
{
#pragma unroll
	for (int d = 1; d < 32; d <<= 1)
	{
		const uint32_t n = __shfl_up_sync(0xffffffffu, v, d);
		if (lane >= static_cast<uint32_t>(d))
			v += n;
	}
	return v;
}

// exclusive scan over a 256-thread block; returns the exclusive prefix, *block_total = sum
__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* block_total)
{
	__shared__ uint32_t warp_sums[8];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t inc = warp_inclusive_scan(v, lane);
	if (lane == 31)
		warp_sums[warp] = inc;
	__syncthreads();
	uint32_t prefix = 0, total = 0;
#pragma unroll
	for (int w = 0; w < 8; ++w)
	{
		const uint32_t s = warp_sums[w];
		if (w < static_cast<int>(warp))
			prefix += s;
		total += s;
	}
	*block_total = total;
	return prefix + inc - v;
}

__global__ void __launch_bounds__(256) count_kernel(const uint32_t* __restrict__ seg_counts, uint32_t num_segs,
	uint32_t cluster_count, uint32_t cap, uint32_t* __restrict__ counts, uint32_t* __restrict__ block_sums,
	DeviceStats* __restrict__ stats)
{
	const uint32_t c = blockIdx.x * 256 + threadIdx.x;
	uint32_t v = 0;
	if (c < cluster_count)
	{
		for (uint32_t s = 0; s < num_segs; ++s)
			v += seg_counts[static_cast<size_t>(s) * cluster_count + c];
		v = min(v, cap);
		counts[c] = v;
	}
	uint32_t total;
	block_exclusive_scan_256(v, &total);
	const uint32_t over = __syncthreads_count(c < cluster_count && v >= cap);
	if (threadIdx.x == 0)
	{
		block_sums[blockIdx.x] = total;
		if (over)
			atomicAdd(&stats->overflow_clusters, over);
	}
}

// one block; exclusive scan of the per-block sums, total of the shard
__global__ void __launch_bounds__(256) scan_kernel(const uint32_t* __restrict__ block_sums, uint32_t num_blocks,
	uint32_t* __restrict__ block_offsets, uint32_t* __restrict__ shard_total, DeviceStats* __restrict__ stats)
{
	uint32_t carry = 0;
	for (uint32_t b0 = 0; b0 < num_blocks; b0 += 256)
	{
		const uint32_t i = b0 + threadIdx.x;
		const uint32_t v = i < num_blocks ? block_sums[i] : 0;
		uint32_t total;
		const uint32_t ex = block_exclusive_scan_256(v, &total);
		if (i < num_blocks)
			block_offsets[i] = carry + ex;
		carry += total;
		__syncthreads();
	}
	if (threadIdx.x == 0)
	{
		*shard_total = carry;
		stats->total_indices = carry;
	}
}

// grid = (ceil(C/256), 8): every CTA redoes the (cheap) 256-wide scan of its cluster block, then copies the
// lists of the 32 clusters blockIdx.y selects, 4 clusters per warp -- 8x the CTAs of a 1-D launch, so the
// dependent seg_counts -> seg_lists -> indices chains of different clusters overlap.
__global__ void __launch_bounds__(256) scatter_kernel(const ScatterParams p)
{
	__shared__ uint32_t s_off[256];
	__shared__ uint32_t s_cnt[256];
	const uint32_t c = blockIdx.x * 256 + threadIdx.x;
	const uint32_t v = c < p.cluster_count ? p.counts[c] : 0;
	uint32_t total;
	const uint32_t ex = block_exclusive_scan_256(v, &total);
	const uint32_t base = (p.index_base ? *p.index_base : 0u) + p.block_offsets[blockIdx.x];
	const uint32_t off = base + ex;
	s_off[threadIdx.x] = off;
	s_cnt[threadIdx.x] = v;
	if (blockIdx.y == 0)
	{
		if (c < p.cluster_count)
		{
			// LightCulling.comp:86-87
			*reinterpret_cast<uint2*>(p.cells + p.cluster_begin + c) = make_uint2(off, v);
		}
		if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 255 && p.frame_total)
			*p.frame_total = off + v; // globalIndex after the frame (LightCulling.comp:21,79)
	}
	__syncthreads();
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (uint32_t j = 0; j < 4; ++j)
	{
		const uint32_t t = blockIdx.y * 32 + warp * 4 + j;
		const uint32_t cc = blockIdx.x * 256 + t;
		if (cc >= p.cluster_count)
			break;
		const uint32_t n = s_cnt[t];
		uint32_t* dst = p.indices + s_off[t];
		uint32_t written = 0;
		for (uint32_t s = 0; s < p.num_segs && written < n; ++s)
		{
			const uint32_t m = min(p.seg_counts[static_cast<size_t>(s) * p.cluster_count + cc], n - written);
			const uint32_t* srcl = p.seg_lists + (static_cast<size_t>(s) * p.cluster_count + cc) * p.cap;
			for (uint32_t i = lane; i < m; i += 32)
				dst[written + i] = srcl[i];
			written += m;
		}
	}
}

// ======================================================================================================
// FP32 issue-rate microbenchmark: 8 independent mul/add chains per thread, explicit .rn (never
// contracted), enough warps to saturate every SMSP.  Used only to state the roofline denominator.
// ======================================================================================================
__global__ void __launch_bounds__(256) fp32_issue_rate_kernel(float* __restrict__ sink, uint32_t iters, float c, float d)
{
	float a[8];
#pragma unroll
	for (int i = 0; i < 8; ++i)
		a[i] = static_cast<float>(threadIdx.x + i);
	for (uint32_t it = 0; it < iters; ++it)
	{
#pragma unroll
		for (int i = 0; i < 8; ++i)
			a[i] = __fmul_rn(a[i], c);
#pragma unroll
		for (int i = 0; i < 8; ++i)
			a[i] = __fadd_rn(a[i], d);
	}
	float s = 0.0f;
#pragma unroll
	for (int i = 0; i < 8; ++i)
		s += a[i];
	if (s == 12345.678f)
		sink[0] = s;
}

cudaError_t launch_fp32_issue_rate(float* sink, uint32_t blocks, uint32_t iters, cudaStream_t stream)
{
	fp32_issue_rate_kernel<<<blocks, 256, 0, stream>>>(sink, iters, 0.999f, 0.001f);
	return cudaGetLastError();
}

// ======================================================================================================
// launchers
// ======================================================================================================
cudaError_t launch_aabb_generate(const hvcf_cluster_forward_ubo* ubo, const float* zplanes, hvcf_aabb* out,
	uint32_t cluster_count, cudaStream_t stream)
{
	if (cluster_count == 0)
		return cudaSuccess;
	aabb_generate_kernel<<<(cluster_count + 127) / 128, 128, 0, stream>>>(ubo, zplanes, out, cluster_count);
	return cudaGetLastError();
}

cudaError_t launch_light_transform(const hvcf_cluster_forward_ubo* ubo, const hvcf_light_data* lights,
	uint32_t num_lights, uint32_t padded_lights, float4* out, uint32_t* zero, uint32_t zero_words, DeviceStats* stats,
	cudaStream_t stream, const LightAnim* anim)
{
	const uint32_t pairs = padded_lights / 2;
	if (pairs == 0 && zero_words == 0 && !stats)
		return cudaSuccess;
	uint32_t blocks = pairs ? (pairs + 255) / 256 : 1;
	const uint32_t zero_blocks = (zero_words + 1023) / 1024 < 592u ? (zero_words + 1023) / 1024 : 592u; // a clear-only launch still spreads out
	if (blocks < zero_blocks)
		blocks = zero_blocks;
	if (anim)
		light_transform_kernel<true><<<blocks, 256, 0, stream>>>(ubo, lights, num_lights, pairs, out, zero, zero_words, stats, *anim);
	else
		light_transform_kernel<false><<<blocks, 256, 0, stream>>>(ubo, lights, num_lights, pairs, out, zero, zero_words, stats, LightAnim{});
	return cudaGetLastError();
}

__global__ void __launch_bounds__(256) capture_base_y_kernel(const hvcf_light_data* __restrict__ lights, uint32_t n, float* __restrict__ base_y)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n)
		base_y[i] = lights[i].position[1];
}

cudaError_t launch_capture_base_y(const hvcf_light_data* lights, uint32_t num_lights, float* base_y, cudaStream_t stream)
{
	if (num_lights == 0)
		return cudaSuccess;
	capture_base_y_kernel<<<(num_lights + 255) / 256, 256, 0, stream>>>(lights, num_lights, base_y);
	return cudaGetLastError();
}

// duplicates in `indices` resolve to the LAST occurrence, as a sequence of UpdateLightPosition calls would
__global__ void __launch_bounds__(256) scatter_positions_kernel(hvcf_light_data* __restrict__ lights, uint32_t num_lights,
	const uint32_t* __restrict__ indices, const float* __restrict__ xyz, uint32_t count, float* __restrict__ base_y)
{
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= count)
		return;
	const uint32_t i = indices[k];
	if (i >= num_lights)
		return; // ResourcesLight::UpdateLightPosition ignores an out-of-range index
	for (uint32_t j = k + 1; j < count && j < k + 64; ++j) // cheap local duplicate check; the API documents uniqueness beyond it
		if (indices[j] == i)
			return;
	*reinterpret_cast<float4*>(lights[i].position) = make_float4(xyz[3 * k], xyz[3 * k + 1], xyz[3 * k + 2], 1.0f);
	if (base_y)
		base_y[i] = xyz[3 * k + 1];
}

cudaError_t launch_scatter_positions(hvcf_light_data* lights, uint32_t num_lights, const uint32_t* indices, const float* xyz,
	uint32_t count, float* base_y, cudaStream_t stream)
{
	if (count == 0)
		return cudaSuccess;
	scatter_positions_kernel<<<(count + 255) / 256, 256, 0, stream>>>(lights, num_lights, indices, xyz, count, base_y);
	return cudaGetLastError();
}

template <int K, bool PACKED, int WARPS>
static cudaError_t launch_cull_t(const CullParams& p, uint32_t num_segs, cudaStream_t stream)
{
	static bool configured = false;
	const size_t smem_bytes = static_cast<size_t>(kStages) * kTileBytes;
	if (!configured)
	{
		cudaError_t e = cudaFuncSetAttribute(cull_kernel<K, PACKED, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
			static_cast<int>(smem_bytes));
		if (e != cudaSuccess)
			return e;
		configured = true;
	}
	const uint32_t per_cta = WARPS * K;
	dim3 grid((p.cluster_count + per_cta - 1) / per_cta, num_segs, 1);
	cull_kernel<K, PACKED, WARPS><<<grid, WARPS * 32, smem_bytes, stream>>>(p);
	return cudaGetLastError();
}

cudaError_t launch_cull(const CullParams& p, uint32_t num_segs, cudaStream_t stream)
{
	if (p.cluster_count == 0 || p.tiles_total == 0 || num_segs == 0)
		return cudaSuccess;
	// K = 2 clusters per warp on packed light pairs: the fastest of the six (K x packed / plain) shapes measured in round 1
	// (profiles/r01b); the others are gone from the build
	return launch_cull_t<2, true, 8>(p, num_segs, stream);
}

cudaError_t launch_count(const uint32_t* seg_counts, uint32_t num_segs, uint32_t cluster_count, uint32_t cap,
	uint32_t* counts, uint32_t* block_sums, DeviceStats* stats, cudaStream_t stream)
{
	if (cluster_count == 0)
		return cudaSuccess;
	count_kernel<<<(cluster_count + 255) / 256, 256, 0, stream>>>(seg_counts, num_segs, cluster_count, cap, counts,
		block_sums, stats);
	return cudaGetLastError();
}

cudaError_t launch_scan(const uint32_t* block_sums, uint32_t num_blocks, uint32_t* block_offsets,
	uint32_t* shard_total, DeviceStats* stats, cudaStream_t stream)
{
	scan_kernel<<<1, 256, 0, stream>>>(block_sums, num_blocks, block_offsets, shard_total, stats);
	return cudaGetLastError();
}

cudaError_t launch_scatter(const ScatterParams& p, cudaStream_t stream)
{
	if (p.cluster_count == 0)
		return cudaSuccess;
	scatter_kernel<<<dim3((p.cluster_count + 255) / 256, 8, 1), 256, 0, stream>>>(p);
	return cudaGetLastError();
}


// ======================================================================================================
// Grid download into mapped pinned host memory (hvcf_download_grid): the index count is only known on the device, so a
// cudaMemcpy of the used part of the index buffer needs a host round trip (copy the count, synchronise, copy the indices).
// This kernel reads the count itself and stores cells, indices[0, total) and the count across PCIe with 16 B stores:
// one stream synchronisation per frame.  At most max_words indices go this way (SM stores across PCIe are a little slower than
// the copy engine on long transfers: 32x18x48 x 1M lights, 12.5 MB of indices, 0.956 against 0.928 ms per step end to end); the
// host copies the rest once it knows the count.  Nothing is copied when the indices do not fit; the host reports the error.
// ======================================================================================================
__global__ void __launch_bounds__(256) grid_download_kernel(const uint4* __restrict__ cells, uint32_t cell_quads, uint32_t cell_tail_words,
	const uint32_t* __restrict__ indices, const uint32_t* __restrict__ total_ptr, uint32_t index_capacity, uint32_t max_words, uint4* __restrict__ cells_dst,
	uint32_t* __restrict__ indices_dst, uint32_t* __restrict__ total_dst)
{
	const uint32_t total = *total_ptr;
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
	if (tid == 0)
		*total_dst = total;
	for (uint32_t i = tid; i < cell_quads; i += nth)
		cells_dst[i] = cells[i];
	if (tid < cell_tail_words)
		reinterpret_cast<uint32_t*>(cells_dst + cell_quads)[tid] = reinterpret_cast<const uint32_t*>(cells + cell_quads)[tid];
	if (total > index_capacity)
		return;
	const bool all = total <= max_words; // max_words is a multiple of 4
	const uint32_t quads = (all ? total : max_words) / 4;
	const uint4* src = reinterpret_cast<const uint4*>(indices);
	uint4* dst = reinterpret_cast<uint4*>(indices_dst);
	uint32_t i = tid;
	for (; i + 3 * nth < quads; i += 4 * nth)
	{
		uint4 v[4];
#pragma unroll
		for (int u = 0; u < 4; ++u)
			v[u] = src[i + u * nth];
#pragma unroll
		for (int u = 0; u < 4; ++u)
			dst[i + u * nth] = v[u];
	}
	for (; i < quads; i += nth)
		dst[i] = src[i];
	if (all && tid < total % 4)
		indices_dst[quads * 4 + tid] = indices[quads * 4 + tid];
}

cudaError_t launch_grid_download(const void* cells, size_t cells_bytes, const uint32_t* indices, const uint32_t* total_ptr,
	uint32_t index_capacity, uint32_t max_words, void* cells_dst, uint32_t* indices_dst, uint32_t* total_dst, uint32_t blocks, cudaStream_t stream)
{
	grid_download_kernel<<<blocks, 256, 0, stream>>>(static_cast<const uint4*>(cells), static_cast<uint32_t>(cells_bytes / 16),
		static_cast<uint32_t>(cells_bytes % 16 / 4), indices, total_ptr, index_capacity, max_words & ~3u, static_cast<uint4*>(cells_dst), indices_dst, total_dst);
	return cudaGetLastError();
}

// every kernel of this file that a frame can launch, force-loaded (HVCF_PRELOAD, hvcf_device.cuh)

cudaError_t preload_basic_kernels()
{
	HVCF_PRELOAD(grid_download_kernel);
	HVCF_PRELOAD(aabb_generate_kernel);
	HVCF_PRELOAD(light_transform_kernel<false>);
	HVCF_PRELOAD(light_transform_kernel<true>);
	HVCF_PRELOAD((cull_kernel<2, true, 8>));
	HVCF_PRELOAD(count_kernel);
	HVCF_PRELOAD(scan_kernel);
	HVCF_PRELOAD(scatter_kernel);
	HVCF_PRELOAD(capture_base_y_kernel);
	HVCF_PRELOAD(scatter_positions_kernel);
	return cudaSuccess;
}

} // namespace hvcf
