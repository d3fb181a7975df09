// hvcf_small.cu -- the whole culling frame as ONE kernel, for frames of a few thousand lights.
//
// The reference's own scene is 1,000 lights on 16x9x24 clusters (AppPBRClusterForward.cpp:81-118, BASELINE.json configs[0]) and
// its frame is one dispatch (PipelineLightCulling.cpp:74-77).  At that size the hierarchical frame -- transform + binning, count
// pass, finalize, write pass: four dependent launches -- is launch latency and little else (0.025 ms for about 1.5 us of
// arithmetic).  Here one kernel does LightCulling.comp:49-88 for the whole grid:
//
//   * a CTA owns 32 consecutive clusters (lane = cluster) and transforms ALL the frame's lights into shared memory itself
//     (LightCulling.comp:51-53 hoisted; the arithmetic of light_transform_kernel -- cheaper than a kernel boundary);
//   * warp w takes the w-th 32nd of the lights, in light order: every lane tests its cluster's box against each of them
//     (sphere_box_hit: the reference's uncontracted float sequence) and keeps the hits as a BITMASK in registers;
//   * counts per (warp, lane) -> exclusive prefix over the warps = the rank of a warp's first hit in its cluster's list; the
//     cluster's count is the sum, capped (LightCulling.comp:71-77: the first `cap` hits in ascending light index);
//   * offsets = exclusive scan of the capped counts in cluster order (deterministic, as everywhere in this library): a
//     decoupled look-back over the CTAs inside the same kernel.  The state words are tagged with a serial that lives on the
//     device and is bumped by the last CTA to leave, so the kernel has no per-frame argument and replays inside a CUDA graph;
//   * every lane then walks its hit bits and stores the light indices at offset + rank while rank < cap, and writes its
//     light cell (LightCulling.comp:81-87).  No second pass over the lights, no scratch lists.
//
// Launched WITHOUT the programmatic-dependent-launch attribute: the serial is read at the start of every CTA and must not be
// read before the previous frame's last CTA has bumped it.
#include "hvcf_hier.cuh"

namespace hvcf
{

constexpr uint32_t kSmallWarps = 32;                                         // 1,024 threads: the frame is ~108 CTAs, so a CTA has an SM to itself
                                                                            // (8 warps: 0.0196 ms per A1k frame; the test phase was 125 serial trips per warp)
constexpr uint32_t kSmallMaxLights = 2048;                                  // 32 KiB of view records in shared memory
constexpr uint32_t kSmallWords = kSmallMaxLights / (kSmallWarps * 32);      // hit-mask words per lane: 2

__global__ void __launch_bounds__(kSmallWarps * 32) small_frame_kernel(const SmallParams p)
{
	__shared__ __align__(16) float4 s_light[kSmallMaxLights];
	__shared__ uint32_t s_cnt[kSmallWarps][32];
	__shared__ uint32_t s_off[32];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t c = blockIdx.x * 32 + lane; // cluster
	const bool valid = c < p.cluster_count;
	const uint32_t serial = *reinterpret_cast<volatile uint32_t*>(p.serial);
	// debug (HVCF_HIER_TRACE=1): globaltimer at the phase boundaries of the first 36 CTAs, 8 words each
	unsigned long long* const tr = (p.trace && blockIdx.x < 36 && threadIdx.x == 0) ? p.trace + 8ull * blockIdx.x : nullptr;
	auto stamp = [&](int i) {
		if (tr)
		{
			unsigned long long t;
			asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
			tr[i] = t;
		}
	};
	stamp(0);

	// ---- every light to view space (LightCulling.comp:51-53, 56) ----
	{
		float view[16];
#pragma unroll
		for (int i = 0; i < 16; ++i)
			view[i] = __ldg(&p.ubo->camera_view[i]);
		// stored as PAIRS of lights, (x0 x1 y0 y1)(z0 z1 r0 r1): the tests below run two lights per instruction (FADD2 / FMUL2)
		float* const f = reinterpret_cast<float*>(s_light);
		const uint32_t padded = (p.num_lights + 1u) & ~1u;
		for (uint32_t i = threadIdx.x; i < padded; i += blockDim.x)
		{
			float x = 0.0f, y = 0.0f, z = 0.0f, r2 = -1.0f; // the odd one out: never a hit
			if (i < p.num_lights)
			{
				const float4 pw = __ldg(reinterpret_cast<const float4*>(p.lights[i].position));
				const float radius = __ldg(&p.lights[i].radius);
				x = mat_row_dot(view, 0, pw.x, pw.y, pw.z, pw.w);
				y = mat_row_dot(view, 1, pw.x, pw.y, pw.z, pw.w);
				z = mat_row_dot(view, 2, pw.x, pw.y, pw.z, pw.w);
				r2 = __fmul_rn(radius, radius);
			}
			float* o = f + (i >> 1) * 8 + (i & 1u);
			o[0] = x, o[2] = y, o[4] = z, o[6] = r2;
		}
	}
	float lo[3] = {0.0f, 0.0f, 0.0f}, hi[3] = {0.0f, 0.0f, 0.0f};
	if (valid)
	{
		const float4* a = reinterpret_cast<const float4*>(p.aabbs + c);
		const float4 mn = __ldg(a), mx = __ldg(a + 1);
		lo[0] = mn.x, lo[1] = mn.y, lo[2] = mn.z;
		hi[0] = mx.x, hi[1] = mx.y, hi[2] = mx.z;
	}
	__syncthreads();
	stamp(1);

	// ---- this warp's lights against the lane's cluster: hit bits in registers ----
	const uint32_t per = 2u * ((p.num_lights + 2u * kSmallWarps - 1) / (2u * kSmallWarps)); // lights per warp: even, <= 64
	const uint32_t l0 = min(warp * per, p.num_lights), l1 = min(l0 + per, p.num_lights);
	uint32_t bits[kSmallWords];
	uint32_t cnt = 0;
#pragma unroll
	for (uint32_t w = 0; w < kSmallWords; ++w)
	{
		uint32_t m = 0;
		const uint32_t b0 = l0 + w * 32;
		if (valid && b0 < l1)
		{
			const uint32_t np = (min(32u, l1 - b0) + 1u) >> 1; // pairs; l0 and b0 are even
#pragma unroll 4
			for (uint32_t j = 0; j < np; ++j)
			{
				const float4 a = s_light[b0 + 2 * j], b = s_light[b0 + 2 * j + 1]; // broadcast
				bool h0, h1;
				sphere_box_hit2(a, b, lo, hi, h0, h1);
				m |= (h0 ? 1u : 0u) << (2 * j) | (h1 ? 2u : 0u) << (2 * j);
			}
		}
		bits[w] = m;
		cnt += __popc(m);
	}
	s_cnt[warp][lane] = cnt;
	__syncthreads();
	stamp(2);
	uint32_t before = 0, total = 0; // hits of the warps before this one / of all warps, for this lane's cluster
#pragma unroll
	for (uint32_t w = 0; w < kSmallWarps; ++w)
	{
		const uint32_t v = s_cnt[w][lane];
		before += w < warp ? v : 0u;
		total += v;
	}
	const uint32_t capped = min(total, p.cap);

	// ---- offsets: exclusive scan over the clusters.  Inside the CTA by warp 0's shuffles; across CTAs a decoupled look-back in
	// which ALL 1,024 threads look at one predecessor each, so 1,024 CTAs back cost one round trip to L2 (the warp-wide version
	// of hier_finalize_kernel took four dependent round trips for the 108 CTAs of grid A: 4.4 of the kernel's 11 us) ----
	__shared__ uint32_t s_agg, s_prefix, s_found;
	__shared__ uint32_t s_wsum[kSmallWarps], s_winc[kSmallWarps];
	const unsigned long long tag = static_cast<unsigned long long>(serial) << 34;
	uint32_t inc = capped;
	if (warp == 0)
	{
#pragma unroll
		for (int d = 1; d < 32; d <<= 1)
		{
			const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
			if (lane >= static_cast<uint32_t>(d))
				inc += t;
		}
		if (lane == 31)
		{
			s_agg = inc;
			// state word: [serial:30 | flag:2 | value:32]; flag 1 = aggregate of this CTA, 2 = inclusive prefix
			st_release_u64(&p.scan_state[blockIdx.x], tag | ((blockIdx.x == 0 ? 2ull : 1ull) << 32) | inc);
		}
	}
	if (threadIdx.x == 0)
		s_prefix = 0u;
	__syncthreads();
	for (int base = static_cast<int>(blockIdx.x); base > 0; base -= static_cast<int>(kSmallWarps * 32))
	{
		const int i = base - 1 - static_cast<int>(threadIdx.x); // thread 0 looks at the nearest predecessor
		unsigned long long w;
		do
			w = i >= 0 ? ld_acquire_u64(&p.scan_state[i]) : (tag | (2ull << 32));
		while ((w >> 34) != serial || ((w >> 32) & 3ull) == 0);
		const uint32_t incl = __ballot_sync(0xffffffffu, ((w >> 32) & 3ull) == 2);
		const uint32_t upto = incl ? static_cast<uint32_t>(__ffs(incl)) - 1u : 31u; // the nearest inclusive prefix ends the walk
		const uint32_t sum = __reduce_add_sync(0xffffffffu, lane <= upto ? static_cast<uint32_t>(w) : 0u);
		if (lane == 0)
			s_wsum[warp] = sum, s_winc[warp] = incl ? 1u : 0u;
		__syncthreads();
		if (threadIdx.x == 0)
		{
			uint32_t acc = s_prefix, found = 0u;
			for (uint32_t ww = 0; ww < kSmallWarps && !found; ++ww)
			{
				acc += s_wsum[ww];
				found = s_winc[ww];
			}
			s_prefix = acc;
			s_found = found;
		}
		__syncthreads();
		if (s_found)
			break;
	}
	const uint32_t prefix = s_prefix;
	if (warp == 0)
	{
		const uint32_t agg = s_agg;
		if (lane == 0 && blockIdx.x != 0)
			st_release_u64(&p.scan_state[blockIdx.x], tag | (2ull << 32) | (prefix + agg));
		const uint32_t off = prefix + inc - capped;
		s_off[lane] = off;
		if (valid)
			*reinterpret_cast<uint2*>(p.cells + c) = make_uint2(off, capped); // LightCulling.comp:86-87
		const uint32_t over = __popc(__ballot_sync(0xffffffffu, valid && total >= p.cap));
		if (lane == 0)
		{
			p.over_part[blockIdx.x] = over; // summed by the last CTA to leave
			if (blockIdx.x == gridDim.x - 1)
			{
				const uint32_t all = prefix + agg;
				*p.frame_total = all; // globalIndex after the frame (LightCulling.comp:21,79)
				*p.shard_total = all;
				p.stats->total_indices = all;
			}
		}
	}
	__syncthreads();
	stamp(3);

	// ---- the hits, at their final positions ----
	if (valid && before < p.cap)
	{
		uint32_t* dst = p.indices + s_off[lane];
		uint32_t rank = before;
#pragma unroll
		for (uint32_t w = 0; w < kSmallWords; ++w)
		{
			uint32_t m = bits[w];
			while (m && rank < p.cap)
			{
				const uint32_t j = __ffs(m) - 1;
				m &= m - 1;
				dst[rank++] = l0 + w * 32 + j;
			}
		}
	}

	// ---- the last CTA to leave sums the per-CTA overflow counts, writes the frame's work counters (brute force: every pair is an
	// exact test) and bumps the serial, so that the next frame's tags differ; it also clears its own ticket counter ----
	__shared__ uint32_t s_last;
	__syncthreads();
	stamp(4);
	if (threadIdx.x == 0)
	{
		__threadfence();
		s_last = atomicAdd(p.done, 1u) == gridDim.x - 1 ? 1u : 0u;
	}
	__syncthreads();
	if (s_last)
	{
		__threadfence();
		uint32_t sum = 0;
		for (uint32_t i = threadIdx.x; i < gridDim.x; i += blockDim.x)
			sum += *reinterpret_cast<volatile uint32_t*>(&p.over_part[i]);
		sum = __reduce_add_sync(0xffffffffu, sum);
		if (lane == 0)
			s_cnt[0][warp] = sum;
		__syncthreads();
		if (threadIdx.x == 0)
		{
			uint32_t all = 0;
			for (uint32_t w = 0; w < kSmallWarps; ++w)
				all += s_cnt[0][w];
			DeviceStats* st = p.stats;
			st->overflow_clusters = all;
			st->executed_tests = static_cast<unsigned long long>(p.cluster_count) * p.num_lights;
			st->prefilter_tests = st->bin_tests = st->write_tests = 0ull;
			*p.done = 0u;
			*p.serial = serial >= 0x3fffffffu ? 1u : serial + 1u;
		}
	}
	stamp(5);
}

uint32_t small_frame_max_lights() { return kSmallMaxLights; }

cudaError_t launch_small_frame(const SmallParams& p, cudaStream_t stream)
{
	if (p.cluster_count == 0 || p.num_lights > kSmallMaxLights)
		return cudaErrorInvalidValue;
	small_frame_kernel<<<(p.cluster_count + 31) / 32, kSmallWarps * 32, 0, stream>>>(p);
	return cudaGetLastError();
}

cudaError_t preload_small_kernels()
{
	HVCF_PRELOAD(small_frame_kernel);
	return cudaSuccess;
}

} // namespace hvcf
