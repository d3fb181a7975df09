// hvcf_shard.cu -- the 2-D decomposition of the sharded grid: clusters sharded by z-slab AND lights owned in shards.
//
// The reference is single-GPU.  With clusters sharded by z-slab, a rank needs only the lights whose
// sphere touches its slab (LightCulling.comp:49-57 against the union of the slab's boxes).  So instead of delivering every
// light to every rank (a broadcast from one root, or an all-gather), rank r OWNS the LightData of a contiguous range of
// lights, and per frame
//   shard_route_kernel        transforms its own lights to view space ONCE (LightCulling.comp:51-53 hoisted), tests each
//                             against the N slab union boxes (N tests per light -- the level above the cluster groups of the
//                             hierarchy) and stores every survivor's 16 B view record + 16-bit index straight into the owning
//                             rank's INBOX over NVLink peer memory.  Survivors are compacted in light order, one route block
//                             (B tiles of 1,024 lights) per CTA, into a fixed-capacity region per (block, destination): no
//                             size exchange, no atomics, and regions in block order = ascending light index.  The last CTA
//                             raises the rank's "routed" flag on every peer (release, system scope).
//   shard_wait_routed_kernel  waits until every rank's share has arrived (bounded; a missing peer is an error code).
// The owner then runs level-0 binning on its inbox only (hier_bin2_kernel<true>), and the count / write passes walk the
// survivor slots as before; a light's index is (first light of its route block) + (16-bit index inside the block).
//
// Exactness: the slab box contains every cluster box of the slab and the test is the same monotone uncontracted float sequence
// (sphere_box_hit), so routing can only deliver more lights than a slab needs, never fewer (hvcf_hier.cu header).  Light order
// is never disturbed, so each cluster still gets the first <= cap hits in ascending light index (LightCulling.comp:71-77).
#include "hvcf_hier.cuh"

#include <cooperative_groups.h>

namespace hvcf
{

constexpr uint32_t kRouteWarps = 8;
constexpr unsigned long long kRouteTimeoutNs = 4000000000ull; // 4 s, as in hvcf_peer.cu

__device__ __forceinline__ unsigned long long shard_ld_acquire_sys(const unsigned long long* p)
{
	unsigned long long v;
	asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void shard_st_release_sys(unsigned long long* p, unsigned long long v)
{
	asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ======================================================================================================
// union box of every rank's cluster range; runs when the AABBs change.  One CTA per rank.
// layout as cb_union: { lo.xyz, 0, hi.xyz, 0 }; a range holding a NaN box passes everything.
// ======================================================================================================
__global__ void __launch_bounds__(256) slab_unions_kernel(const hvcf_aabb* __restrict__ aabbs, const uint32_t* __restrict__ ranges,
	float* __restrict__ slab_box)
{
	__shared__ float s_u[8][6];
	__shared__ int s_nan;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const float kInf = __int_as_float(0x7f800000);
	const uint32_t begin = ranges[2 * blockIdx.x], end = ranges[2 * blockIdx.x + 1];
	if (threadIdx.x == 0)
		s_nan = 0;
	__syncthreads();
	float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
	bool nan = false;
	for (uint32_t c = begin + threadIdx.x; c < end; c += blockDim.x)
	{
		const float4* a = reinterpret_cast<const float4*>(aabbs + c);
		const float4 mn = a[0], mx = a[1];
		nan = nan || !(mn.x == mn.x && mn.y == mn.y && mn.z == mn.z && mx.x == mx.x && mx.y == mx.y && mx.z == mx.z);
		lo[0] = fminf(lo[0], mn.x), lo[1] = fminf(lo[1], mn.y), lo[2] = fminf(lo[2], mn.z);
		hi[0] = fmaxf(hi[0], mx.x), hi[1] = fmaxf(hi[1], mx.y), hi[2] = fmaxf(hi[2], mx.z);
	}
	if (__any_sync(0xffffffffu, nan) && lane == 0)
		s_nan = 1;
#pragma unroll
	for (int i = 0; i < 3; ++i)
	{
		lo[i] = warp_min_f(lo[i]);
		hi[i] = warp_max_f(hi[i]);
	}
	if (lane == 0)
	{
		s_u[warp][0] = lo[0], s_u[warp][1] = lo[1], s_u[warp][2] = lo[2];
		s_u[warp][3] = hi[0], s_u[warp][4] = hi[1], s_u[warp][5] = hi[2];
	}
	__syncthreads();
	if (threadIdx.x < 3)
	{
		float l = kInf, h = -kInf;
		for (int w = 0; w < 8; ++w)
		{
			l = fminf(l, s_u[w][threadIdx.x]);
			h = fmaxf(h, s_u[w][3 + threadIdx.x]);
		}
		float* u = slab_box + static_cast<size_t>(blockIdx.x) * 8;
		u[threadIdx.x] = s_nan ? -kInf : l;
		u[4 + threadIdx.x] = s_nan ? kInf : h;
		if (threadIdx.x == 0)
			u[3] = u[7] = 0.0f;
	}
}

// ---- TMA bulk store shared -> (peer) global, bulk-group completion (SASS UBLKCP / UTMASTG family) -----------------------
__device__ __forceinline__ void bulk_store_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes)
{
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() // at most N of the most recent groups may still be READING shared memory
{
	asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// ======================================================================================================
// transform + slab test + route.  A route block of B tiles is a THREAD-BLOCK CLUSTER of B CTAs, one tile of 1,024 lights
// each, so every tile of the rank's share is in flight at once.  Inside a CTA, warp w owns the 128 consecutive lights
// [w * 128, w * 128 + 128) as four rounds of 32 (lane = light), so (CTA rank, warp, round, lane) order = light order.
// Per destination: the ballots ARE the survivor bitmask of the tile (32 words); warp totals -> CTA totals in shared memory
// -> cluster barrier -> every CTA reads the totals of the CTAs before it through distributed shared memory = its first
// position in the (block, destination) region.  The CTA's survivors for one destination are compacted into a shared-memory
// staging buffer and leave as ONE TMA bulk store (cp.async.bulk shared -> peer global) of count x 16 B, plus one of 128 B for
// the bitmask: the copy engine cuts them into full-size NVLink writes.  (A first version stored the records and a 16-bit
// index per record straight from registers, 16 + 2 B per lane: 0.25 ms per frame for 2 M lights at 8 GPUs whatever the
// parallelism -- NVLink packet-bound, profiles/r02g.)  The destination rebuilds a record's in-block light index from the
// bitmask: the j-th record of the region is the j-th set bit.
// ======================================================================================================
__global__ void __launch_bounds__(kRouteWarps * 32) shard_route_kernel(const RouteParams p)
{
	namespace cg = cooperative_groups;
	__shared__ __align__(128) float4 s_stage[2][kTileLights];
	__shared__ __align__(128) uint32_t s_mask[kMaxPeers][32];
	__shared__ float s_box[kMaxPeers * kSlabBoxes][6];
	__shared__ uint32_t s_wt[kRouteWarps][kMaxPeers];
	__shared__ uint32_t s_tot[kMaxPeers];   // this CTA's survivors per destination; read by the CTAs after it in the cluster
	__shared__ uint32_t s_cbase[kMaxPeers]; // this CTA's first position in the (block, destination) region
	__shared__ uint32_t s_btot[kMaxPeers];  // the whole block's survivors per destination
	cg::cluster_group cluster = cg::this_cluster();
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t k = cluster.block_rank();            // tile inside the route block
	const uint32_t blk = p.block0 + blockIdx.x / p.block_tiles;
	// the first kernel of the frame also clears the early-exit counters and the work counters (ResetGlobalIndex,
	// PipelineLightCulling.cpp:38-43, is part of that); its own test count is a product, stored rather than accumulated
	{
		const uint32_t nthreads = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
		for (uint32_t i = tid; i < p.zero_words; i += nthreads)
			p.zero[i] = 0u;
		if (tid == 0 && p.stats)
		{
			DeviceStats* st = p.stats;
			const uint32_t first = p.block0 * p.block_tiles * kTileLights;
			const unsigned long long mine = static_cast<unsigned long long>(p.light_end > first ? p.light_end - first : 0u) * p.world * p.boxes;
			if (p.phase == 0)
			{
				st->executed_tests = 0ull;
				st->total_indices = 0u;
				st->overflow_clusters = 0u;
				st->write_tests = 0ull;
				st->bin_tests = st->prefilter_tests = mine;
				stats_next_epoch(st);
			}
			else
			{
				atomicAdd(&st->bin_tests, mine);
				atomicAdd(&st->prefilter_tests, mine);
			}
		}
	}
	const uint32_t nbox = p.world * p.boxes;
	if (threadIdx.x < nbox * 6)
	{
		const uint32_t d = threadIdx.x / 6, c = threadIdx.x % 6;
		s_box[d][c] = __ldcg(&p.slab_box[d * 8 + (c < 3 ? c : c + 1)]); // (second phase: written into this rank's mailbox by the peers)
	}
	__syncthreads();
	// the view matrix goes to registers once: read through the pointer it was re-loaded for every light (the kernel also stores to
	// global memory, so the compiler may not keep it) -- 12 of the 14 global loads per light, profiles/r03k
	float view[16];
#pragma unroll
	for (int i = 0; i < 16; ++i)
		view[i] = __ldg(&p.ubo->camera_view[i]);
	const size_t slot = static_cast<size_t>(blk) * p.block_tiles * kTileLights;
	const uint32_t lt = lanemask_lt();
	const uint32_t light0 = (blk * p.block_tiles + k) * kTileLights + warp * 128u;
	float4 v[4];
	uint32_t hm[4];
	bool have[4];
#pragma unroll
	for (int it = 0; it < 4; ++it)
	{
		const uint32_t i = light0 + it * 32 + lane;
		hm[it] = 0u;
		v[it] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
		have[it] = i < p.light_end;
		if (have[it])
		{
			const float4 pw = __ldg(reinterpret_cast<const float4*>(p.lights[i].position));
			const float radius = __ldg(&p.lights[i].radius);
			v[it].x = mat_row_dot(view, 0, pw.x, pw.y, pw.z, pw.w); // LightCulling.comp:53
			v[it].y = mat_row_dot(view, 1, pw.x, pw.y, pw.z, pw.w);
			v[it].z = mat_row_dot(view, 2, pw.x, pw.y, pw.z, pw.w);
			v[it].w = __fmul_rn(radius, radius);                    // LightCulling.comp:56
		}
	}
	// box by box, each against the lane's four lights (the box is read from shared memory once, not once per light)
	for (uint32_t bi = 0; bi < nbox; ++bi)
	{
		const float lo[3] = {s_box[bi][0], s_box[bi][1], s_box[bi][2]}, hi[3] = {s_box[bi][3], s_box[bi][4], s_box[bi][5]};
		const uint32_t bit = 1u << (bi / p.boxes);
#pragma unroll
		for (int it = 0; it < 4; ++it)
			hm[it] |= (have[it] && sphere_box_hit(v[it].x, v[it].y, v[it].z, v[it].w, lo, hi)) ? bit : 0u;
	}
	uint32_t my_tot = 0u; // lane d: survivors of this warp for destination d
	for (uint32_t d = 0; d < p.world; ++d)
	{
		uint32_t t = 0u;
#pragma unroll
		for (int it = 0; it < 4; ++it)
		{
			const uint32_t b = __ballot_sync(0xffffffffu, (hm[it] >> d) & 1u);
			if (lane == 0)
				s_mask[d][warp * 4 + it] = b; // bit l of word (w * 4 + it) = light w * 128 + it * 32 + l of the tile
			t += __popc(b);
		}
		if (lane == d)
			my_tot = t;
	}
	if (lane < p.world)
		s_wt[warp][lane] = my_tot;
	__syncthreads();
	if (threadIdx.x < p.world)
	{
		uint32_t t = 0u;
		for (uint32_t w = 0; w < kRouteWarps; ++w)
			t += s_wt[w][threadIdx.x];
		s_tot[threadIdx.x] = t;
	}
	cluster.sync(); // every CTA's totals are in its shared memory
	uint32_t wbase = 0u; // lane d: first position of this warp inside the CTA's run for destination d
	if (lane < p.world)
	{
		for (uint32_t w = 0; w < warp; ++w)
			wbase += s_wt[w][lane];
		if (warp == 0)
		{
			uint32_t cb = 0u, all = 0u;
			for (uint32_t r = 0; r < p.block_tiles; ++r)
			{
				const uint32_t t = *cluster.map_shared_rank(&s_tot[lane], r);
				cb += r < k ? t : 0u;
				all += t;
			}
			s_cbase[lane] = cb;
			s_btot[lane] = all; // the block's records for this destination: none -> the owner never looks at the block's bitmasks
			if (k + 1 == p.block_tiles && all != 0u)
				p.cnt[lane][blk] = all; // the region's count (the owner cleared its counts after the last frame of this parity: zero is not sent)
		}
	}
	cluster.sync(); // every CTA has read what it needs from the others (and s_cbase is in place): from here on the CTAs are independent
	// ---- staging rounds: consecutive destinations are packed into the staging buffer while they fit (2,048 records; a tile's
	// survivors over ALL destinations usually do), one proxy fence + one barrier per round, then thread d hands destination d's
	// run to the copy engine.  Every destination that gets any record of the BLOCK also gets the tile's 128-byte bitmask (zeros
	// included: the owner sums the block's masks). ----
	float4* const stage = &s_stage[0][0];
	constexpr uint32_t kStageCap = 2u * kTileLights;
	uint32_t d0 = 0u;
	bool first_round = true;
	while (d0 < p.world)
	{
		uint32_t d1 = d0, fill = 0u;
		while (d1 < p.world && (d1 == d0 || fill + s_tot[d1] <= kStageCap))
			fill += s_tot[d1++];
		if (!first_round) // the copy engine must have read the previous round out of the staging buffer
		{
			if (threadIdx.x < p.world)
				bulk_wait_read<0>();
			__syncthreads();
		}
		first_round = false;
		uint32_t off = 0u; // first staging slot of destination d (CTA-uniform)
		for (uint32_t d = d0; d < d1; ++d)
		{
			const uint32_t tot = s_tot[d];
			if (tot != 0u)
			{
				uint32_t wb = off + __shfl_sync(0xffffffffu, wbase, d);
#pragma unroll
				for (int it = 0; it < 4; ++it)
				{
					const uint32_t b = s_mask[d][warp * 4 + it];
					if ((b >> lane) & 1u)
						stage[wb + __popc(b & lt)] = v[it];
					wb += __popc(b);
				}
			}
			off += tot;
		}
		fence_proxy_async_smem(); // generic-proxy writes to shared memory -> visible to the copy engine
		__syncthreads();
		if (threadIdx.x >= d0 && threadIdx.x < d1)
		{
			const uint32_t d = threadIdx.x;
			uint32_t o = 0u;
			for (uint32_t e = d0; e < d; ++e)
				o += s_tot[e];
			if (s_tot[d] != 0u)
				bulk_store_s2g(p.rec[d] + slot + s_cbase[d], stage + o, s_tot[d] * 16u);
			if (s_btot[d] != 0u) // (the second phase of a two-phase frame sends nothing at all to the slabs that are full)
				bulk_store_s2g(p.mask[d] + (static_cast<size_t>(blk) * p.block_tiles + k) * 32u, s_mask[d], 128u);
			bulk_commit();
		}
		d0 = d1;
	}
	// ---- the threads that issued bulk stores only wait until the copy engine has READ the staging buffer (the CTA may then
	// leave).  Nobody fences or counts here: the rank's flag is raised by the NEXT kernel of the stream (shard_wait_routed_kernel),
	// and a kernel boundary orders every store of this grid -- the bulk copies included -- before it.  (A first version waited for
	// the completion of its remote stores, fenced at system scope and counted itself off in every CTA: that tail, not the routing,
	// was most of the kernel's 0.3 ms for 8M lights, profiles/r02z.) ----
	if (threadIdx.x < p.world)
		bulk_wait_read<0>();
}

// one warp: lane r raises this rank's flag of the phase on rank r (what = 0 / 2: the route kernel of the phase, if the rank had
// anything to route, is the previous kernel of the stream) and waits for rank r's share of the lights, or (what = 1) only waits
// for rank r's live boxes
__global__ void __launch_bounds__(32) shard_wait_routed_kernel(const PeerParams p, uint32_t parity, uint32_t what)
{
	const uint32_t r = threadIdx.x;
	if (r < p.world)
	{
		PeerMailbox* mine = p.mail[p.rank];
		// the route kernel before this one in the stream has completed, and with it every store into the peers' inboxes: fence at
		// system scope, then tell rank r (release) that this rank's share of the phase has arrived
		if (what != 1u)
		{
			__threadfence_system();
			shard_st_release_sys(what == 0 ? &p.mail[r]->routed[parity][p.rank] : &p.mail[r]->routed2[parity][p.rank], static_cast<unsigned long long>(p.epoch) << 32);
		}
		const unsigned long long* word = what == 0 ? &mine->routed[parity][r] : (what == 1 ? &mine->alive_flag[parity][r] : &mine->routed2[parity][r]);
		unsigned long long t0;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
		for (;;)
		{
			const unsigned long long v = shard_ld_acquire_sys(word);
			if ((v >> 32) == p.epoch)
				break;
			unsigned long long t1;
			asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
			if (t1 - t0 > kRouteTimeoutNs)
			{
				atomicCAS(&mine->error, 0u, (3u << 28) | (r << 20) | (p.epoch & 0xfffffu)); // kind 3 = routed lights (or the live boxes between its two shares)
				break;
			}
			__nanosleep(200);
		}
	}
	__threadfence_system();
}

// ======================================================================================================
// Between the two phases of a routed frame: the rounds counted so far have filled most clusters (on the reference scene at 16M
// lights every cluster in front of slice 64 after 70,000 lights), and hier_alive_kernel has left, per group of this slab, whether
// any cluster is still below its cap and the union box of those that are.  This kernel folds the live groups' boxes into
// kSlabBoxes boxes -- by the side of the view axis their centre lies on: the clusters that stay alive longest sit at the left
// and right edge of the lit region, and one box around both would span everything between them -- and stores them into every
// rank's mailbox, followed by a release flag.  The rest of the frame's lights is routed against these boxes: a slab that is full
// receives nothing any more, and the others only the lights that can still matter (exactly conservative, as every box test here).
// One CTA.
// ======================================================================================================
__global__ void __launch_bounds__(256) slab_alive_kernel(const PeerParams p, uint32_t parity, const uint32_t* group_alive, const float* alive_union, uint32_t ncb)
{
	__shared__ float s_u[8][kSlabBoxes][6];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const float kInf = __int_as_float(0x7f800000);
	float lo[kSlabBoxes][3], hi[kSlabBoxes][3];
#pragma unroll
	for (uint32_t b = 0; b < kSlabBoxes; ++b)
		for (int i = 0; i < 3; ++i)
			lo[b][i] = kInf, hi[b][i] = -kInf;
	for (uint32_t g = threadIdx.x; g < ncb; g += blockDim.x)
	{
		if (__ldcg(&group_alive[g]) == 0u)
			continue;
		const float4* u = reinterpret_cast<const float4*>(alive_union + static_cast<size_t>(g) * 8);
		const float4 a = __ldcg(u), b4 = __ldcg(u + 1);
		// a group holding a NaN box is (-inf, +inf): its centre is NaN, the comparison false, it lands in box 1 and makes that box pass everything
		const uint32_t side = (a.x + b4.x) < 0.0f ? 0u : 1u;
#pragma unroll
		for (uint32_t b = 0; b < kSlabBoxes; ++b)
			if (b == side)
			{
				lo[b][0] = fminf(lo[b][0], a.x), lo[b][1] = fminf(lo[b][1], a.y), lo[b][2] = fminf(lo[b][2], a.z);
				hi[b][0] = fmaxf(hi[b][0], b4.x), hi[b][1] = fmaxf(hi[b][1], b4.y), hi[b][2] = fmaxf(hi[b][2], b4.z);
			}
	}
#pragma unroll
	for (uint32_t b = 0; b < kSlabBoxes; ++b)
		for (int i = 0; i < 3; ++i)
		{
			const float l = warp_min_f(lo[b][i]), h = warp_max_f(hi[b][i]);
			if (lane == 0)
				s_u[warp][b][i] = l, s_u[warp][b][3 + i] = h;
		}
	__syncthreads();
	if (threadIdx.x < p.world)
	{
		float* dst = &p.mail[threadIdx.x]->alive_box[parity][p.rank][0][0];
		for (uint32_t b = 0; b < kSlabBoxes; ++b)
		{
			float l[3] = {kInf, kInf, kInf}, h[3] = {-kInf, -kInf, -kInf};
			for (int w = 0; w < 8; ++w)
				for (int i = 0; i < 3; ++i)
					l[i] = fminf(l[i], s_u[w][b][i]), h[i] = fmaxf(h[i], s_u[w][b][3 + i]);
			float* o = dst + b * 8;
			o[0] = l[0], o[1] = l[1], o[2] = l[2], o[3] = 0.0f;
			o[4] = h[0], o[5] = h[1], o[6] = h[2], o[7] = 0.0f;
		}
		__threadfence_system();
		shard_st_release_sys(&p.mail[threadIdx.x]->alive_flag[parity][p.rank], static_cast<unsigned long long>(p.epoch) << 32);
	}
}

cudaError_t launch_slab_unions(const hvcf_aabb* aabbs, const uint32_t* d_ranges, uint32_t world, float* slab_box, cudaStream_t stream)
{
	if (world == 0)
		return cudaSuccess;
	slab_unions_kernel<<<world, 256, 0, stream>>>(aabbs, d_ranges, slab_box);
	return cudaGetLastError();
}

cudaError_t launch_shard_route(const RouteParams& p, cudaStream_t stream)
{
	if (p.blocks == 0 || p.world == 0 || p.boxes == 0 || p.boxes > kSlabBoxes || p.world > kMaxPeers || p.block_tiles == 0 || p.block_tiles > kMaxRouteBlockTiles ||
		(p.block_tiles & (p.block_tiles - 1u)))
		return cudaErrorInvalidValue;
	if (p.block_tiles > 8) // clusters of more than 8 CTAs are a per-kernel opt-in
	{
		static bool allowed = false;
		if (!allowed)
		{
			const cudaError_t e = cudaFuncSetAttribute(shard_route_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
			if (e != cudaSuccess)
				return e;
			allowed = true;
		}
	}
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = dim3(p.blocks * p.block_tiles, 1, 1);
	cfg.blockDim = dim3(kRouteWarps * 32, 1, 1);
	cfg.stream = stream;
	cudaLaunchAttribute attr{};
	attr.id = cudaLaunchAttributeClusterDimension;
	attr.val.clusterDim.x = p.block_tiles;
	attr.val.clusterDim.y = 1;
	attr.val.clusterDim.z = 1;
	cfg.attrs = &attr;
	cfg.numAttrs = 1;
	return cudaLaunchKernelEx(&cfg, shard_route_kernel, p);
}

cudaError_t launch_shard_wait_routed(const PeerParams& p, uint32_t parity, cudaStream_t stream, uint32_t what)
{
	shard_wait_routed_kernel<<<1, 32, 0, stream>>>(p, parity, what);
	return cudaGetLastError();
}

cudaError_t launch_slab_alive(const PeerParams& p, uint32_t parity, const uint32_t* group_alive, const float* alive_union, uint32_t ncb, cudaStream_t stream)
{
	slab_alive_kernel<<<1, 256, 0, stream>>>(p, parity, group_alive, alive_union, ncb);
	return cudaGetLastError();
}

cudaError_t preload_shard_kernels()
{
	HVCF_PRELOAD(slab_unions_kernel);
	HVCF_PRELOAD(shard_route_kernel);
	HVCF_PRELOAD(shard_wait_routed_kernel);
	HVCF_PRELOAD(slab_alive_kernel);
	return cudaSuccess;
}

} // namespace hvcf
