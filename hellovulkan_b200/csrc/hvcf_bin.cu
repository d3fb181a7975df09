// hvcf_bin.cu -- HVCF_MODE_HIERARCHICAL, second generation: level 0 of the cluster hierarchy is a streaming
// "bin" kernel of its own, and the count / write passes are barrier-free warp programs.
//
//   hier_unions_kernel   once per AABB generation: union box of every cluster group (8 x K blocks x 32 clusters)
//   hier_bin_kernel      per frame: each warp tests one TMA-staged light tile against ONE group's union box
//                        (lane = light pair, FADD2/FMUL2, ballot/popc) and writes the survivors, in light order,
//                        into that (group, tile) slot.  A slot holds kBinCap records; when more lights survive
//                        the filter is not selective there and the consumers read the raw tile instead.
//   hier2_kernel<K,W>    per frame, twice (count, then write): a WARP owns K blocks of 32 clusters (lane =
//                        cluster) for one light segment and walks the group's survivor slots tile by tile:
//                        level 1 (lane = surviving light vs the block union box) -> level 2 (lane = cluster,
//                        candidate light broadcast from a per-warp shared-memory slab).  No CTA barrier, no
//                        shared tile ring: warps only ever wait for their own loads.
//
// Exactness, ordering, early exit and the count -> finalize -> write scheme are those of hvcf_hier.cu (see its
// header): union boxes are exactly conservative because they run the same uncontracted float sequence, light
// order is never disturbed, and cum[cluster][chunk] lower bounds only remove work that cannot matter.
#include "hvcf_hier.cuh"

namespace hvcf
{

// Loads of the level-0 slots (bin_count / bin_rec / bin_idx, unit_any) in the count / write passes.  They were written by an
// EARLIER kernel of the same programmatic-dependent-launch chain and are read only after griddepcontrol.wait, i.e. after that
// kernel has completed and its stores are visible; a grid's L1 starts invalidated and no line of these buffers can be filled before
// the wait, so the read-only path (ld.global.nc, L1-cached) cannot return a stale line.  It matters: the four warp programs of a
// CTA belong to one group and read the same slot records (A65k 0.0446 ms per frame against 0.0511 with ld.global.cg).  The PTX
// wording ("read-only for the lifetime of the kernel") is stricter than that argument; -DHVCF_CHAIN_LDCG builds the L2-only loads.
#ifdef HVCF_CHAIN_LDCG
#define HVCF_CHAIN_LD(p) __ldcg(p)
#else
#define HVCF_CHAIN_LD(p) __ldg(p)
#endif

constexpr uint32_t kBinWarps = 8;  // hier_bin_kernel: groups per CTA (one per warp)
constexpr uint32_t kMainWarps = 4; // hier2_kernel: warp programs per CTA

// ======================================================================================================
// union box of every cluster group; runs when the AABBs change.  One CTA of 256 threads per group.
// layout: cb_union[group][8] = { lo.xyz, 0, hi.xyz, 0 }; a group holding a NaN box passes everything.
// ======================================================================================================
__global__ void __launch_bounds__(256) hier_unions_kernel(const uint32_t* __restrict__ lane_cluster, const hvcf_aabb* __restrict__ aabbs,
	uint32_t cluster_begin, uint32_t K, float* __restrict__ cb_union)
{
	__shared__ float s_u[8][6];
	__shared__ int s_nan;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const float kInf = __int_as_float(0x7f800000);
	if (threadIdx.x == 0)
		s_nan = 0;
	__syncthreads();
	float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
	bool nan = false;
	for (uint32_t k = 0; k < K; ++k)
	{
		const uint32_t c = lane_cluster[(static_cast<size_t>((blockIdx.x * 8 + warp) * K + k) << 5) + lane];
		if (c == kSkip)
			continue;
		const float4* a = reinterpret_cast<const float4*>(aabbs + cluster_begin + c);
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
		float* u = cb_union + static_cast<size_t>(blockIdx.x) * 8;
		u[threadIdx.x] = s_nan ? -kInf : l;
		u[4 + threadIdx.x] = s_nan ? kInf : h;
		if (threadIdx.x == 0)
			u[3] = u[7] = 0.0f;
	}
}

// ======================================================================================================
// Rounds (frames of many tiles; hvcf_api.cu plans them): the light stream is cut into a few rounds of growing length, and each
// round is binned and counted by launches of its own.  BETWEEN launches cum[cluster][*] is exact -- every count of the earlier
// rounds has been added -- so a cluster whose row sums to >= cap provably has its first `cap` hits among the lights of those
// rounds (LightCulling.comp:71-77), and a GROUP without any cluster below its cap needs no light of any later round.  This
// kernel publishes that per group, together with the union box of the group's LIVE clusters; the binning kernels of the next
// round leave the dead groups' slots empty and test the others against that tighter box.  It is the count pass's per-cluster early exit lifted to level 0:
// on the reference scene at 16M lights every cluster in front of slice 64 is full after the first 70,000 lights.
// One CTA per group (8 warp programs x K blocks x 32 lanes).
// ======================================================================================================
__global__ void __launch_bounds__(256) hier_alive_kernel(const uint32_t* __restrict__ lane_cluster, const hvcf_aabb* __restrict__ aabbs,
	uint32_t cluster_begin, const uint32_t* cum, uint32_t K, uint32_t cap, uint32_t* __restrict__ group_alive, float* __restrict__ alive_union)
{
	__shared__ float s_u[8][6];
	__shared__ int s_nan;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const float kInf = __int_as_float(0x7f800000);
	grid_dep_launch();
	if (threadIdx.x == 0)
		s_nan = 0;
	grid_dep_wait(); // cum: the count pass of the previous round
	float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
	bool alive = false, nan = false;
	for (uint32_t k = 0; k < K; ++k)
	{
		const uint32_t c = lane_cluster[(static_cast<size_t>((blockIdx.x * 8 + warp) * K + k) << 5) + lane];
		if (c == kSkip)
			continue;
		const uint4* row = reinterpret_cast<const uint4*>(cum + static_cast<size_t>(c) * kHMaxChunks);
		const uint4 a = __ldcg(row), b = __ldcg(row + 1);
		if ((a.x + a.y) + (a.z + a.w) + (b.x + b.y) + (b.z + b.w) >= cap)
			continue; // full: no later light can enter its list
		alive = true;
		const float4* bx = reinterpret_cast<const float4*>(aabbs + cluster_begin + c);
		const float4 mn = bx[0], mx = bx[1];
		nan = nan || !(mn.x == mn.x && mn.y == mn.y && mn.z == mn.z && mx.x == mx.x && mx.y == mx.y && mx.z == mx.z);
		lo[0] = fminf(lo[0], mn.x), lo[1] = fminf(lo[1], mn.y), lo[2] = fminf(lo[2], mn.z);
		hi[0] = fmaxf(hi[0], mx.x), hi[1] = fmaxf(hi[1], mx.y), hi[2] = fmaxf(hi[2], mx.z);
	}
	const int any = __syncthreads_or(alive ? 1 : 0);
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
	if (threadIdx.x == 0)
		group_alive[blockIdx.x] = any ? 1u : 0u;
	if (threadIdx.x < 3)
	{
		// the union box of the LIVE clusters only (layout and NaN rule of hier_unions_kernel): it contains every box that can still
		// take a light, so the level-0 test against it stays exactly conservative -- and it is much tighter than the group's box
		// where a group straddles the edge of the lit region (its full clusters are the ones inside)
		float l = kInf, h = -kInf;
		for (int w = 0; w < 8; ++w)
		{
			l = fminf(l, s_u[w][threadIdx.x]);
			h = fmaxf(h, s_u[w][3 + threadIdx.x]);
		}
		float* u = alive_union + static_cast<size_t>(blockIdx.x) * 8;
		u[threadIdx.x] = s_nan ? -kInf : l;
		u[4 + threadIdx.x] = s_nan ? kInf : h;
		if (threadIdx.x == 0)
			u[3] = u[7] = 0.0f;
	}
}

// ======================================================================================================
// level 0: grid = (ceil(ncb / 8), tiles).  The tile is staged once per CTA by a TMA bulk copy; warp w tests
// all of it against group cbg * 8 + w and appends survivors to the (group, tile) slot in light order.
// Measured and rejected on B200 (B1M, this kernel 57 us warm): a two-level variant (super-groups of 4 groups first:
// 15 + 4.8 tests per light instead of 60, half the instructions) ran in the same 57 us (two more CTA barriers, shared-
// memory gathers), and making its CTAs persistent over tiles with a two-deep TMA ring made it slower (65 us: 4 CTAs/SM).
// ======================================================================================================
__global__ void __launch_bounds__(kBinWarps * 32) hier_bin_kernel(const BinParams p)
{
	__shared__ __align__(128) unsigned char s_tile[kTileBytes];
	__shared__ __align__(8) uint64_t bar;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t tile_id = p.tile_begin + blockIdx.y, cb = blockIdx.x * kBinWarps + warp;
	grid_dep_launch();
	if (threadIdx.x == 0)
	{
		mbar_init(&bar, 1);
		mbar_fence_init();
	}
	__syncthreads();
	grid_dep_wait(); // the tile is written by the transform kernel of this frame
	// rounds: groups without a cluster below its cap keep an empty slot; a CTA whose groups are all dead does not stage the tile
	uint32_t alive_mask = 0xffu;
	if (p.group_alive)
	{
		const uint32_t g = blockIdx.x * kBinWarps + (lane & 7u);
		alive_mask = __ballot_sync(0xffffffffu, g < p.ncb && __ldcg(&p.group_alive[g]) != 0u) & 0xffu;
		if (((alive_mask >> warp) & 1u) == 0u && cb < p.ncb && lane == 0)
			p.bin_count[static_cast<size_t>(cb) * p.tiles_cap + tile_id] = 0u;
		if (alive_mask == 0u)
			return;
	}
	if (threadIdx.x == 0)
	{
		mbar_arrive_expect_tx(&bar, kTileBytes);
		bulk_copy_g2s(s_tile, reinterpret_cast<const unsigned char*>(p.lights_view) + static_cast<size_t>(tile_id) * kTileBytes, kTileBytes, &bar);
	}
	float lo[3] = {0.0f, 0.0f, 0.0f}, hi[3] = {0.0f, 0.0f, 0.0f};
	if (cb < p.ncb)
	{
		const float4* u = reinterpret_cast<const float4*>(p.cb_union + static_cast<size_t>(cb) * 8); // (rounds: written by hier_alive_kernel just before)
		const float4 a = p.group_alive ? __ldcg(u) : __ldg(u), b = p.group_alive ? __ldcg(u + 1) : __ldg(u + 1);
		lo[0] = a.x, lo[1] = a.y, lo[2] = a.z;
		hi[0] = b.x, hi[1] = b.y, hi[2] = b.z;
	}
	mbar_wait(&bar, 0);
	if (cb >= p.ncb || ((alive_mask >> warp) & 1u) == 0u)
		return;
	const float4* tile = reinterpret_cast<const float4*>(s_tile);
	const size_t slot = (static_cast<size_t>(cb) * p.tiles_cap + tile_id) * kBinCap;
	float4* const rec = p.bin_rec + slot;
	uint16_t* const idx = p.bin_idx + slot;
	const uint32_t lt = lanemask_lt();
	uint32_t cnt = 0;
#pragma unroll 4
	for (uint32_t g = 0; g < kTileGroups; ++g)
	{
		const float4 a = tile[g * 64 + lane], b = tile[g * 64 + 32 + lane];
		bool h0, h1;
		sphere_box_hit2(a, b, lo, hi, h0, h1);
		const uint32_t m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
		if ((m0 | m1) == 0)
			continue;
		const uint32_t d0 = cnt + __popc(m0 & lt), d1 = cnt + __popc(m0) + __popc(m1 & lt);
		if (h0 && d0 < kBinCap)
		{
			rec[d0] = make_float4(a.x, a.z, b.x, b.z);
			idx[d0] = static_cast<uint16_t>(g * 64 + lane);
		}
		if (h1 && d1 < kBinCap)
		{
			rec[d1] = make_float4(a.y, a.w, b.y, b.w);
			idx[d1] = static_cast<uint16_t>(g * 64 + 32 + lane);
		}
		cnt += __popc(m0) + __popc(m1);
	}
	if (lane == 0)
	{
		p.bin_count[static_cast<size_t>(cb) * p.tiles_cap + tile_id] = cnt;
		atomicAdd(&p.stats->prefilter_tests, static_cast<unsigned long long>(kTileLights));
		atomicAdd(&p.stats->bin_tests, static_cast<unsigned long long>(kTileLights));
	}
}

// ======================================================================================================
// levels 1 + 2, count pass (WRITE = false) and write pass (WRITE = true).
// grid = segments x (ncb * 8 / kMainWarps) CTAs of kMainWarps independent warp programs; warp program
// `unit` owns blocks unit * K .. unit * K + K - 1 (lane = cluster) and belongs to group unit / 8.
// ======================================================================================================
// Measured and rejected on B200 (r01y kernels, per-kernel events): one warp program serving 2 or 4 segments in a loop (stride
// gridDim.y) with the lane's cluster indices and boxes loaded once.  The loop-carried state costs registers (write pass K=2
// 64 -> 96, count pass K=2 spills 48 B) and the gain of the saved round trips does not pay for it: B1M count pass 97 -> 101 us,
// write pass 57 -> 77 us with 2 segments per program; A65k and A1M within +-1 us.
// Also rejected: exact "owns output" flags per (segment, warp program) set by hier_finalize_kernel instead of the count pass's
// "counted anything" word -- B1M write pass 57.3 -> 55.8 us but finalize 25.5 -> 28.6 us: almost every program that counted owns output.
// Also rejected (profiles/r01zh_l2_pairs_rejected.txt, "default" = this variant): level 2 on candidate PAIRS -- the batch's level-1
// survivors compacted in light order into (x0 x1 y0 y1)(z0 z1 r0 r1) records and tested two per trip with sphere_box_hit2
// (12.5 instead of 24 instructions per candidate).  Byte-identical (205 GPU tests), but a batch of 64 survivor records holds
// only ~5 level-2 candidates on average, so the compaction (scattered STS, per-block masks in compacted order) costs more than
// the halved tests save, and the loop-invariant (lo, lo) / (hi, hi) register pairs cost occupancy: A65k 0.0521 -> 0.0541 ms,
// A1M 0.130 -> 0.137, B1M 0.216 -> 0.237 (count pass 98 -> 115 us).
// Also rejected (profiles/r01zi_early_loads_rejected.txt, "default" = this variant): the lane's cluster indices and boxes loaded
// ABOVE griddepcontrol.wait (they are not written inside the frame's PDL chain; it takes ld.volatile to keep ptxas from sinking
// the loads below ACQBULK).  No gain at A65k / A1M, B1M 0.215 -> 0.219 ms: the warp programs that leave right after the wait
// (nothing counted, empty slots) now pay for loads they never use.
__device__ __forceinline__ size_t any_slot(uint32_t seg, uint32_t unit) { return static_cast<size_t>(seg) * (gridDim.x * kMainWarps) + unit; }

// 16-byte asynchronous copy global -> shared (SASS LDGSTS); src_bytes < 16 zero-fills the rest
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src, uint32_t src_bytes)
{
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src),
		"r"(src_bytes)
		: "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
	asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// PIPE: the walk over the survivor slots is software-pipelined (cp.async double buffer, see below); chosen by the host for frames
// with many tiles per light segment
// SUB: sub-tile light segments (p.parts > 1); a template parameter so that whole-tile frames run exactly the code they ran before
// INBOX: the tiles are inbox tiles of a routed frame (hvcf_shard.cu); a template parameter so that every other frame runs exactly the code it ran before
template <int K, bool WRITE, bool PIPE, bool SUB, bool INBOX = false>
__global__ void __launch_bounds__(kMainWarps * 32) hier2_kernel(const HierParams p)
{
	__shared__ __align__(16) float4 s_rec2[kMainWarps][PIPE ? 2 : 1][64];
	__shared__ __align__(16) uint16_t s_idx2[kMainWarps][PIPE ? 2 : 1][64];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t seg = p.seg_begin + blockIdx.y;
	const uint32_t unit = blockIdx.x * kMainWarps + warp;
	const uint32_t cb = unit >> 3;
	// sub-tile segments (frames of few tiles): segment = (tile, part); part q of P takes the batches [q nb / P, (q + 1) nb / P) of
	// the tile's slot (nb = its batches of 64 records).  Slots are in light order, so segments still are; the dense slots'
	// serial walks -- the tail of the count pass at grid A x 65,536 lights, profiles/r01zl -- are cut into P shorter ones.
	const uint32_t parts = SUB ? p.parts : 1u;
	uint32_t part = 0, t_begin = p.tile_begin + (seg - p.seg_begin) * p.tiles_per_seg;
	if (parts > 1)
	{
		part = seg % parts;
		t_begin = seg / parts; // one tile per segment
	}
	const uint32_t t_end = min(t_begin + p.tiles_per_seg, p.tiles_total);
	const uint32_t ntiles = t_end - t_begin;
	grid_dep_launch();
	grid_dep_wait();
	unsigned long long t_start = 0;
	if (p.trace && threadIdx.x == 0)
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
	// the count pass left one word per (segment, warp program): nothing counted there means nothing to write.  One load
	// instead of the whole prologue -- 63 % of the write pass's warp programs at B1M ran ~190 instructions of prologue only
	// to find that they own no output (profiles/r01q).  Segment 0 stays: it writes the light cells.
	// (the slot index is recomputed where it is needed: keeping it live cost the count pass two registers = one CTA per SM)
	// rounds: the group held its cap everywhere before this round -- nothing to count, nothing to write, and nobody reads this
	// program's counts (hier_finalize_kernel stops at the round before which a cluster was full)
	if (p.round_alive && !p.trace && __ldcg(&p.round_alive[cb]) == 0u)
		return;
	if (WRITE && seg != 0 && !p.trace && HVCF_CHAIN_LD(&p.unit_any[any_slot(seg, unit)]) == 0u)
		return;
	// count pass: a group whose level-0 slots are empty for every tile of this segment has nothing to test -- zero counts,
	// no boxes, no early-exit counters, no union reductions (HVCF_EMPTY_SKIP=0 at build time switches the check off)
#ifndef HVCF_EMPTY_SKIP
#define HVCF_EMPTY_SKIP 1
#endif
	if (HVCF_EMPTY_SKIP && !WRITE && !p.trace)
	{
		const size_t d0 = static_cast<size_t>(cb) * p.tiles_cap + t_begin;
		uint32_t some = 0;
		if (parts > 1)
		{
			const uint32_t n0 = HVCF_CHAIN_LD(&p.bin_count[d0]);
			const uint32_t nb = n0 > kBinCap ? (INBOX ? (__ldcg(&p.tile_cnt[t_begin]) + 63u) / 64u : kTileGroups) : (n0 + 63) / 64;
			some = (part + 1) * nb / parts > part * nb / parts ? 1u : 0u;
		}
		else // (macro tiles: only the count of a macro tile's first tile means anything)
			for (uint32_t t = lane << p.macro_shift; t < ntiles; t += 32u << p.macro_shift)
				some |= HVCF_CHAIN_LD(&p.bin_count[d0 + t]);
		if (!__any_sync(0xffffffffu, some != 0))
		{
#pragma unroll
			for (int k = 0; k < K; ++k)
			{
				const uint32_t c = __ldg(&p.lane_cluster[(static_cast<size_t>(unit * K + k) << 5) + lane]);
				if (c != kSkip)
					p.seg_counts[static_cast<size_t>(seg) * p.cluster_count + c] = 0u;
			}
			if (lane == 0)
				p.unit_any[any_slot(seg, unit)] = 0u;
			return;
		}
	}

	// ---- this lane's K clusters (one per block) -------------------------------------------------------
	const float kInf = __int_as_float(0x7f800000);
	float lo[K][3], hi[K][3];
	uint32_t cidx[K];    // cluster index inside the shard, kSkip = no cluster on this lane
	uint32_t cnt[K];     // pass 1: hits counted in this segment; pass 2: entries written
	uint32_t limit[K];   // pass 1: hits still missing before the cap; pass 2: entries this lane owns
	uint32_t wbase[K];   // pass 2: first output slot
	bool active[K];
	const uint32_t chunk = p.round_chunk != 0xffffffffu ? p.round_chunk : hier_chunk_of(seg);
	const uint32_t index_base = (WRITE && p.index_base) ? *p.index_base : 0u;
#pragma unroll
	for (int k = 0; k < K; ++k)
	{
		cidx[k] = __ldg(&p.lane_cluster[(static_cast<size_t>(unit * K + k) << 5) + lane]);
		const bool valid = cidx[k] != kSkip;
		cnt[k] = 0;
		limit[k] = 0;
		wbase[k] = 0;
		active[k] = false;
		lo[k][0] = lo[k][1] = lo[k][2] = kInf;
		hi[k][0] = hi[k][1] = hi[k][2] = -kInf;
		if (valid)
		{
			// every load below is independent of the others: one memory round trip for the whole prologue
			const float4* a = reinterpret_cast<const float4*>(p.aabbs + p.cluster_begin + cidx[k]);
			const float4 mn = __ldg(a), mx = __ldg(a + 1);
			if (WRITE)
			{
				const size_t sc = static_cast<size_t>(seg) * p.cluster_count + cidx[k];
				const uint32_t n = p.seg_counts[sc], pfx = p.seg_prefix[sc];
				const uint32_t off = p.offsets[cidx[k]] + index_base;
				if (seg == 0) // LightCulling.comp:86-87
					*reinterpret_cast<uint2*>(p.cells + p.cluster_begin + cidx[k]) = make_uint2(off, p.counts[cidx[k]]);
				if (n != 0 && pfx != kSkip && pfx < p.cap)
				{
					active[k] = true;
					limit[k] = min(n, p.cap - pfx);
					wbase[k] = off + pfx;
				}
			}
			else
			{
				const uint32_t prior = hier_prior_hits(p.cum, cidx[k], chunk);
				if (prior < p.cap)
				{
					active[k] = true;
					limit[k] = p.cap - prior;
				}
			}
			if (active[k])
			{
				lo[k][0] = mn.x, lo[k][1] = mn.y, lo[k][2] = mn.z;
				hi[k][0] = mx.x, hi[k][1] = mx.y, hi[k][2] = mx.z;
			}
		}
	}
	if (WRITE && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && p.frame_total)
		*p.frame_total = index_base + *p.shard_total; // globalIndex after the frame (LightCulling.comp:21,79)

	// ---- nothing to do for this (unit, segment)?  Most warp programs of the write pass own no output (clusters fill
	// their quota in the early segments) and many of the count pass already know their clusters are full: leave before
	// the union reductions and the tile walk.  (Measured on B200, B1M: 0.295 -> 0.283 ms/frame.  Staging the survivor
	// slots up front with TMA bulk copies was also tried here and LOST -- 0.362 ms: most slots hold a handful of
	// records, so speculative 1 KiB reads per (unit, tile) cost more L2 traffic than the round trips they hide.)
	{
		bool want = false;
#pragma unroll
		for (int k = 0; k < K; ++k)
			want = want || active[k];
		if (!p.trace && !__any_sync(0xffffffffu, want))
		{
			if (!WRITE)
			{
#pragma unroll
				for (int k = 0; k < K; ++k)
					if (cidx[k] != kSkip)
						p.seg_counts[static_cast<size_t>(seg) * p.cluster_count + cidx[k]] = 0u;
				if (lane == 0)
					p.unit_any[any_slot(seg, unit)] = 0u;
			}
			return;
		}
	}

	// ---- block union boxes (level 1) over the lanes that still want hits ------------------------------
	float ulo[K][3], uhi[K][3];
	uint32_t alive[K]; // lanes of the block that still want hits
	bool blk_nan[K];   // the block holds a NaN box: its union stays 'pass everything'
	bool warp_done = true;
#pragma unroll
	for (int k = 0; k < K; ++k)
	{
		alive[k] = __ballot_sync(0xffffffffu, active[k]);
		// a NaN bound makes every light hit that cluster (both shader comparisons are false): pass everything
		const bool nan = active[k] && !(lo[k][0] == lo[k][0] && lo[k][1] == lo[k][1] && lo[k][2] == lo[k][2] &&
			hi[k][0] == hi[k][0] && hi[k][1] == hi[k][1] && hi[k][2] == hi[k][2]);
		blk_nan[k] = __any_sync(0xffffffffu, nan);
#pragma unroll
		for (int i = 0; i < 3; ++i)
		{
			ulo[k][i] = blk_nan[k] ? -kInf : warp_min_f(lo[k][i]);
			uhi[k][i] = blk_nan[k] ? kInf : warp_max_f(hi[k][i]);
		}
		warp_done = warp_done && alive[k] == 0;
	}

	unsigned long long n_l1 = 0, n_l2 = 0; // warp-uniform test counters (lane tests)
	const size_t desc0 = static_cast<size_t>(cb) * p.tiles_cap + t_begin;
	float4* const my_rec = s_rec2[warp][0];
	uint16_t* const my_idx = s_idx2[warp][0];
	// one batch of up to 64 records in ascending light order: r0 = record of lane, r1 = record of 32 + lane.  `staged`: the
	// records (and, where indices are needed, their tile indices) already sit in rec / idxs in exactly that order.
	auto batch = [&](const float4& r0, const float4& r1, uint32_t id0, uint32_t id1, uint32_t tile_base, float4* rec, uint16_t* idxs,
		bool staged) {
		// ---- level 1: lane = light, against the union box of each block -----------------------------
		uint32_t mlo[K], mhi[K];
		uint32_t any = 0;
#pragma unroll
		for (int k = 0; k < K; ++k)
		{
			mlo[k] = mhi[k] = 0;
			if (alive[k])
			{
				mlo[k] = __ballot_sync(0xffffffffu, sphere_box_hit(r0.x, r0.y, r0.z, r0.w, ulo[k], uhi[k]));
				mhi[k] = __ballot_sync(0xffffffffu, sphere_box_hit(r1.x, r1.y, r1.z, r1.w, ulo[k], uhi[k]));
				n_l1 += 64;
				any |= mlo[k] | mhi[k];
			}
		}
		if (any == 0)
			return;
		if (!staged)
		{
			__syncwarp(); // the previous batch's broadcasts are done (and, for a raw batch, every lane has read its pair)
			rec[lane] = r0;
			rec[32 + lane] = r1;
			idxs[lane] = static_cast<uint16_t>(id0);
			idxs[32 + lane] = static_cast<uint16_t>(id1);
			__syncwarp();
		}
		// ---- level 2: lane = cluster, candidates broadcast from the slab -----------------------------
		warp_done = true;
#pragma unroll
		for (int k = 0; k < K; ++k)
		{
			if (alive[k] == 0)
				continue;
			n_l2 += 32u * (__popc(mlo[k]) + __popc(mhi[k]));
#pragma unroll
			for (uint32_t half = 0; half < 2; ++half)
			{
				uint32_t m = half ? mhi[k] : mlo[k];
				// kCand candidates per trip while the mask holds that many: a trip is latency-bound (LDS -> 6-deep
				// FP chain), so independent chains are free; the remainder goes one by one
				while (__popc(m) >= static_cast<int>(kCand))
				{
					uint32_t j[kCand];
					float4 c[kCand];
#pragma unroll
					for (uint32_t q = 0; q < kCand; ++q)
					{
						j[q] = half * 32 + __ffs(m) - 1;
						m &= m - 1;
						c[q] = rec[j[q]]; // broadcast
					}
#pragma unroll
					for (uint32_t q = 0; q < kCand; ++q)
					{
						const bool hit = sphere_box_hit(c[q].x, c[q].y, c[q].z, c[q].w, lo[k], hi[k]) && active[k];
						if (WRITE)
						{
							if (hit && cnt[k] < limit[k])
							{
								p.indices[wbase[k] + cnt[k]] = tile_base + idxs[j[q]];
								++cnt[k];
							}
						}
						else
						{
							cnt[k] += hit ? 1u : 0u;
						}
					}
				}
				while (m)
				{
					const uint32_t j = half * 32 + __ffs(m) - 1;
					m &= m - 1;
					const float4 c = rec[j]; // broadcast
					const bool hit = sphere_box_hit(c.x, c.y, c.z, c.w, lo[k], hi[k]) && active[k];
					if (WRITE)
					{
						if (hit && cnt[k] < limit[k])
						{
							p.indices[wbase[k] + cnt[k]] = tile_base + idxs[j];
							++cnt[k];
						}
					}
					else
					{
						cnt[k] += hit ? 1u : 0u;
					}
				}
			}
			// early exit per lane: a cluster that holds its quota leaves the block's union box, so the remaining
			// lanes see fewer candidates; a block with no lane left is finished
			const uint32_t now_alive = __ballot_sync(0xffffffffu, active[k] && cnt[k] < limit[k]);
			if (now_alive != alive[k])
			{
				alive[k] = now_alive;
				active[k] = (now_alive >> lane) & 1u;
				if (now_alive && !blk_nan[k])
				{
#pragma unroll
					for (int a = 0; a < 3; ++a)
					{
						ulo[k][a] = warp_min_f(active[k] ? lo[k][a] : kInf);
						uhi[k][a] = warp_max_f(active[k] ? hi[k][a] : -kInf);
					}
				}
			}
			warp_done = warp_done && alive[k] == 0;
		}
	};
	// ---- the walk, software-pipelined: while batch i is tested, batch i + 1 (the next 64 records of the slot, or the first
	// batch of the next non-empty tile) is already on its way into the other shared-memory buffer (cp.async, SASS LDGSTS: no
	// registers held).  Measured on B200 (profiles/r01zj): grid C x 16M lights on one GPU, 32 tiles per segment, 3.63 -> 3.23 ms
	// per frame; no change at 1-2 tiles per segment (A65k, A1M) and a loss at 4 (B1M 0.215 -> 0.223 ms: 12 registers more, and
	// most slots there hold a single batch) -- hence a template parameter, set by the host from the tiles per segment.
	if constexpr (PIPE)
	{
		// macro tiles (p.macro_shift = log2 M; hier_bin2_kernel): a slot belongs to M consecutive tiles, its count sits at the first
		// of them, it holds M x kBinCap records and its 16-bit indices run over the M tiles.  M = 1: a slot per tile, as ever.
		const uint32_t ms = p.macro_shift, mstep = 1u << ms; // (routed frames: a route block's tiles, one slot per group and block)
		uint32_t it_t = 0, it_b = 0, it_n0 = 0, it_nb = 0; // the batch to FETCH next: (first) tile inside the segment, batch, slot count, batches
		bool it_raw = false;                                // the slot overflowed: its tiles are read raw
		uint32_t cnt_chunk = 0xffffffffu, my_cnt = 0;
		auto seek = [&]() -> bool { // first non-empty slot at or after it_t
			for (; it_t < ntiles; it_t += mstep)
			{
				const uint32_t mi = it_t >> ms;
				if ((mi >> 5) != cnt_chunk)
				{
					cnt_chunk = mi >> 5;
					const uint32_t tt = (cnt_chunk * 32 + lane) << ms;
					my_cnt = tt < ntiles ? HVCF_CHAIN_LD(&p.bin_count[desc0 + tt]) : 0u;
				}
				it_n0 = __shfl_sync(0xffffffffu, my_cnt, mi & 31u);
				if (it_n0)
				{
					const uint32_t meff = min(mstep, ntiles - it_t);
					it_raw = it_n0 > meff * kBinCap;
					it_nb = it_raw ? (INBOX ? (__ldcg(&p.tile_cnt[t_begin + it_t]) + 63u) / 64u : meff * kTileGroups) : (it_n0 + 63) / 64;
					it_b = 0;
					return true;
				}
			}
			return false;
		};
		auto fetch = [&](uint32_t buf) { // batch (it_t, it_b) -> buffer buf
			float4* const rec = s_rec2[warp][buf];
			const uint32_t tile_id = t_begin + it_t;
			if (it_raw) // raw tile(s): the lane's packed pair, converted when it is consumed
			{
				const float4* const lv = p.lights_view + static_cast<size_t>(tile_id) * (kTileBytes / 16) + it_b * 64;
				if constexpr (INBOX) // inbox tile: plain records, `valid` of them; zero fill = (0, 0, 0, r2 = 0) never hits
				{
					const uint32_t valid = __ldcg(&p.tile_cnt[tile_id]), i0 = it_b * 64 + lane, i1 = i0 + 32;
					cp_async16(rec + lane, lv + (i0 < valid ? lane : 0u), i0 < valid ? 16u : 0u);
					cp_async16(rec + 32 + lane, lv + (i1 < valid ? 32 + lane : 0u), i1 < valid ? 16u : 0u);
					if (WRITE && lane < 8)
						cp_async16(reinterpret_cast<uint4*>(s_idx2[warp][buf]) + lane,
							reinterpret_cast<const uint4*>(p.tile_ids + static_cast<size_t>(tile_id) * kTileLights + it_b * 64) + lane, 16u);
				}
				else
				{
					cp_async16(rec + lane, lv + lane, 16u);
					cp_async16(rec + 32 + lane, lv + 32 + lane, 16u);
				}
			}
			else
			{
				const size_t slot = (desc0 + it_t) * kBinCap;
				const uint32_t i0 = it_b * 64 + lane, i1 = i0 + 32;
				// beyond the slot's count: zero fill = (0, 0, 0, r2 = 0), and no squared distance is < 0
				cp_async16(rec + lane, p.bin_rec + slot + (i0 < it_n0 ? i0 : 0u), i0 < it_n0 ? 16u : 0u);
				cp_async16(rec + 32 + lane, p.bin_rec + slot + (i1 < it_n0 ? i1 : 0u), i1 < it_n0 ? 16u : 0u);
				if (WRITE && lane < 8) // 64 tile indices = 8 x 16 B (slots are whole: reading past the count is in bounds)
					cp_async16(reinterpret_cast<uint4*>(s_idx2[warp][buf]) + lane, reinterpret_cast<const uint4*>(p.bin_idx + slot + it_b * 64) + lane, 16u);
			}
			cp_async_commit();
		};
		bool have = !warp_done && seek();
		uint32_t cur = 0;
		if (have)
			fetch(0);
		while (have && !warp_done)
		{
			const uint32_t c_t = it_t, c_b = it_b;
			const bool c_raw = it_raw;
			bool have_next = true;
			if (++it_b >= it_nb)
			{
				it_t += mstep;
				have_next = seek();
			}
			__syncwarp(); // everyone is done with the buffer the next batch goes into
			if (have_next)
			{
				fetch(cur ^ 1u);
				cp_async_wait<1>();
			}
			else
				cp_async_wait<0>();
			__syncwarp();
			float4* const rec = s_rec2[warp][cur];
			uint16_t* const idxs = s_idx2[warp][cur];
			const uint32_t tile_base = (INBOX ? ((t_begin + c_t) & p.tile_mask) : (t_begin + c_t)) * kTileLights;
			if (INBOX && c_raw) // raw inbox tile: already plain records + their indices, as a slot batch
				batch(rec[lane], rec[32 + lane], 0u, 0u, tile_base, rec, idxs, true);
			else if (c_raw)
			{
				const float4 a = rec[lane], b = rec[32 + lane];
				batch(make_float4(a.x, a.z, b.x, b.z), make_float4(a.y, a.w, b.y, b.w), c_b * 64 + lane, c_b * 64 + 32 + lane, tile_base, rec, idxs, false);
			}
			else
				batch(rec[lane], rec[32 + lane], 0u, 0u, tile_base, rec, idxs, true);
			cur ^= 1u;
			have = have_next;
		}
		cp_async_wait<0>();
	}
	else
	{
	for (uint32_t tb = 0; tb < ntiles && !warp_done; tb += 32)
	{
		// survivor counts of the next 32 tiles in one coalesced load
		const uint32_t my_cnt = tb + lane < ntiles ? HVCF_CHAIN_LD(&p.bin_count[desc0 + tb + lane]) : 0u;
		const uint32_t tn = min(32u, ntiles - tb);
		for (uint32_t ti = 0; ti < tn && !warp_done; ++ti)
		{
			const uint32_t n0 = __shfl_sync(0xffffffffu, my_cnt, ti);
			if (n0 == 0)
				continue;
			const uint32_t tile_id = t_begin + tb + ti;
			const uint32_t tile_base = (INBOX ? (tile_id & p.tile_mask) : tile_id) * kTileLights;
			// a slot that overflowed means the group's filter is not selective for this tile: read the raw tile
			const bool raw = n0 > kBinCap;
			const uint32_t raw_valid = (INBOX && raw) ? __ldcg(&p.tile_cnt[tile_id]) : kTileLights; // inbox tiles are partly filled
			const uint32_t nb = raw ? (INBOX ? (raw_valid + 63u) / 64u : kTileGroups) : (n0 + 63) / 64;
			const float4* const lv = p.lights_view + static_cast<size_t>(tile_id) * (kTileBytes / 16);
			const size_t slot = (desc0 + tb + ti) * kBinCap;
			uint32_t b_begin = 0, b_end = nb;
			if (parts > 1) // (the division stays off the path of whole-tile segments)
			{
				b_begin = part * nb / parts;
				b_end = (part + 1) * nb / parts;
			}
			for (uint32_t bi = b_begin; bi < b_end && !warp_done; ++bi)
			{
				// ---- a batch of up to 64 lights in ascending order: records r0 (lane) and r1 (32 + lane) ----
				float4 r0, r1;
				uint32_t id0, id1;
				if (INBOX && raw) // inbox tile: plain records and their in-block indices
				{
					const uint32_t i0 = bi * 64 + lane, i1 = i0 + 32;
					const uint16_t* const ids = p.tile_ids + static_cast<size_t>(tile_id) * kTileLights;
					r0 = r1 = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
					id0 = id1 = 0;
					if (i0 < raw_valid)
					{
						r0 = lv[i0];
						id0 = ids[i0];
					}
					if (i1 < raw_valid)
					{
						r1 = lv[i1];
						id1 = ids[i1];
					}
				}
				else if (raw)
				{
					const float4 a = HVCF_CHAIN_LD(lv + bi * 64 + lane), b = HVCF_CHAIN_LD(lv + bi * 64 + 32 + lane);
					r0 = make_float4(a.x, a.z, b.x, b.z);
					r1 = make_float4(a.y, a.w, b.y, b.w);
					id0 = bi * 64 + lane;
					id1 = id0 + 32;
				}
				else
				{
					const uint32_t i0 = bi * 64 + lane, i1 = i0 + 32;
					r0 = r1 = make_float4(0.0f, 0.0f, 0.0f, -1.0f); // r2 = -1: never a hit
					id0 = id1 = 0;
					if (i0 < n0)
					{
						r0 = HVCF_CHAIN_LD(p.bin_rec + slot + i0);
						id0 = HVCF_CHAIN_LD(p.bin_idx + slot + i0);
					}
					if (i1 < n0)
					{
						r1 = HVCF_CHAIN_LD(p.bin_rec + slot + i1);
						id1 = HVCF_CHAIN_LD(p.bin_idx + slot + i1);
					}
				}
				batch(r0, r1, id0, id1, tile_base, my_rec, my_idx, false);
			}
		}
	}
	}
	if (!WRITE)
	{
		bool counted = false;
#pragma unroll
		for (int k = 0; k < K; ++k)
			counted = counted || cnt[k] != 0;
		const bool any_counted = __any_sync(0xffffffffu, counted);
		if (lane == 0)
			p.unit_any[any_slot(seg, unit)] = any_counted ? 1u : 0u;
#pragma unroll
		for (int k = 0; k < K; ++k)
			if (cidx[k] != kSkip)
			{
				p.seg_counts[static_cast<size_t>(seg) * p.cluster_count + cidx[k]] = cnt[k];
				if (cnt[k])
					atomicAdd(&p.cum[static_cast<size_t>(cidx[k]) * kHMaxChunks + chunk], cnt[k]);
			}
	}
	if (p.trace)
	{
		// debug only (HVCF_HIER_TRACE): per CTA {start ns, end ns, SM id, level-2 lane tests of warp 0}
		__syncthreads();
		if (threadIdx.x == 0)
		{
			unsigned long long t_end;
			uint32_t smid;
			asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
			asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
			const size_t cta = static_cast<size_t>(blockIdx.y) * gridDim.x + blockIdx.x, ctas = static_cast<size_t>(gridDim.x) * gridDim.y;
			unsigned long long* t = p.trace + 4ull * ((WRITE ? ctas : 0) + cta);
			t[0] = t_start, t[1] = t_end, t[2] = smid, t[3] = n_l2;
		}
	}
	if (lane == 0 && (n_l1 | n_l2))
	{
		if (WRITE) // the write pass repeats part of the count pass's tests: counted apart, so that per-kernel rates stay honest
			atomicAdd(&p.stats->write_tests, n_l1 + n_l2);
		else
		{
			atomicAdd(&p.stats->prefilter_tests, n_l1);
			atomicAdd(&p.stats->executed_tests, n_l2);
		}
	}
}

// ======================================================================================================
// launchers
// ======================================================================================================
static cudaLaunchAttribute pdl_attr()
{
	cudaLaunchAttribute a{};
	a.id = cudaLaunchAttributeProgrammaticStreamSerialization;
	a.val.programmaticStreamSerializationAllowed = 1;
	return a;
}

// ======================================================================================================
// Two-level level 0 ("bin2"), for shards with many cluster groups (grid C: 216 groups; one level costs groups x lights
// tests and was 2.9 of the 5.4 ms of the 16M-light frame, profiles/r01d).  Groups are taken SG at a time into
// super-groups (consecutive groups are spatial neighbours: blocks run x, then y, then z).  One CTA per light tile:
//   phase A  warp w tests the whole tile against super-group w, w + 8, ... (lane = light pair, FADD2/FMUL2, ballot/popc)
//            and compacts the survivors' tile indices, in light order, into a shared-memory list per super-group;
//   phase B  warp w takes groups w, w + 8, ... (one of every super-group when SG == 8, so the warps stay balanced), walks
//            its super-group's list (lane = surviving light, gathered from the staged tile) against the group's own
//            union box and writes the (group, tile) slot exactly as hier_bin_kernel does.
// The super box contains the group boxes and the test is the same monotone float sequence, so phase A can only pass
// more lights than phase B needs: the slots are byte-identical to the one-level kernel's.  The tile is staged once
// per tile instead of once per 8 groups.
// ======================================================================================================
__global__ void hier_super_unions_kernel(const float* __restrict__ cb_union, uint32_t ncb, uint32_t sg, float* __restrict__ sb_union)
{
	const uint32_t s = blockIdx.x, c = threadIdx.x; // c: 0..2 lo, 4..6 hi
	if (c >= 8)
		return;
	const bool is_hi = c >= 4;
	float v = is_hi ? -__int_as_float(0x7f800000) : __int_as_float(0x7f800000);
	for (uint32_t g = s * sg; g < min(ncb, (s + 1) * sg); ++g)
	{
		const float u = cb_union[static_cast<size_t>(g) * 8 + c];
		v = is_hi ? fmaxf(v, u) : fminf(v, u);
	}
	sb_union[static_cast<size_t>(s) * 8 + c] = (c & 3) == 3 ? 0.0f : v;
}

// ROUTED (hvcf_shard.cu): the tile is a tile of this rank's INBOX -- up to 1,024 plain {x, y, z, r*r} records that the light
// owners compacted in light order, `valid` of them in use -- and the slot's index of a record is the light's 16-bit index inside
// its route block (rebuilt from the block's survivor bitmask, in_mask), not its position.  Same two phases, same slots; an empty tile only zeroes its slot counts.
//
// MACRO TILES (p.macro = M > 1; late rounds of a frame cut into rounds): one CTA takes M consecutive tiles, one after the other
// (TMA double buffer), and APPENDS every group's survivors to ONE slot per (group, macro tile) -- the M tiles' slots are
// contiguous in memory, so the macro slot starts at the first tile's slot and holds M x kBinCap records; a record's 16-bit index
// is (tile inside the macro tile) << 10 | (light inside the tile).  Where few lights survive (the far field of the reference
// scene: 98 % of the lights produce 0.1 % of the hits) the count / write passes then visit one dense slot instead of M nearly
// empty ones; the per-visit latency was their whole cost there (profiles/r02m).  Only the slot count of the macro tile's FIRST
// tile is written and read.
//
// Rounds (p.group_alive): only the super-groups that still hold a live group are tested, dead groups keep an empty slot, and a
// tile whose groups are all dead is not even staged.
template <bool ROUTED>
__global__ void __launch_bounds__(kBinWarps * 32) hier_bin2_kernel(const BinParams p)
{
	extern __shared__ __align__(128) unsigned char s_dyn[];
	const uint32_t M = p.macro; // (routed frames: 1, or the tiles of a route block -- one slot per group and BLOCK)
	const uint32_t nbuf = M > 1u ? 2u : 1u;
	const uint32_t LS = p.list_supers; // super-group survivor lists held at a time
	unsigned char* const s_tiles = s_dyn;                                                          // nbuf x kTileBytes
	uint32_t* const s_mask = reinterpret_cast<uint32_t*>(s_dyn + nbuf * kTileBytes);               // [ns][32] survivor bits of the tile per live super-group
	uint16_t* const s_list = reinterpret_cast<uint16_t*>(s_mask + static_cast<size_t>(p.ns) * 32); // [LS][kTileLights]
	uint32_t* const s_cnt = reinterpret_cast<uint32_t*>(s_list + static_cast<size_t>(LS) * kTileLights); // [LS]
	uint32_t* const s_gcnt = s_cnt + ((LS + 3u) & ~3u);                                            // [ncb] running slot counts
	uint32_t* const s_galive = s_gcnt + ((p.ncb + 3u) & ~3u);                                      // [ceil(ncb / 32)] live-group bits
	float4* const s_sbox = reinterpret_cast<float4*>(s_galive + (((p.ncb + 31u) / 32u + 3u) & ~3u)); // [ns][2] super-group boxes {lo, hi}, live ones first
	float4* const s_gbox = s_sbox + 2u * p.ns;                                                     // [live groups][2] group boxes, in group order
	uint16_t* const s_glist = reinterpret_cast<uint16_t*>(s_gbox + 2u * p.ncb);                    // [live groups] group ids
	uint16_t* const s_ids = s_glist + ((p.ncb + 7u) & ~7u);                                        // ROUTED: [kTileLights]
	__shared__ __align__(8) uint64_t bar[2];
	__shared__ uint32_t s_smask;
	__shared__ uint32_t s_super[32]; // the i-th live super-group
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t tile_id = p.tile_begin + blockIdx.x * M;                 // first tile of this CTA
	uint32_t tM = min(M, p.tile_end - tile_id);                             // its tiles
	grid_dep_launch();
	if (threadIdx.x == 0)
	{
		mbar_init(&bar[0], 1);
		mbar_init(&bar[1], 1);
		mbar_fence_init();
		s_smask = 0u;
	}
	__syncthreads();
	grid_dep_wait(); // the tile is written by the transform kernel of this frame (or arrived from another rank)
	uint32_t valid = kTileLights; // records of the current tile
	uint32_t left = 0;            // ROUTED: records of the route block from this CTA's first tile on
	if constexpr (ROUTED)
	{
		const uint32_t in_block = (tile_id & (p.block_tiles - 1u)) * kTileLights;
		const uint32_t c = p.in_cnt[tile_id / p.block_tiles];
		left = c > in_block ? c - in_block : 0u;
		if (M == 1u)
			left = min(left, kTileLights);
		valid = min(left, kTileLights);
		tM = min(tM, (left + kTileLights - 1u) / kTileLights); // only the tiles of the block's region that hold records
		if (threadIdx.x == 0)
			p.tile_cnt[tile_id] = left; // (a macro tile's raw walk runs over all of its records)
		if (left == 0u)
		{
			for (uint32_t cb = threadIdx.x; cb < p.ncb; cb += blockDim.x)
				p.bin_count[static_cast<size_t>(cb) * p.tiles_cap + tile_id] = 0u;
			return;
		}
	}
	// ---- which groups are alive: bits in shared memory, live super-groups as a mask; dead groups get their empty slot here ----
	{
		uint32_t mine = 0u;
		for (uint32_t cb0 = 0; cb0 < p.ncb; cb0 += blockDim.x)
		{
			const uint32_t cb = cb0 + threadIdx.x;
			const bool a = cb < p.ncb && (!p.group_alive || __ldcg(&p.group_alive[cb]) != 0u);
			const uint32_t bits = __ballot_sync(0xffffffffu, a);
			if (cb < p.ncb)
			{
				if (lane == 0)
					s_galive[cb >> 5] = bits;
				s_gcnt[cb] = 0u;
				if (a)
					mine |= 1u << (cb / p.sg);
				else
					p.bin_count[static_cast<size_t>(cb) * p.tiles_cap + tile_id] = 0u;
			}
		}
		mine = __reduce_or_sync(0xffffffffu, mine);
		if (lane == 0 && mine)
			atomicOr(&s_smask, mine);
	}
	__syncthreads();
	const uint32_t smask = s_smask;
	if (smask == 0u)
		return;
	const unsigned char* const src = reinterpret_cast<const unsigned char*>(p.lights_view);
	if (threadIdx.x == 0)
	{
		mbar_arrive_expect_tx(&bar[0], ROUTED ? valid * 16u : kTileBytes);
		bulk_copy_g2s(s_tiles, src + static_cast<size_t>(tile_id) * kTileBytes, ROUTED ? valid * 16u : kTileBytes, &bar[0]);
	}
	const uint32_t n_as = __popc(smask); // live super-groups
	// ---- the live groups, compacted in group order, with their boxes (one global load per group and CTA: a load per group and
	// tile inside phase B was that phase's latency chain), and the live super-groups' boxes: the static ones, or -- rounds, where
	// the group boxes have shrunk to their live clusters -- the union of their live groups' boxes ----
	uint32_t n_ag = 0;
	{
		const uint32_t words = (p.ncb + 31u) / 32u;
		for (uint32_t w = 0; w < words; ++w)
			n_ag += __popc(s_galive[w]);
		for (uint32_t cb = threadIdx.x; cb < p.ncb; cb += blockDim.x)
			if ((s_galive[cb >> 5] >> (cb & 31u)) & 1u)
			{
				uint32_t pos = __popc(s_galive[cb >> 5] & ((1u << (cb & 31u)) - 1u));
				for (uint32_t w = 0; w < (cb >> 5); ++w)
					pos += __popc(s_galive[w]);
				const float4* u = reinterpret_cast<const float4*>(p.cb_union + static_cast<size_t>(cb) * 8);
				s_glist[pos] = static_cast<uint16_t>(cb);
				s_gbox[2 * pos] = __ldcg(u);
				s_gbox[2 * pos + 1] = __ldcg(u + 1);
			}
		if (threadIdx.x < n_as)
			s_super[threadIdx.x] = __fns(smask, 0u, static_cast<int>(threadIdx.x) + 1);
		for (uint32_t sidx = threadIdx.x; sidx < n_as * 8u; sidx += blockDim.x)
		{
			const uint32_t si = sidx >> 3, c = sidx & 7u;
			const uint32_t sgi = __fns(smask, 0u, static_cast<int>(si) + 1);
			float v;
			if (!p.group_alive)
				v = __ldg(p.sb_union + static_cast<size_t>(sgi) * 8 + c);
			else
			{
				const bool is_hi = c >= 4;
				v = is_hi ? -__int_as_float(0x7f800000) : __int_as_float(0x7f800000);
				for (uint32_t g = sgi * p.sg; g < min(p.ncb, (sgi + 1u) * p.sg); ++g)
					if ((s_galive[g >> 5] >> (g & 31u)) & 1u)
					{
						const float x = __ldcg(p.cb_union + static_cast<size_t>(g) * 8 + c);
						v = is_hi ? fmaxf(v, x) : fminf(v, x);
					}
				if ((c & 3u) == 3u)
					v = 0.0f;
			}
			reinterpret_cast<float*>(s_sbox + 2 * si)[c] = v;
		}
	}
	__shared__ uint32_t s_wsum[kBinWarps];
	__shared__ uint32_t s_carry;
	// ROUTED: the in-block light index of tile t's records.  The j-th record of the block's region is the j-th set bit of the
	// block's survivor bitmask (block_tiles x 32 words, one bit per light of the block in light order).  Each thread takes the
	// words t, t + 256, ...: an exclusive prefix of the popcounts gives the rank of a word's first set bit.  Also kept in
	// tile_ids for the raw walk of the count / write passes (a slot that overflowed reads the inbox records themselves).
	auto rebuild_ids = [&](uint32_t t, uint32_t t_valid) {
		const uint32_t words = p.block_tiles * 32u;
		const uint32_t* mask = p.in_mask + static_cast<size_t>(t & ~(p.block_tiles - 1u)) * 32u;
		const uint32_t r_begin = (t & (p.block_tiles - 1u)) * kTileLights, r_end = r_begin + t_valid;
		if (threadIdx.x == 0)
			s_carry = 0u;
		__syncthreads();
		for (uint32_t w0 = 0; w0 < words; w0 += blockDim.x)
		{
			const uint32_t wi = w0 + threadIdx.x;
			uint32_t word = wi < words ? mask[wi] : 0u;
			const uint32_t c = __popc(word);
			uint32_t inc = c;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1)
			{
				const uint32_t tt = __shfl_up_sync(0xffffffffu, inc, o);
				if (lane >= static_cast<uint32_t>(o))
					inc += tt;
			}
			if (lane == 31)
				s_wsum[warp] = inc;
			__syncthreads();
			uint32_t before = s_carry;
			for (uint32_t w = 0; w < warp; ++w)
				before += s_wsum[w];
			uint32_t rank = before + inc - c; // rank of this word's first set bit inside the block's region
			if (rank < r_end && rank + c > r_begin)
				while (word)
				{
					const uint32_t bit = __ffs(word) - 1;
					word &= word - 1;
					if (rank >= r_begin && rank < r_end)
						s_ids[rank - r_begin] = static_cast<uint16_t>(wi * 32u + bit);
					++rank;
				}
			__syncthreads();
			if (threadIdx.x == blockDim.x - 1)
				s_carry = before + inc;
			__syncthreads();
		}
		for (uint32_t i = threadIdx.x; i < t_valid; i += blockDim.x)
			p.tile_ids[static_cast<size_t>(t) * kTileLights + i] = s_ids[i];
	};
	if constexpr (ROUTED)
		rebuild_ids(tile_id, valid);
	__syncthreads(); // the compacted groups, the boxes (and the in-block indices) are in place
	const uint32_t lt = lanemask_lt();
	const uint32_t slot_cap = (ROUTED ? M : tM) * kBinCap; // records of a (group, macro tile) slot
	unsigned long long tests = 0;
	for (uint32_t m = 0; m < tM; ++m)
	{
		const uint32_t buf = nbuf == 2u ? (m & 1u) : 0u;
		if constexpr (ROUTED)
		{
			valid = min(left - m * kTileLights, kTileLights);
			if (m != 0u)
				rebuild_ids(tile_id + m, valid); // (ends with a barrier-free store loop; the barrier below the next phase A1 orders s_ids)
		}
		const uint32_t ngroups = ROUTED ? (valid + 63u) / 64u : kTileGroups;
		// the next tile goes into the other buffer: every warp left it at the barrier that ended tile m - 1
		if (m + 1u < tM && threadIdx.x == 0)
		{
			const uint32_t bytes = ROUTED ? min(left - (m + 1u) * kTileLights, kTileLights) * 16u : kTileBytes;
			mbar_arrive_expect_tx(&bar[buf ^ 1u], bytes);
			bulk_copy_g2s(s_tiles + (buf ^ 1u) * kTileBytes, src + static_cast<size_t>(tile_id + m + 1u) * kTileBytes, bytes, &bar[buf ^ 1u]);
		}
		mbar_wait(&bar[buf], (m >> 1) & 1u);
		const float4* tile = reinterpret_cast<const float4*>(s_tiles + buf * kTileBytes);

		// ---- phase A1: warp w takes the 128 lights of pair rounds 2 w and 2 w + 1 against EVERY live super-group (the same
		// work for every warp, whatever the number of live super-groups); the ballots are the tile's survivor bits per super-group ----
#pragma unroll
		for (uint32_t gg = 0; gg < 2; ++gg)
		{
			const uint32_t g = warp * 2u + gg;
			if (g >= ngroups)
			{
				for (uint32_t si = lane; si < n_as; si += 32)
					s_mask[si * 32 + 2 * g] = s_mask[si * 32 + 2 * g + 1] = 0u;
				continue;
			}
			float4 a, b;
			bool v0 = true, v1 = true;
			if constexpr (ROUTED)
			{
				// plain records of lights g * 64 + lane and g * 64 + 32 + lane, paired in registers into the packed operand shape
				const float4 r0 = tile[g * 64 + lane], r1 = tile[g * 64 + 32 + lane];
				a = make_float4(r0.x, r1.x, r0.y, r1.y);
				b = make_float4(r0.z, r1.z, r0.w, r1.w);
				v0 = g * 64 + lane < valid; // beyond the tile's records: whatever shared memory held
				v1 = g * 64 + 32 + lane < valid;
			}
			else
				a = tile[g * 64 + lane], b = tile[g * 64 + 32 + lane];
			for (uint32_t si = 0; si < n_as; ++si)
			{
				const float4 a4 = s_sbox[2 * si], b4 = s_sbox[2 * si + 1];
				const float lo[3] = {a4.x, a4.y, a4.z}, hi[3] = {b4.x, b4.y, b4.z};
				bool h0, h1;
				sphere_box_hit2(a, b, lo, hi, h0, h1);
				const uint32_t m0 = __ballot_sync(0xffffffffu, h0 && v0), m1 = __ballot_sync(0xffffffffu, h1 && v1);
				if (lane == 0)
				{
					s_mask[si * 32 + 2 * g] = m0;
					s_mask[si * 32 + 2 * g + 1] = m1;
				}
			}
			tests += static_cast<unsigned long long>(n_as) * (ROUTED ? min(64u, valid - g * 64u) : 64u);
		}
		__syncthreads();
		const float* tile_f = reinterpret_cast<const float*>(tile);
		for (uint32_t c0 = 0; c0 < n_as; c0 += LS)
		{
			// ---- phase A2: the bits of up to LS live super-groups become index lists in light order (lane = word: words 2 g and
			// 2 g + 1 hold lights g * 64 .. g * 64 + 63) ----
			for (uint32_t si = c0 + warp; si < min(c0 + LS, n_as); si += kBinWarps)
			{
				uint32_t word = s_mask[si * 32 + lane];
				const uint32_t c = __popc(word);
				uint32_t inc = c;
#pragma unroll
				for (int o = 1; o < 32; o <<= 1)
				{
					const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
					if (lane >= static_cast<uint32_t>(o))
						inc += t;
				}
				uint16_t* out = s_list + static_cast<size_t>(si - c0) * kTileLights + (inc - c);
				while (word)
				{
					const uint32_t bit = __ffs(word) - 1;
					word &= word - 1;
					*out++ = static_cast<uint16_t>(lane * 32u + bit);
				}
				if (lane == 31)
					s_cnt[si - c0] = inc;
			}
			__syncthreads();

			// ---- phase B: every live group of these super-groups against its super-group's survivors ----
			for (uint32_t gi = warp; gi < n_ag; gi += kBinWarps)
			{
				const uint32_t cb = s_glist[gi];
				const uint32_t sgi = cb / p.sg;
				const uint32_t si = __popc(smask & ((1u << sgi) - 1u));
				if (si < c0 || si >= c0 + LS)
					continue;
				const uint32_t ns = s_cnt[si - c0];
				if (ns == 0u)
					continue;
				const uint16_t* const list = s_list + static_cast<size_t>(si - c0) * kTileLights;
				const float4 a4 = s_gbox[2 * gi], b4 = s_gbox[2 * gi + 1];
				const float lo[3] = {a4.x, a4.y, a4.z}, hi[3] = {b4.x, b4.y, b4.z};
				const size_t slot = (static_cast<size_t>(cb) * p.tiles_cap + tile_id) * kBinCap;
				float4* const rec = p.bin_rec + slot;
				uint16_t* const idx = p.bin_idx + slot;
				uint32_t cnt = s_gcnt[cb];
				const uint32_t cnt_in = cnt;
				// two list entries per lane per trip (i0 + lane, then i0 + 32 + lane: ascending), tested as a pair on FADD2/FMUL2 like the
				// tile pairs of phase A; a missing entry is the never-hit record (r2 = -1)
				for (uint32_t i0 = 0; i0 < ns; i0 += 64)
				{
					const uint32_t ia = i0 + lane, ib = ia + 32;
					uint32_t ta = 0, tb = 0;
					float4 pa = make_float4(0.0f, 0.0f, 0.0f, 0.0f), pb = make_float4(0.0f, 0.0f, -1.0f, -1.0f); // (x0,x1,y0,y1), (z0,z1,r0,r1)
					if (ia < ns)
					{
						ta = list[ia];
						if constexpr (ROUTED)
						{
							const float4 r = tile[ta];
							pa.x = r.x, pa.z = r.y, pb.x = r.z, pb.z = r.w;
						}
						else
						{
							// packed culling layout: light t sits in pair (t & 31) of group t >> 6, half (t >> 5) & 1
							const uint32_t f = (((ta >> 6) * 64 + (ta & 31)) << 2) + ((ta >> 5) & 1);
							pa.x = tile_f[f], pa.z = tile_f[f + 2], pb.x = tile_f[f + 128], pb.z = tile_f[f + 130];
						}
					}
					if (ib < ns)
					{
						tb = list[ib];
						if constexpr (ROUTED)
						{
							const float4 r = tile[tb];
							pa.y = r.x, pa.w = r.y, pb.y = r.z, pb.w = r.w;
						}
						else
						{
							const uint32_t f = (((tb >> 6) * 64 + (tb & 31)) << 2) + ((tb >> 5) & 1);
							pa.y = tile_f[f], pa.w = tile_f[f + 2], pb.y = tile_f[f + 128], pb.w = tile_f[f + 130];
						}
					}
					bool h0, h1;
					sphere_box_hit2(pa, pb, lo, hi, h0, h1);
					const uint32_t m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
					if ((m0 | m1) != 0)
					{
						const uint32_t d0 = cnt + __popc(m0 & lt), d1 = cnt + __popc(m0) + __popc(m1 & lt);
						if (h0 && d0 < slot_cap)
						{
							rec[d0] = make_float4(pa.x, pa.z, pb.x, pb.z);
							idx[d0] = ROUTED ? s_ids[ta] : static_cast<uint16_t>((m << 10) | ta);
						}
						if (h1 && d1 < slot_cap)
						{
							rec[d1] = make_float4(pa.y, pa.w, pb.y, pb.w);
							idx[d1] = ROUTED ? s_ids[tb] : static_cast<uint16_t>((m << 10) | tb);
						}
						cnt += __popc(m0) + __popc(m1);
					}
				}
				if (lane == 0 && cnt != cnt_in)
					s_gcnt[cb] = cnt;
				tests += ns;
			}
			__syncthreads(); // the lists (and, after the last chunk, the masks and two tiles on the tile buffer) are free again
		}
	}
	for (uint32_t cb = threadIdx.x; cb < p.ncb; cb += blockDim.x)
		if ((s_galive[cb >> 5] >> (cb & 31u)) & 1u)
			p.bin_count[static_cast<size_t>(cb) * p.tiles_cap + tile_id] = s_gcnt[cb];
	if (lane == 0 && tests)
	{
		atomicAdd(&p.stats->prefilter_tests, tests);
		atomicAdd(&p.stats->bin_tests, tests);
	}
}

cudaError_t launch_hier_alive(const uint32_t* lane_cluster, const hvcf_aabb* aabbs, uint32_t cluster_begin, const uint32_t* cum, uint32_t K, uint32_t ncb,
	uint32_t cap, uint32_t* group_alive, float* alive_union, cudaStream_t stream)
{
	if (ncb == 0)
		return cudaSuccess;
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = dim3(ncb, 1, 1);
	cfg.blockDim = dim3(256, 1, 1);
	cfg.stream = stream;
	cudaLaunchAttribute attr{};
	attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr.val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = &attr;
	cfg.numAttrs = 1;
	return cudaLaunchKernelEx(&cfg, hier_alive_kernel, lane_cluster, aabbs, cluster_begin, cum, K, cap, group_alive, alive_union);
}

cudaError_t launch_hier_unions(const uint32_t* lane_cluster, const hvcf_aabb* aabbs, uint32_t cluster_begin, uint32_t K, uint32_t ncb,
	float* cb_union, cudaStream_t stream)
{
	if (ncb == 0)
		return cudaSuccess;
	hier_unions_kernel<<<ncb, 256, 0, stream>>>(lane_cluster, aabbs, cluster_begin, K, cb_union);
	return cudaGetLastError();
}

// ======================================================================================================
// transform + level 0 in ONE kernel, for frames with few lights (grid A x 65,536: the two kernels are 4 + 7 us of pure launch
// and memory latency in a 57 us frame).  Every CTA of a tile transforms the tile's 1,024 lights itself (LightCulling.comp:51-53,
// the arithmetic of light_transform_kernel, 28 flops per light -- cheaper than a round trip through HBM) straight into the
// shared-memory tile; the CTA of the first group column also stores the tile to lights_view for the passes that follow.
// It is the first kernel of the frame, so it also clears the early-exit counters and the work counters; its own test count
// is a product (every light against every group), stored rather than accumulated.
// ======================================================================================================
constexpr uint32_t kXBinMinCtas = 32;
__global__ void __launch_bounds__(kBinWarps * 32) hier_xbin_kernel(const XBinParams q)
{
	__shared__ __align__(128) float4 s_tile4[kTileBytes / 16];
	const BinParams& p = q.bin;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t tile_id = blockIdx.y, cb = blockIdx.x * kBinWarps + warp;
	grid_dep_launch();
	{
		const uint32_t nthreads = gridDim.x * gridDim.y * blockDim.x;
		const uint32_t tid = (blockIdx.y * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
		for (uint32_t i = tid; i < q.zero_words; i += nthreads)
			q.zero[i] = 0u;
		if (tid == 0)
		{
			DeviceStats* st = p.stats;
			st->executed_tests = 0ull;
			st->total_indices = 0u;
			st->overflow_clusters = 0u;
			st->write_tests = 0ull;
			st->bin_tests = st->prefilter_tests = static_cast<unsigned long long>(p.ncb) * q.tiles * kTileLights;
			stats_next_epoch(st);
		}
	}
	if (tile_id >= q.tiles) // CTAs beyond the last tile only help with the clearing (frames of a few tiles)
		return;
	const float* view = q.ubo->camera_view;
	// 512 lane pairs per tile, two per thread
#pragma unroll
	for (uint32_t r = 0; r < 2; ++r)
	{
		const uint32_t pr = r * 256 + threadIdx.x; // pair inside the tile
		const uint32_t g = pr >> 5, l = pr & 31;
		float x[2], y[2], z[2], r2[2];
#pragma unroll
		for (int h = 0; h < 2; ++h)
		{
			const uint32_t i = tile_id * kTileLights + g * kGroupLights + h * 32 + l;
			if (i < q.num_lights)
			{
				const float4 pw = *reinterpret_cast<const float4*>(q.lights[i].position);
				const float radius = q.lights[i].radius;
				x[h] = mat_row_dot(view, 0, pw.x, pw.y, pw.z, pw.w);
				y[h] = mat_row_dot(view, 1, pw.x, pw.y, pw.z, pw.w);
				z[h] = mat_row_dot(view, 2, pw.x, pw.y, pw.z, pw.w);
				r2[h] = __fmul_rn(radius, radius);
			}
			else
			{
				x[h] = y[h] = z[h] = 0.0f;
				r2[h] = -1.0f; // pad lights cannot hit
			}
		}
		s_tile4[g * 64 + l] = make_float4(x[0], x[1], y[0], y[1]);
		s_tile4[g * 64 + 32 + l] = make_float4(z[0], z[1], r2[0], r2[1]);
	}
	__syncthreads();
	if (blockIdx.x == 0)
	{
		float4* out = const_cast<float4*>(p.lights_view) + static_cast<size_t>(tile_id) * (kTileBytes / 16);
		for (uint32_t i = threadIdx.x; i < kTileBytes / 16; i += blockDim.x)
			out[i] = s_tile4[i];
	}
	if (cb >= p.ncb)
		return;
	float lo[3], hi[3];
	{
		const float4* u = reinterpret_cast<const float4*>(p.cb_union + static_cast<size_t>(cb) * 8);
		const float4 a = __ldg(u), b = __ldg(u + 1);
		lo[0] = a.x, lo[1] = a.y, lo[2] = a.z;
		hi[0] = b.x, hi[1] = b.y, hi[2] = b.z;
	}
	const size_t slot = (static_cast<size_t>(cb) * p.tiles_cap + tile_id) * kBinCap;
	float4* const rec = p.bin_rec + slot;
	uint16_t* const idx = p.bin_idx + slot;
	const uint32_t lt = lanemask_lt();
	uint32_t cnt = 0;
#pragma unroll 4
	for (uint32_t g = 0; g < kTileGroups; ++g)
	{
		const float4 a = s_tile4[g * 64 + lane], b = s_tile4[g * 64 + 32 + lane];
		bool h0, h1;
		sphere_box_hit2(a, b, lo, hi, h0, h1);
		const uint32_t m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
		if ((m0 | m1) == 0)
			continue;
		const uint32_t d0 = cnt + __popc(m0 & lt), d1 = cnt + __popc(m0) + __popc(m1 & lt);
		if (h0 && d0 < kBinCap)
		{
			rec[d0] = make_float4(a.x, a.z, b.x, b.z);
			idx[d0] = static_cast<uint16_t>(g * 64 + lane);
		}
		if (h1 && d1 < kBinCap)
		{
			rec[d1] = make_float4(a.y, a.w, b.y, b.w);
			idx[d1] = static_cast<uint16_t>(g * 64 + 32 + lane);
		}
		cnt += __popc(m0) + __popc(m1);
	}
	if (lane == 0)
		p.bin_count[static_cast<size_t>(cb) * p.tiles_cap + tile_id] = cnt;
}

cudaError_t launch_hier_xbin(const XBinParams& q, uint32_t tiles, cudaStream_t stream)
{
	if (q.bin.ncb == 0 || tiles == 0)
		return cudaSuccess;
	XBinParams qq = q;
	qq.tiles = tiles;
	const uint32_t gx = (q.bin.ncb + kBinWarps - 1) / kBinWarps;
	const uint32_t min_rows = q.zero_words ? (kXBinMinCtas + gx - 1) / gx : 0u;
	hier_xbin_kernel<<<dim3(gx, tiles > min_rows ? tiles : min_rows, 1), kBinWarps * 32, 0, stream>>>(qq);
	return cudaGetLastError();
}

cudaError_t launch_hier_super_unions(const float* cb_union, uint32_t ncb, uint32_t sg, uint32_t ns, float* sb_union, cudaStream_t stream)
{
	if (ns == 0)
		return cudaSuccess;
	hier_super_unions_kernel<<<ns, 32, 0, stream>>>(cb_union, ncb, sg, sb_union);
	return cudaGetLastError();
}

size_t hier_bin2_smem_bytes(uint32_t ns, uint32_t ncb, bool routed, uint32_t macro, uint32_t list_supers)
{
	return static_cast<size_t>(macro > 1 ? 2 : 1) * kTileBytes + static_cast<size_t>(ns) * 32 * sizeof(uint32_t) +
		static_cast<size_t>(list_supers) * kTileLights * sizeof(uint16_t) + static_cast<size_t>((list_supers + 3u) & ~3u) * sizeof(uint32_t) +
		static_cast<size_t>((ncb + 3u) & ~3u) * sizeof(uint32_t) + static_cast<size_t>(((ncb + 31u) / 32u + 3u) & ~3u) * sizeof(uint32_t) +
		(static_cast<size_t>(ns) + ncb) * 32u + static_cast<size_t>((ncb + 7u) & ~7u) * sizeof(uint16_t) + (routed ? kTileLights * sizeof(uint16_t) : 0);
}

static cudaError_t launch_hier_bin2(const BinParams& p0, uint32_t tiles, cudaStream_t stream)
{
	BinParams p = p0;
	const bool routed = p.in_mask != nullptr;
	if (p.macro == 0)
		p.macro = 1;
	if (routed && p.macro != 1 && p.macro != p.block_tiles)
		return cudaErrorInvalidValue; // a routed frame's macro tile is a route block
	if (p.macro > 64 || (p.macro & (p.macro - 1u)))
		return cudaErrorInvalidValue; // a record's 16-bit index holds 6 + 10 bits
	p.tile_end = p.tile_begin + tiles;
	// survivor lists of at most 16 super-groups at a time (8 with the second tile buffer of macro tiles): 3 CTAs per SM.  Late
	// rounds rarely have more live super-groups than that; the first round takes two or three passes over its lists.
	p.list_supers = p.ns < (p.macro > 1 ? 8u : 16u) ? p.ns : (p.macro > 1 ? 8u : 16u);
	const size_t smem = hier_bin2_smem_bytes(p.ns, p.ncb, routed, p.macro, p.list_supers);
	if (smem > 227u * 1024u || p.ncb > 0xffffu)
		return cudaErrorInvalidValue; // (a grid of more than ~5,000 cluster groups: the group tables no longer fit)
	static size_t configured[2] = {0, 0};
	if (smem > configured[routed])
	{
		cudaError_t e = routed ? cudaFuncSetAttribute(hier_bin2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem))
			: cudaFuncSetAttribute(hier_bin2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
		if (e != cudaSuccess)
			return e;
		configured[routed] = smem;
	}
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = dim3((tiles + p.macro - 1u) / p.macro, 1, 1);
	cfg.blockDim = dim3(kBinWarps * 32, 1, 1);
	cfg.dynamicSmemBytes = smem;
	cfg.stream = stream;
	cudaLaunchAttribute attr = pdl_attr();
	cfg.attrs = &attr;
	cfg.numAttrs = 1;
	return routed ? cudaLaunchKernelEx(&cfg, hier_bin2_kernel<true>, p) : cudaLaunchKernelEx(&cfg, hier_bin2_kernel<false>, p);
}

cudaError_t launch_hier_bin(const BinParams& p, uint32_t tiles, cudaStream_t stream)
{
	if (p.ncb == 0 || tiles == 0)
		return cudaSuccess;
	if (p.ns)
		return launch_hier_bin2(p, tiles, stream);
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = dim3((p.ncb + kBinWarps - 1) / kBinWarps, tiles, 1);
	cfg.blockDim = dim3(kBinWarps * 32, 1, 1);
	cfg.stream = stream;
	cudaLaunchAttribute attr = pdl_attr();
	cfg.attrs = &attr;
	cfg.numAttrs = 1;
	return cudaLaunchKernelEx(&cfg, hier_bin_kernel, p);
}

template <int K, bool WRITE, bool PIPE, bool SUB, bool INBOX = false>
static cudaError_t launch_hier2_t(const HierParams& p, cudaStream_t stream)
{
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = dim3(p.ncb * 8 / kMainWarps, p.seg_end - p.seg_begin, 1); // x fastest: all units of a segment, then the next segment
	cfg.blockDim = dim3(kMainWarps * 32, 1, 1);
	cfg.stream = stream;
	cudaLaunchAttribute attr = pdl_attr();
	cfg.attrs = &attr;
	cfg.numAttrs = 1;
	return cudaLaunchKernelEx(&cfg, hier2_kernel<K, WRITE, PIPE, SUB, INBOX>, p);
}

cudaError_t launch_hier2(int K, bool write, const HierParams& p, cudaStream_t stream)
{
	if (p.ncb == 0 || p.seg_end <= p.seg_begin)
		return cudaSuccess;
	const bool sub = p.parts > 1;
	if (sub && p.pipe != 0)
		return cudaErrorInvalidValue; // the host keeps sub-tile segments to the plain walk
	const bool pipe = p.pipe != 0;
	if (p.tile_ids) // routed frame: whole-tile segments
	{
		if (sub)
			return cudaErrorInvalidValue;
		switch (K * 4 + (write ? 1 : 0) + (pipe ? 2 : 0))
		{
		case 4: return launch_hier2_t<1, false, false, false, true>(p, stream);
		case 5: return launch_hier2_t<1, true, false, false, true>(p, stream);
		case 6: return launch_hier2_t<1, false, true, false, true>(p, stream);
		case 7: return launch_hier2_t<1, true, true, false, true>(p, stream);
		case 8: return launch_hier2_t<2, false, false, false, true>(p, stream);
		case 9: return launch_hier2_t<2, true, false, false, true>(p, stream);
		case 10: return launch_hier2_t<2, false, true, false, true>(p, stream);
		case 11: return launch_hier2_t<2, true, true, false, true>(p, stream);
		case 16: return launch_hier2_t<4, false, false, false, true>(p, stream);
		case 17: return launch_hier2_t<4, true, false, false, true>(p, stream);
		case 18: return launch_hier2_t<4, false, true, false, true>(p, stream);
		case 19: return launch_hier2_t<4, true, true, false, true>(p, stream);
		default: return cudaErrorInvalidValue;
		}
	}
	switch (K * 8 + (write ? 1 : 0) + (pipe ? 2 : 0) + (sub ? 4 : 0))
	{
	case 8: return launch_hier2_t<1, false, false, false>(p, stream);
	case 9: return launch_hier2_t<1, true, false, false>(p, stream);
	case 10: return launch_hier2_t<1, false, true, false>(p, stream);
	case 11: return launch_hier2_t<1, true, true, false>(p, stream);
	case 12: return launch_hier2_t<1, false, false, true>(p, stream);
	case 13: return launch_hier2_t<1, true, false, true>(p, stream);
	case 16: return launch_hier2_t<2, false, false, false>(p, stream);
	case 17: return launch_hier2_t<2, true, false, false>(p, stream);
	case 18: return launch_hier2_t<2, false, true, false>(p, stream);
	case 19: return launch_hier2_t<2, true, true, false>(p, stream);
	case 20: return launch_hier2_t<2, false, false, true>(p, stream);
	case 21: return launch_hier2_t<2, true, false, true>(p, stream);
	case 32: return launch_hier2_t<4, false, false, false>(p, stream);
	case 33: return launch_hier2_t<4, true, false, false>(p, stream);
	case 34: return launch_hier2_t<4, false, true, false>(p, stream);
	case 35: return launch_hier2_t<4, true, true, false>(p, stream);
	case 36: return launch_hier2_t<4, false, false, true>(p, stream);
	case 37: return launch_hier2_t<4, true, false, true>(p, stream);
	default: return cudaErrorInvalidValue;
	}
}


// every kernel of this file that a frame can launch, force-loaded (HVCF_PRELOAD, hvcf_device.cuh)

cudaError_t preload_bin_kernels()
{
	HVCF_PRELOAD(hier_unions_kernel);
	HVCF_PRELOAD(hier_alive_kernel);
	HVCF_PRELOAD(hier_super_unions_kernel);
	HVCF_PRELOAD(hier_bin_kernel);
	HVCF_PRELOAD(hier_bin2_kernel<false>);
	HVCF_PRELOAD(hier_bin2_kernel<true>);
	HVCF_PRELOAD(hier_xbin_kernel);
	HVCF_PRELOAD((hier2_kernel<1, false, false, false>));
	HVCF_PRELOAD((hier2_kernel<1, false, true, false>));
	HVCF_PRELOAD((hier2_kernel<1, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<1, false, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<1, false, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<1, true, false, false>));
	HVCF_PRELOAD((hier2_kernel<1, true, true, false>));
	HVCF_PRELOAD((hier2_kernel<1, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<1, true, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<1, true, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<2, false, false, false>));
	HVCF_PRELOAD((hier2_kernel<2, false, true, false>));
	HVCF_PRELOAD((hier2_kernel<2, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<2, false, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<2, false, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<2, true, false, false>));
	HVCF_PRELOAD((hier2_kernel<2, true, true, false>));
	HVCF_PRELOAD((hier2_kernel<2, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<2, true, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<2, true, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<4, false, false, false>));
	HVCF_PRELOAD((hier2_kernel<4, false, true, false>));
	HVCF_PRELOAD((hier2_kernel<4, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<4, false, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<4, false, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<4, true, false, false>));
	HVCF_PRELOAD((hier2_kernel<4, true, true, false>));
	HVCF_PRELOAD((hier2_kernel<4, true, false, true>));
	HVCF_PRELOAD((hier2_kernel<4, true, false, false, true>));
	HVCF_PRELOAD((hier2_kernel<4, true, true, false, true>));
	return cudaSuccess;
}

} // namespace hvcf
