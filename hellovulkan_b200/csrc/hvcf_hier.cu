// hvcf_hier.cu -- HVCF_MODE_HIERARCHICAL: hierarchical light culling that produces the same bytes as the
// brute force path (and as LightCulling.comp:59-88) while running a small fraction of its tests.
// This file holds the scheme's description and proof sketch, the finalize kernel and the host-side plans; the
// level-0 binning and the count / write passes are in hvcf_bin.cu.
//
// The hierarchy is over CLUSTERS, not lights, so light order -- and with it the reference's "first 150
// hits in ascending light index" -- is preserved for free:
//
//   level 0  group: union box of a cluster group (8 warp programs x K blocks x 32 clusters).  Every light of a
//                   TMA-staged tile is tested against it (lane = light pair, FADD2/FMUL2); survivors are compacted,
//                   in light order, into the (group, tile) slot (hvcf_bin.cu: hier_bin_kernel / hier_bin2_kernel).
//   level 1  block: union box of one 8x4x1 block of clusters.  Lane = surviving light; ballot -> candidates.
//   level 2  exact: lane = cluster.  Each candidate light is broadcast from shared memory and tested
//                   against the lane's own AABB with the reference's arithmetic.
//
// The union tests use the SAME uncontracted float sequence as the exact test.  Rounding is monotone, so
// lo_u <= lo_c and hi_u >= hi_c give d_u <= d_c per axis and sq_u <= sq_c in floating point: a light that
// hits a cluster always passes the unions -- the filter is exactly conservative, no epsilon.
//
// Ordered output without scratch lists: pass 1 (WRITE = false) only COUNTS hits per (light segment,
// cluster); hier_finalize_kernel turns the counts into per-segment prefixes, capped counts and an
// exclusive scan over clusters (single pass, decoupled look-back); pass 2 (WRITE = true) re-runs the
// hierarchy over the lanes that own output (tight unions, so it is far cheaper than pass 1) and stores
// each index straight at offset[cluster] + prefix[segment] + k.  No global atomic append, no sorting.
//
// Early exit (LightCulling.comp:71, `lightIter < MAX_LIGHTS_PER_CLUSTER`) across concurrently running
// segments: segments are grouped into chunks of doubling size; a finishing CTA adds its counts to
// cum[cluster][chunk]; a starting CTA sums cum over the EARLIER chunks and drops every cluster that already
// holds `cap` hits.  Whatever it reads is a lower bound of the hits in earlier segments (counts only grow
// towards the true value), and a lower bound >= cap proves that the first `cap` entries lie before this
// segment -- so knowledge only ever removes work that cannot matter, and the result does not depend on
// scheduling.  No flags, no fences.
#include "hvcf_hier.cuh"

#include <algorithm>

#include <vector>

namespace hvcf
{

constexpr uint32_t kHWarps = 8; // warp programs per cluster group

// ======================================================================================================
// finalize: per-segment counts -> per-segment prefixes, capped counts, exclusive scan over clusters.
// CTA = 32 clusters (lane) x 8 segment groups (warp): every count is loaded exactly once, all loads of a
// thread are independent, rows are 128 B coalesced.  The scan across CTAs is a single pass with a
// warp-wide decoupled look-back; the state words carry the frame's epoch, so they never need clearing.
// ======================================================================================================
constexpr uint32_t kFinWarps = 8;

__global__ void __launch_bounds__(kFinWarps * 32) hier_finalize_kernel(const FinalizeParams p)
{
	__shared__ uint32_t s_part[kFinWarps][32];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t c = blockIdx.x * 32 + lane;
	const bool valid = c < p.cluster_count;
	grid_dep_launch();
	grid_dep_wait();
	// rounds: this cluster's counts are valid (and needed) only up to the round before which it was full
	uint32_t seg_limit = p.num_segs;
	if (p.num_rounds >= 2 && valid)
	{
		const uint4* row = reinterpret_cast<const uint4*>(p.cum + static_cast<size_t>(c) * kHMaxChunks);
		const uint4 a = __ldcg(row), b = __ldcg(row + 1);
		const uint32_t h[kHMaxChunks] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
		uint32_t before = 0;
#pragma unroll
		for (uint32_t r = 0; r < kHMaxChunks; ++r)
		{
			if (r < p.num_rounds && before >= p.cap && seg_limit == p.num_segs)
				seg_limit = p.round_seg_begin[r];
			before += h[r];
		}
	}
	const uint32_t group = (p.num_segs + kFinWarps - 1) / kFinWarps;
	const uint32_t s0 = min(min(warp * group, p.num_segs), seg_limit), s1 = min(min(warp * group + group, p.num_segs), seg_limit);
	// two sweeps over this warp's segments instead of holding all of them in registers (the 32-deep register array cost
	// 120 registers = 2 CTAs per SM; the counts were written by the previous kernel and sit in L2, so the second read is
	// cheap and the occupancy is what hides the latency): sweep 1 sums, sweep 2 turns counts into prefixes
	const uint32_t* col = p.seg_counts + c;
	uint32_t part = 0;
	if (valid)
	{
		uint32_t s = s0;
		for (; s + 4 <= s1; s += 4)
		{
			const uint32_t a = col[static_cast<size_t>(s) * p.cluster_count], b = col[static_cast<size_t>(s + 1) * p.cluster_count];
			const uint32_t d = col[static_cast<size_t>(s + 2) * p.cluster_count], e = col[static_cast<size_t>(s + 3) * p.cluster_count];
			part += (a + b) + (d + e);
		}
		for (; s < s1; ++s)
			part += col[static_cast<size_t>(s) * p.cluster_count];
	}
	s_part[warp][lane] = part;
	__syncthreads();
	uint32_t run = 0, total = 0;
#pragma unroll
	for (uint32_t w = 0; w < kFinWarps; ++w)
	{
		const uint32_t v = s_part[w][lane];
		run += w < warp ? v : 0u;
		total += v;
	}
	// warp 0 publishes this CTA's aggregate BEFORE the second sweep: the CTAs behind it in the look-back chain need nothing else
	// from us, so their wait shrinks by the length of our sweep (A65k: the chain is the kernel)
	const uint32_t epoch = p.stats->epoch; // bumped by the first kernel of this frame (DeviceStats)
	const unsigned long long tag = static_cast<unsigned long long>(epoch) << 34;
	uint32_t v = 0, inc = 0, agg = 0, over = 0;
	if (warp == 0)
	{
		v = valid ? min(total, p.cap) : 0u;
		inc = v;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1)
		{
			const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
			if (lane >= static_cast<uint32_t>(d))
				inc += t;
		}
		agg = __shfl_sync(0xffffffffu, inc, 31);
		over = __popc(__ballot_sync(0xffffffffu, valid && v >= p.cap));
		// state word: [epoch:30 | flag:2 | value:32]; flag 1 = aggregate of this CTA, 2 = inclusive prefix
		if (lane == 0)
			st_release_u64(&p.scan_state[blockIdx.x], tag | ((blockIdx.x == 0 ? 2ull : 1ull) << 32) | agg);
	}
	if (valid)
	{
		uint32_t* pre = p.seg_prefix + c;
		uint32_t s = s0;
		for (; s + 4 <= s1; s += 4)
		{
			const uint32_t a = col[static_cast<size_t>(s) * p.cluster_count], b = col[static_cast<size_t>(s + 1) * p.cluster_count];
			const uint32_t d = col[static_cast<size_t>(s + 2) * p.cluster_count], e = col[static_cast<size_t>(s + 3) * p.cluster_count];
			pre[static_cast<size_t>(s) * p.cluster_count] = (a == 0 || run >= p.cap) ? kSkip : run;
			run += a;
			pre[static_cast<size_t>(s + 1) * p.cluster_count] = (b == 0 || run >= p.cap) ? kSkip : run;
			run += b;
			pre[static_cast<size_t>(s + 2) * p.cluster_count] = (d == 0 || run >= p.cap) ? kSkip : run;
			run += d;
			pre[static_cast<size_t>(s + 3) * p.cluster_count] = (e == 0 || run >= p.cap) ? kSkip : run;
			run += e;
		}
		for (; s < s1; ++s)
		{
			const uint32_t a = col[static_cast<size_t>(s) * p.cluster_count];
			pre[static_cast<size_t>(s) * p.cluster_count] = (a == 0 || run >= p.cap) ? kSkip : run;
			run += a;
		}
	}
	if (warp != 0)
		return;
	uint32_t prefix = 0;
	for (int base = static_cast<int>(blockIdx.x); base > 0; base -= 32)
	{
		const int i = base - 1 - static_cast<int>(lane); // lane 0 looks at the nearest predecessor
		unsigned long long w;
		bool ok;
		do
		{
			w = i >= 0 ? ld_acquire_u64(&p.scan_state[i]) : (tag | (2ull << 32));
			ok = (w >> 34) == epoch && ((w >> 32) & 3ull) != 0;
		} while (!__all_sync(0xffffffffu, ok));
		const uint32_t inclusive = __ballot_sync(0xffffffffu, ((w >> 32) & 3ull) == 2);
		const uint32_t upto = inclusive ? static_cast<uint32_t>(__ffs(inclusive)) - 1u : 31u; // nearest inclusive prefix ends the walk
		prefix += __reduce_add_sync(0xffffffffu, lane <= upto ? static_cast<uint32_t>(w) : 0u);
		if (inclusive)
			break;
	}
	if (lane == 0)
	{
		if (blockIdx.x != 0)
			st_release_u64(&p.scan_state[blockIdx.x], tag | (2ull << 32) | (prefix + agg));
		if (over)
			atomicAdd(&p.stats->overflow_clusters, over);
		if (blockIdx.x == gridDim.x - 1)
		{
			*p.shard_total = prefix + agg;
			p.stats->total_indices = prefix + agg;
		}
	}
	if (valid)
	{
		p.counts[c] = v;
		p.offsets[c] = prefix + inc - v;
	}
}

// ======================================================================================================
// host side: plan + launchers
// ======================================================================================================
static uint32_t pow2_ceil(uint32_t v)
{
	uint32_t r = 1;
	while (r < v)
		r <<= 1;
	return r;
}

void hier_plan_grid(uint32_t X, uint32_t Y, uint32_t Zs, int k_override, const uint32_t shape_override[3], uint32_t max_lights, HierPlan* out)
{
	HierPlan& h = *out;
	h.bx = pow2_ceil(X) < 8 ? pow2_ceil(X) : 8;
	const uint32_t rem = 32 / h.bx;
	const uint32_t by_pref = h.bx == 8 ? 4 : rem;
	h.by = pow2_ceil(Y) < by_pref ? pow2_ceil(Y) : by_pref;
	h.bz = rem / h.by;
	if (shape_override && shape_override[0] * shape_override[1] * shape_override[2] == 32)
		h.bx = shape_override[0], h.by = shape_override[1], h.bz = shape_override[2];
	h.nbx = (X + h.bx - 1) / h.bx;
	h.nby = (Y + h.by - 1) / h.by;
	h.nbz = (Zs + h.bz - 1) / h.bz;
	h.nblocks = h.nbx * h.nby * h.nbz;
	h.K = h.nblocks >= 2048 ? 4 : (h.nblocks >= 512 ? 2 : 1);
	// with millions of lights the per-(group, light) binning cost outweighs the finer filtering of small groups: bigger
	// groups earlier (measured on z-slabs of grid C with 16M lights, scripts/slab_sweep.py: 360 and 504 blocks K=2 best,
	// 1,584 and 1,800 blocks K=4 best)
	if (max_lights >= (1u << 22))
		h.K = h.nblocks >= 1024 ? 4 : (h.nblocks >= 128 ? 2 : 1);
	if (k_override == 1 || k_override == 2 || k_override == 4)
		h.K = static_cast<uint32_t>(k_override);
	h.ncb = (h.nblocks + kHWarps * h.K - 1) / (kHWarps * h.K);
}

// cluster index (inside the shard) of every lane of every block, kSkip where the block sticks out of the grid or the cluster
// is not the shard's (row-granular shards: the box covers whole slices, the shard may start and end inside one); blocks
// that keep no lane are dropped, the rest keep their order (x, then y, then z: neighbours stay neighbours in a group);
// padded to whole CTAs so the kernel needs no bounds check
void hier_build_lane_map(HierPlan* hp, uint32_t X, uint32_t Y, uint32_t Zs, uint32_t cluster_lo, uint32_t cluster_hi, std::vector<uint32_t>* out)
{
	HierPlan& h = *hp;
	// Block order.  A warp program takes K consecutive blocks and a group 8 K consecutive blocks, so the order decides the shape of
	// the level-0 boxes.  Plain x-then-y order makes a group of a wide grid a flat strip across the whole width (grid C: 64 x 16
	// clusters); with `brick` the blocks of a slice are visited brick by brick (4 x 2K blocks = 32 x 8K clusters for 8 x 4 blocks,
	// 2 x 2 blocks inside a brick first), so a group is compact in view space.  That matters most in the late rounds of a long
	// frame: the clusters still alive there are the ones at the left and right edge of the lit region, and a strip across the whole
	// width cannot shrink to them while a brick at the edge can (profiles/r02q: 8 strips took 4.1 M records in the last round).
	std::vector<uint32_t> order(h.nblocks);
	for (uint32_t b = 0; b < h.nblocks; ++b)
		order[b] = b;
	const bool brick = h.brick && h.nbx >= 8;
	const uint32_t gx = 4, gy = (8 * h.K) / gx >= 1 ? (8 * h.K) / gx : 1; // blocks per brick
	const uint32_t ux = h.K >= 2 ? 2 : 1, uy = h.K >= 4 ? 2 : 1;          // blocks of one warp program
	auto brick_of = [&](uint32_t b) {
		const uint64_t bxi = b % h.nbx, byi = (b / h.nbx) % h.nby, bzi = b / (h.nbx * h.nby);
		return ((byi / gy) * 4096 + bxi / gx) * 65536 + bzi; // brick position first, slice last: consecutive groups -- and with them
		                                                     // the super-groups of the two-level binning -- are the same brick in
		                                                     // consecutive slices, so a super-group at the edge of the lit region
		                                                     // stays at the edge (its box rejects most lights in phase A)
	};
	if (brick)
	{
		auto key = [&](uint32_t b) {
			const uint64_t bxi = b % h.nbx, byi = (b / h.nbx) % h.nby;
			uint64_t k = brick_of(b);
			k = k * 64 + (byi % gy) / uy;
			k = k * 64 + (bxi % gx) / ux;
			k = k * 8 + byi % uy;
			k = k * 8 + bxi % ux;
			return k;
		};
		std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return key(a) < key(b); });
	}
	const uint32_t group_blocks = kHWarps * h.K;
	std::vector<uint32_t> kept; // 32 lanes per kept (or padding) block
	kept.reserve(static_cast<size_t>(h.nblocks) * 32);
	std::vector<uint32_t> cur;  // the lanes of the current brick's kept blocks
	uint32_t nkept = 0;         // blocks emitted so far, padding included
	auto flush_brick = [&]() {
		const uint32_t nb = static_cast<uint32_t>(cur.size() / 32);
		if (nb == 0)
			return;
		// a brick that does not fit into what is left of the current group starts a new one (the gap is padding: blocks without
		// lanes, whose warp programs leave at once).  Otherwise 72 blocks per slice against 32 per group would smear every brick
		// over two groups, and a slice's ragged bottom row would share a group with the next slice's first brick.
		const uint32_t used = nkept % group_blocks;
		if (used != 0 && used + nb > group_blocks && nb <= group_blocks)
		{
			kept.insert(kept.end(), static_cast<size_t>(group_blocks - used) * 32, kSkip);
			nkept += group_blocks - used;
		}
		kept.insert(kept.end(), cur.begin(), cur.end());
		nkept += nb;
		cur.clear();
	};
	uint64_t cur_brick = ~0ull;
	for (uint32_t oi = 0; oi < h.nblocks; ++oi)
	{
		const uint32_t b = order[oi];
		if (brick && brick_of(b) != cur_brick)
		{
			flush_brick();
			cur_brick = brick_of(b);
		}
		const uint32_t bxi = b % h.nbx, byi = (b / h.nbx) % h.nby, bzi = b / (h.nbx * h.nby);
		uint32_t lanes[32];
		bool any = false;
		for (uint32_t lane = 0; lane < 32; ++lane)
		{
			const uint32_t x = bxi * h.bx + lane % h.bx, y = byi * h.by + (lane / h.bx) % h.by, z = bzi * h.bz + lane / (h.bx * h.by);
			lanes[lane] = kSkip;
			if (x < X && y < Y && z < Zs)
			{
				const uint32_t c = x + y * X + z * X * Y;
				if (c >= cluster_lo && c < cluster_hi)
				{
					lanes[lane] = c - cluster_lo;
					any = true;
				}
			}
		}
		if (any)
		{
			if (brick)
				cur.insert(cur.end(), lanes, lanes + 32);
			else
			{
				kept.insert(kept.end(), lanes, lanes + 32);
				++nkept;
			}
		}
	}
	flush_brick();
	h.nblocks = nkept;
	h.ncb = (nkept + group_blocks - 1) / group_blocks;
	if (h.ncb == 0)
		h.ncb = 1;
	out->assign(static_cast<size_t>(h.ncb) * group_blocks * 32, kSkip);
	std::copy(kept.begin(), kept.end(), out->begin());
}

void hier_plan_lights(const HierPlan& h, uint32_t tiles_total, uint32_t max_segs, int tps_override, uint32_t* tiles_per_seg,
	uint32_t* num_segs)
{
	// about 96 eight-unit groups per SM over the whole launch (measured on B200: short segments win -- more CTAs in flight hide the
	// per-tile barrier latency and earlier segments publish their counts sooner); segments are whole tiles
	const uint64_t target = 148ull * 96ull;
	const uint64_t tps = (static_cast<uint64_t>(tiles_total) * h.ncb) / target;
	uint32_t t = 1;
	while (static_cast<uint64_t>(t) * 2 <= tps)
		t <<= 1;
	if (tps_override > 0)
		t = static_cast<uint32_t>(tps_override);
	if (max_segs == 0)
		max_segs = 1;
	while ((tiles_total + t - 1) / t > max_segs)
		t <<= 1;
	*tiles_per_seg = t;
	*num_segs = (tiles_total + t - 1) / t;
}

uint32_t hier_num_chunks(uint32_t num_segs)
{
	if (num_segs <= 4)
		return 1;
	uint32_t c = 0, s = num_segs - 1;
	while (s >= 8)
	{
		s >>= 1;
		++c;
	}
	return c + 2 < kHMaxChunks ? c + 2 : kHMaxChunks; // chunk_of(num_segs - 1) + 1
}

cudaError_t launch_hier_finalize(const FinalizeParams& p, cudaStream_t stream)
{
	if (p.cluster_count == 0)
		return cudaSuccess;
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = dim3((p.cluster_count + 31) / 32, 1, 1);
	cfg.blockDim = dim3(kFinWarps * 32, 1, 1);
	cfg.stream = stream;
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = attr;
	cfg.numAttrs = 1;
	return cudaLaunchKernelEx(&cfg, hier_finalize_kernel, p);
}


// every kernel of this file that a frame can launch, force-loaded (HVCF_PRELOAD, hvcf_device.cuh)

cudaError_t preload_hier_kernels()
{
	HVCF_PRELOAD(hier_finalize_kernel);
	return cudaSuccess;
}

} // namespace hvcf
