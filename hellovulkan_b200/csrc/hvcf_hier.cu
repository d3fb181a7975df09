// hvcf_hier.cu -- HVCF_MODE_HIERARCHICAL: hierarchical light culling that produces the same bytes as the
// brute force path (and as LightCulling.comp:59-88) while running a small fraction of its tests.
//
// The hierarchy is over CLUSTERS, not lights, so light order -- and with it the reference's "first 150
// hits in ascending light index" -- is preserved for free:
//
//   level 0  CTA  : union box of all its clusters (8 warps x K blocks x 32 clusters).  Every light of a
//                   TMA-staged tile is tested once per CTA (lane = light pair, FADD2/FMUL2); survivors are
//                   compacted, in light order, into shared memory (__ballot_sync/__popc + a CTA prefix).
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

constexpr uint32_t kHWarps = 8;
constexpr uint32_t kHStages = 2;
constexpr uint32_t kHGroupsPerWarp = kTileGroups / kHWarps; // 2
static_assert(kTileGroups % kHWarps == 0, "level 0 splits a tile's groups evenly over the warps");

// dynamic shared memory: light tile ring | survivor records (2 buffers) | survivor local indices (2 buffers)
constexpr uint32_t kHSmemRing = kHStages * kTileBytes;
constexpr uint32_t kHSmemS0f = 2 * kTileLights * 16;
constexpr uint32_t kHSmemS0i = 2 * kTileLights * 2;
constexpr uint32_t kHSmemBytes = kHSmemRing + kHSmemS0f + kHSmemS0i;

template <int K, bool WRITE>
__global__ void __launch_bounds__(kHWarps * 32) hier_kernel(const HierParams p)
{
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ __align__(8) uint64_t full_bar[kHStages];
	__shared__ uint32_t s_warp_n[2][kHWarps];
	__shared__ float s_union[kHWarps][6];

	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t cb = blockIdx.x % p.ncb, seg = p.seg_begin + blockIdx.x / p.ncb;
	const uint32_t t_begin = seg * p.tiles_per_seg;
	const uint32_t t_end = min(t_begin + p.tiles_per_seg, p.tiles_total);
	const uint32_t ntiles = t_end - t_begin;
	const unsigned char* src = reinterpret_cast<const unsigned char*>(p.lights_view) + static_cast<size_t>(t_begin) * kTileBytes;
	float4* const s0f = reinterpret_cast<float4*>(smem + kHSmemRing);
	uint16_t* const s0i = reinterpret_cast<uint16_t*>(smem + kHSmemRing + kHSmemS0f);

	grid_dep_launch();
	if (threadIdx.x == 0)
	{
#pragma unroll
		for (uint32_t s = 0; s < kHStages; ++s)
			mbar_init(&full_bar[s], 1);
		mbar_fence_init();
	}
	__syncthreads();
	// everything above overlaps the tail of the previous kernel of the frame (PDL); everything below reads what it wrote
	grid_dep_wait();
	unsigned long long t_start = 0;
	if (p.trace && threadIdx.x == 0)
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
	if (threadIdx.x == 0)
	{
		for (uint32_t s = 0; s < kHStages && s < ntiles; ++s)
		{
			mbar_arrive_expect_tx(&full_bar[s], kTileBytes);
			bulk_copy_g2s(smem + s * kTileBytes, src + static_cast<size_t>(s) * kTileBytes, kTileBytes, &full_bar[s]);
		}
	}

	// ---- this lane's K clusters (one per block), while the first tiles are in flight ----------------
	const float kInf = __int_as_float(0x7f800000);
	float lo[K][3], hi[K][3];
	uint32_t cidx[K];    // cluster index inside the shard, kSkip = no cluster on this lane
	uint32_t cnt[K];     // pass 1: hits counted in this segment; pass 2: entries written
	uint32_t limit[K];   // pass 1: hits still missing before the cap; pass 2: entries this lane owns
	uint32_t wbase[K];   // pass 2: first output slot
	bool active[K];
	const uint32_t chunk = hier_chunk_of(seg);
	const uint32_t index_base = (WRITE && p.index_base) ? *p.index_base : 0u;
#pragma unroll
	for (int k = 0; k < K; ++k)
	{
		// block -> lane -> cluster map, built once per context on the host (hier_build_lane_map)
		cidx[k] = __ldg(&p.lane_cluster[(static_cast<size_t>((cb * kHWarps + warp) * K + k) << 5) + lane]);
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
	if (WRITE && blockIdx.x == 0 && threadIdx.x == 0 && p.frame_total)
		*p.frame_total = index_base + *p.shard_total; // globalIndex after the frame (LightCulling.comp:21,79)

	// ---- union boxes: per block (level 1), then per CTA (level 0) ---------------------------------------
	float ulo[K][3], uhi[K][3];
	bool blk_active[K];
	uint32_t alive[K]; // lanes of the block that still want hits
	bool blk_nan[K];   // the block holds a NaN box: its union stays 'pass everything'
	float wlo0 = kInf, wlo1 = kInf, wlo2 = kInf, whi0 = -kInf, whi1 = -kInf, whi2 = -kInf;
#pragma unroll
	for (int k = 0; k < K; ++k)
	{
		alive[k] = __ballot_sync(0xffffffffu, active[k]);
		blk_active[k] = alive[k] != 0;
		// a NaN bound makes every light hit that cluster (both shader comparisons are false): pass everything
		const bool nan = active[k] && !(lo[k][0] == lo[k][0] && lo[k][1] == lo[k][1] && lo[k][2] == lo[k][2] &&
			hi[k][0] == hi[k][0] && hi[k][1] == hi[k][1] && hi[k][2] == hi[k][2]);
		const bool any_nan = __any_sync(0xffffffffu, nan);
		blk_nan[k] = any_nan;
#pragma unroll
		for (int i = 0; i < 3; ++i)
		{
			ulo[k][i] = any_nan ? -kInf : warp_min_f(lo[k][i]);
			uhi[k][i] = any_nan ? kInf : warp_max_f(hi[k][i]);
		}
		wlo0 = fminf(wlo0, ulo[k][0]), wlo1 = fminf(wlo1, ulo[k][1]), wlo2 = fminf(wlo2, ulo[k][2]);
		whi0 = fmaxf(whi0, uhi[k][0]), whi1 = fmaxf(whi1, uhi[k][1]), whi2 = fmaxf(whi2, uhi[k][2]);
	}
	if (lane == 0)
	{
		s_union[warp][0] = wlo0, s_union[warp][1] = wlo1, s_union[warp][2] = wlo2;
		s_union[warp][3] = whi0, s_union[warp][4] = whi1, s_union[warp][5] = whi2;
	}
	bool warp_done = true;
#pragma unroll
	for (int k = 0; k < K; ++k)
		warp_done = warp_done && !blk_active[k];
	__syncthreads();
	float clo[3] = {kInf, kInf, kInf}, chi[3] = {-kInf, -kInf, -kInf};
#pragma unroll
	for (uint32_t w = 0; w < kHWarps; ++w)
#pragma unroll
		for (int i = 0; i < 3; ++i)
		{
			clo[i] = fminf(clo[i], s_union[w][i]);
			chi[i] = fmaxf(chi[i], s_union[w][3 + i]);
		}

	const uint32_t lt = lanemask_lt();
	unsigned long long n_l0 = 0, n_l1 = 0, n_l2 = 0; // warp-uniform test counters (lane tests)
	uint32_t it = 0;
	for (; it < ntiles; ++it)
	{
		const uint32_t buf = it % kHStages, sb = it & 1u;
		mbar_wait(&full_bar[buf], (it / kHStages) & 1u);
		// ---- level 0: this warp's two groups of the tile against the CTA union ---------------------------
		const float4* tile = reinterpret_cast<const float4*>(smem + buf * kTileBytes);
		float4 ra[kHGroupsPerWarp], rb[kHGroupsPerWarp];
		uint32_t m0[kHGroupsPerWarp], m1[kHGroupsPerWarp];
		uint32_t n_w = 0;
#pragma unroll
		for (uint32_t q = 0; q < kHGroupsPerWarp; ++q)
		{
			const uint32_t g = warp * kHGroupsPerWarp + q;
			ra[q] = tile[g * 64 + lane];
			rb[q] = tile[g * 64 + 32 + lane];
			bool h0, h1;
			sphere_box_hit2(ra[q], rb[q], clo, chi, h0, h1);
			m0[q] = __ballot_sync(0xffffffffu, h0);
			m1[q] = __ballot_sync(0xffffffffu, h1);
			n_w += __popc(m0[q]) + __popc(m1[q]);
		}
		if (lane == 0)
			s_warp_n[sb][warp] = n_w;
		// barrier 1: survivor counts are published, every warp is done reading ring slot `buf`, and the
		// CTA-wide early-exit vote of the previous tile is taken
		if (__syncthreads_and(warp_done))
			break;
		if (threadIdx.x == 0 && it + kHStages < ntiles)
		{
			mbar_arrive_expect_tx(&full_bar[buf], kTileBytes);
			bulk_copy_g2s(smem + buf * kTileBytes, src + static_cast<size_t>(it + kHStages) * kTileBytes, kTileBytes, &full_bar[buf]);
		}
		uint32_t pos = 0, n0 = 0;
#pragma unroll
		for (uint32_t w = 0; w < kHWarps; ++w)
		{
			const uint32_t nw = s_warp_n[sb][w];
			pos += w < warp ? nw : 0u;
			n0 += nw;
		}
		n_l0 += kTileLights / kHWarps;
		if (n0 == 0)
			continue; // uniform over the CTA: nobody reaches barrier 2 of this tile
		float4* const sf = s0f + sb * kTileLights;
		uint16_t* const si = s0i + sb * kTileLights;
#pragma unroll
		for (uint32_t q = 0; q < kHGroupsPerWarp; ++q)
		{
			const uint32_t g = warp * kHGroupsPerWarp + q;
			if (m0[q] & (1u << lane))
			{
				const uint32_t d = pos + __popc(m0[q] & lt);
				sf[d] = make_float4(ra[q].x, ra[q].z, rb[q].x, rb[q].z);
				si[d] = static_cast<uint16_t>(g * 64 + lane);
			}
			pos += __popc(m0[q]);
			if (m1[q] & (1u << lane))
			{
				const uint32_t d = pos + __popc(m1[q] & lt);
				sf[d] = make_float4(ra[q].y, ra[q].w, rb[q].y, rb[q].w);
				si[d] = static_cast<uint16_t>(g * 64 + 32 + lane);
			}
			pos += __popc(m1[q]);
		}
		__syncthreads(); // barrier 2: the survivor list of this tile is complete
		if (warp_done)
			continue;
		// ---- level 1 (lane = surviving light) and level 2 (lane = cluster) ---------------------------------
		const uint32_t tile_base = (t_begin + it) * kTileLights;
		for (uint32_t i = 0; i < n0 && !warp_done; i += 32)
		{
			const uint32_t idx = i + lane;
			const float4 r = idx < n0 ? sf[idx] : make_float4(0.0f, 0.0f, 0.0f, -1.0f);
			warp_done = true;
#pragma unroll
			for (int k = 0; k < K; ++k)
			{
				if (!blk_active[k])
					continue;
				uint32_t m = __ballot_sync(0xffffffffu, sphere_box_hit(r.x, r.y, r.z, r.w, ulo[k], uhi[k]));
				n_l1 += 32;
				n_l2 += 32u * __popc(m);
				// kCand candidates per trip: a trip is latency-bound (LDS -> 6-deep FP chain), so independent chains are free
				while (m)
				{
					uint32_t j[kCand];
					bool take[kCand];
#pragma unroll
					for (uint32_t q = 0; q < kCand; ++q)
					{
						take[q] = m != 0;
						j[q] = take[q] ? i + __ffs(m) - 1 : j[0];
						m &= m - 1; // 0 & 0xffffffff == 0 once the mask is empty
					}
					float4 c[kCand];
#pragma unroll
					for (uint32_t q = 0; q < kCand; ++q)
						c[q] = sf[j[q]]; // broadcasts
#pragma unroll
					for (uint32_t q = 0; q < kCand; ++q)
					{
						const bool hit = sphere_box_hit(c[q].x, c[q].y, c[q].z, c[q].w, lo[k], hi[k]) && active[k] && take[q];
						if (WRITE)
						{
							if (hit && cnt[k] < limit[k])
							{
								p.indices[wbase[k] + cnt[k]] = tile_base + si[j[q]];
								++cnt[k];
							}
						}
						else
							cnt[k] += hit ? 1u : 0u;
					}
				}
				// early exit inside the tile, per lane: a cluster that holds its quota leaves the block's union box,
				// so the remaining lanes see fewer candidates; a block with no lane left stops here
				const uint32_t now_alive = __ballot_sync(0xffffffffu, active[k] && cnt[k] < limit[k]);
				if (now_alive != alive[k])
				{
					alive[k] = now_alive;
					active[k] = (now_alive >> lane) & 1u;
					blk_active[k] = now_alive != 0;
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
				warp_done = warp_done && !blk_active[k];
			}
		}
	}
	// bulk copies still in flight must land before this CTA's shared memory is handed on
	if (threadIdx.x == 0)
	{
		// tiles [0, issued) were requested; tiles [0, it] have been waited for (tile `it` only if the loop broke)
		const uint32_t issued = min(it + kHStages, ntiles);
		for (uint32_t j = it + 1; j < issued; ++j)
			mbar_wait(&full_bar[j % kHStages], (j / kHStages) & 1u);
	}
	if (!WRITE)
	{
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
			unsigned long long* t = p.trace + 4ull * (static_cast<size_t>(WRITE ? gridDim.x : 0) + blockIdx.x);
			t[0] = t_start, t[1] = t_end, t[2] = smid, t[3] = n_l2;
		}
	}
	if (lane == 0 && (n_l0 | n_l1 | n_l2))
	{
		if (WRITE)
			atomicAdd(&p.stats->write_tests, n_l0 + n_l1 + n_l2);
		else
		{
			atomicAdd(&p.stats->prefilter_tests, n_l0 + n_l1);
			atomicAdd(&p.stats->executed_tests, n_l2);
		}
	}
}

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
	const uint32_t group = (p.num_segs + kFinWarps - 1) / kFinWarps;
	const uint32_t s0 = min(warp * group, p.num_segs), s1 = min(s0 + group, p.num_segs);
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
	if (valid && !p.skip_prefix)
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
	std::vector<uint32_t> kept;
	kept.reserve(static_cast<size_t>(h.nblocks) * 32);
	uint32_t nkept = 0;
	for (uint32_t b = 0; b < h.nblocks; ++b)
	{
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
			kept.insert(kept.end(), lanes, lanes + 32);
			++nkept;
		}
	}
	h.nblocks = nkept;
	h.ncb = (nkept + kHWarps * h.K - 1) / (kHWarps * h.K);
	if (h.ncb == 0)
		h.ncb = 1;
	out->assign(static_cast<size_t>(h.ncb) * kHWarps * h.K * 32, kSkip);
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

template <int K, bool WRITE>
static cudaError_t launch_hier_t(const HierParams& p, cudaStream_t stream)
{
	static bool configured = false;
	if (!configured)
	{
		cudaError_t e = cudaFuncSetAttribute(hier_kernel<K, WRITE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
			static_cast<int>(kHSmemBytes));
		if (e != cudaSuccess)
			return e;
		configured = true;
	}
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = dim3(p.ncb * (p.seg_end - p.seg_begin), 1, 1);
	cfg.blockDim = dim3(kHWarps * 32, 1, 1);
	cfg.dynamicSmemBytes = kHSmemBytes;
	cfg.stream = stream;
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = attr;
	cfg.numAttrs = 1;
	return cudaLaunchKernelEx(&cfg, hier_kernel<K, WRITE>, p);
}

cudaError_t launch_hier(int K, bool write, const HierParams& p, cudaStream_t stream)
{
	if (p.ncb == 0 || p.seg_end <= p.seg_begin)
		return cudaSuccess;
	switch (K * 2 + (write ? 1 : 0))
	{
	case 2: return launch_hier_t<1, false>(p, stream);
	case 3: return launch_hier_t<1, true>(p, stream);
	case 4: return launch_hier_t<2, false>(p, stream);
	case 5: return launch_hier_t<2, true>(p, stream);
	case 8: return launch_hier_t<4, false>(p, stream);
	case 9: return launch_hier_t<4, true>(p, stream);
	default: return cudaErrorInvalidValue;
	}
}

// ======================================================================================================
// copy pass (instead of the write pass when the count pass kept its hits in seg_lists): CTA = 32 clusters (lane) x 8
// segment groups (warp), the shape of hier_finalize_kernel.  Sweep 1 sums this warp's segments, the warps exchange their
// sums, sweep 2 walks the segments again with the running prefix and copies each (segment, cluster) list -- its first
// min(count, cap - prefix) entries -- to indices[offset + prefix ...].  A count may exceed what the list holds (hits are
// counted past the lane's limit inside a batch) but never the part that is needed: limit = cap - (a lower bound of the
// earlier hits) >= cap - prefix.  Warp 0 also writes the light cells (LightCulling.comp:86-87).
// ======================================================================================================
__global__ void __launch_bounds__(kFinWarps * 32) hier_copy_kernel(const CopyParams p)
{
	__shared__ uint32_t s_part[kFinWarps][32];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t c = blockIdx.x * 32 + lane;
	const bool valid = c < p.cluster_count;
	grid_dep_launch();
	grid_dep_wait();
	const uint32_t base = p.index_base ? *p.index_base : 0u;
	const uint32_t group = (p.num_segs + kFinWarps - 1) / kFinWarps;
	const uint32_t s0 = min(warp * group, p.num_segs), s1 = min(s0 + group, p.num_segs);
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
	uint32_t run = 0;
#pragma unroll
	for (uint32_t w = 0; w < kFinWarps; ++w)
		run += w < warp ? s_part[w][lane] : 0u;
	if (blockIdx.x == 0 && threadIdx.x == 0 && p.frame_total)
		*p.frame_total = base + *p.shard_total; // globalIndex after the frame (LightCulling.comp:21,79)
	if (!valid)
		return;
	const uint32_t off = p.offsets[c] + base;
	if (warp == 0)
		*reinterpret_cast<uint2*>(p.cells + p.cluster_begin + c) = make_uint2(off, p.counts[c]);
	if (part == 0 || run >= p.cap)
		return;
	uint32_t* const dst0 = p.indices + off;
	for (uint32_t s = s0; s < s1 && run < p.cap; ++s)
	{
		const uint32_t a = col[static_cast<size_t>(s) * p.cluster_count];
		if (a == 0)
			continue;
		const uint32_t n = min(a, p.cap - run);
		const uint32_t* src = p.seg_lists + (static_cast<size_t>(s) * p.cluster_count + c) * p.cap;
		uint32_t* dst = dst0 + run;
		uint32_t i = 0;
		for (; i + 4 <= n; i += 4)
		{
			const uint32_t v0 = __ldcs(src + i), v1 = __ldcs(src + i + 1), v2 = __ldcs(src + i + 2), v3 = __ldcs(src + i + 3);
			dst[i] = v0, dst[i + 1] = v1, dst[i + 2] = v2, dst[i + 3] = v3;
		}
		for (; i < n; ++i)
			dst[i] = __ldcs(src + i);
		run += a;
	}
}

cudaError_t launch_hier_copy(const CopyParams& p, cudaStream_t stream)
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
	return cudaLaunchKernelEx(&cfg, hier_copy_kernel, p);
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
	HVCF_PRELOAD((hier_kernel<1, false>));
	HVCF_PRELOAD((hier_kernel<1, true>));
	HVCF_PRELOAD((hier_kernel<2, false>));
	HVCF_PRELOAD((hier_kernel<2, true>));
	HVCF_PRELOAD((hier_kernel<4, false>));
	HVCF_PRELOAD((hier_kernel<4, true>));
	HVCF_PRELOAD(hier_finalize_kernel);
	HVCF_PRELOAD(hier_copy_kernel);
	return cudaSuccess;
}

} // namespace hvcf
