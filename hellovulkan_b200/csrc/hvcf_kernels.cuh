// hvcf_kernels.cuh -- launch interface between the C ABI (hvcf_api.cu) and the kernels.
#pragma once

#include "hvcf_device.cuh"

#include <vector>

namespace hvcf
{

struct CullParams
{
	const float4* lights_view;   // culling layout, tiles_total * kTileBytes
	const hvcf_aabb* aabbs;      // whole grid
	uint32_t cluster_begin;      // shard, global cluster ids
	uint32_t cluster_count;      // clusters in the shard
	uint32_t tiles_total;
	uint32_t tiles_per_seg;
	uint32_t cap;                // max_lights_per_cluster
	uint32_t* seg_counts;        // [num_segs][cluster_count]
	uint32_t* seg_lists;         // [num_segs][cluster_count][cap]
	DeviceStats* stats;
};

struct ScatterParams
{
	const uint32_t* counts;        // [cluster_count]
	const uint32_t* block_offsets; // [ceil(cluster_count/256)]
	const uint32_t* index_base;    // device scalar or NULL
	const uint32_t* seg_counts;
	const uint32_t* seg_lists;
	uint32_t num_segs;
	uint32_t cluster_begin;
	uint32_t cluster_count;
	uint32_t cap;
	hvcf_light_cell* cells;        // whole grid
	uint32_t* indices;             // whole grid capacity
	uint32_t* frame_total;         // globalIndexCountBuffers_[frame] or NULL
};

// Device-side light animation of the fly-through (BASELINE.json configs[4]):
//   position.y = base_y[i] + amplitude * table[(step + i) & mask]
// evaluated with one rounded multiply and one rounded add, so that a host evaluation of the same expression is
// bit-identical; the transform kernel applies it on the fly and writes the moved position back into the light
// buffer (what ResourcesLight::UpdateLightPosition does per light, ResourcesLight.cpp:21-29).
struct LightAnim
{
	const float* base_y;          // [num_lights] position.y at hvcf_set_light_animation time
	const float* table;           // [mask + 1]
	uint32_t mask;
	uint32_t step;
	float amplitude;
	hvcf_light_data* lights_rw;   // same buffer as `lights`
};

cudaError_t launch_aabb_generate(const hvcf_cluster_forward_ubo* ubo, const float* zplanes, hvcf_aabb* out,
	uint32_t cluster_count, cudaStream_t stream);
// also clears `zero_words` words at `zero` and the work counters `stats` (either may be null / 0)
cudaError_t launch_light_transform(const hvcf_cluster_forward_ubo* ubo, const hvcf_light_data* lights,
	uint32_t num_lights, uint32_t padded_lights, float4* out, uint32_t* zero, uint32_t zero_words, DeviceStats* stats,
	cudaStream_t stream, const LightAnim* anim = nullptr);
cudaError_t launch_capture_base_y(const hvcf_light_data* lights, uint32_t num_lights, float* base_y, cudaStream_t stream);
// lights[indices[k]].position = (xyz[3k], xyz[3k+1], xyz[3k+2], 1): a batch of ResourcesLight::UpdateLightPosition
cudaError_t launch_scatter_positions(hvcf_light_data* lights, uint32_t num_lights, const uint32_t* indices, const float* xyz,
	uint32_t count, float* base_y, cudaStream_t stream);
// variant: 0..2 = packed FP32x2 with 1/2/4 clusters per warp, 3..5 = scalar FP32 with 1/2/4
constexpr uint32_t kCullClustersPerCta = 16; // cull_kernel<2, true, 8>: 8 warps x 2 clusters
cudaError_t launch_cull(const CullParams& p, uint32_t num_segs, cudaStream_t stream);
cudaError_t launch_count(const uint32_t* seg_counts, uint32_t num_segs, uint32_t cluster_count, uint32_t cap,
	uint32_t* counts, uint32_t* block_sums, DeviceStats* stats, cudaStream_t stream);
cudaError_t launch_scan(const uint32_t* block_sums, uint32_t num_blocks, uint32_t* block_offsets,
	uint32_t* shard_total, DeviceStats* stats, cudaStream_t stream);
cudaError_t launch_scatter(const ScatterParams& p, cudaStream_t stream);
// 16 lane-ops per thread per iteration (8 FMUL + 8 FADD)
cudaError_t launch_fp32_issue_rate(float* sink, uint32_t blocks, uint32_t iters, cudaStream_t stream);

// ---- the whole frame as one kernel, for frames of at most small_frame_max_lights() lights (hvcf_small.cu) ----
struct SmallParams
{
	const hvcf_cluster_forward_ubo* ubo;
	const hvcf_light_data* lights;
	uint32_t num_lights;
	const hvcf_aabb* aabbs;         // the context's clusters (whole grid: unsharded contexts only)
	uint32_t cluster_count;
	uint32_t cap;
	hvcf_light_cell* cells;
	uint32_t* indices;
	uint32_t* frame_total;          // globalIndexCountBuffers_[frame]
	uint32_t* shard_total;
	DeviceStats* stats;
	unsigned long long* scan_state; // [ceil(cluster_count / 32)] look-back words, tagged with *serial
	uint32_t* over_part;            // [ceil(cluster_count / 32)] clusters at the cap, per CTA
	uint32_t* serial;               // device-resident frame serial of this kernel (never 0), bumped by the last CTA
	uint32_t* done;                 // ticket counter
	unsigned long long* trace = nullptr; // debug: phase timestamps of the first CTAs
};
uint32_t small_frame_max_lights();
cudaError_t launch_small_frame(const SmallParams& p, cudaStream_t stream);
cudaError_t preload_small_kernels();

// ---- HVCF_MODE_HIERARCHICAL (hvcf_hier.cu) -------------------------------------------------------------
constexpr uint32_t kHMaxChunks = 8; // early-exit chunks per cluster: 256 segments at most (hier_num_chunks(256) == 7)
struct HierPlan
{
	uint32_t bx, by, bz;      // cluster block shape, bx * by * bz == 32 (one lane per cluster)
	uint32_t nbx, nby, nbz;   // blocks per axis of the shard
	uint32_t nblocks;
	uint32_t K;               // blocks per warp (1, 2 or 4)
	uint32_t ncb;             // CTAs per light segment
	uint32_t brick = 1;       // wide grids: blocks ordered brick by brick inside a slice (hier_build_lane_map); 0 = plain x-then-y order
};

constexpr uint32_t kHMaxParts = 8;   // most sub-tile light segments per tile
struct HierParams
{
	uint32_t parts = 1;            // > 1 (only with one tile per segment): a tile's survivor slot is cut into `parts` batch ranges, one light segment each
	uint32_t pipe = 0;             // 1: software-pipelined slot walk (hier2_kernel<..., PIPE>)
	const float4* lights_view;
	const hvcf_aabb* aabbs;        // whole grid
	const uint32_t* lane_cluster;  // [ncb * 8 * K][32] shard-local cluster index of each lane, 0xffffffff = none
	uint32_t cluster_begin;        // shard, global cluster ids
	uint32_t cluster_count;
	uint32_t ncb;
	uint32_t tiles_total, tiles_per_seg, num_segs; // tiles_total: one past the last tile of this launch's segments
	uint32_t seg_begin, seg_end;   // segments of this launch
	uint32_t tile_begin = 0;       // first tile of segment seg_begin: segment s covers tiles tile_begin + (s - seg_begin) * tiles_per_seg ...
	uint32_t macro_shift = 0;      // log2 of the tiles per macro tile of this launch's slots (pipelined walk only)
	const uint32_t* round_alive = nullptr; // [ncb] rounds after the first: 0 = the group was full before this launch's round; its warp programs leave at once and write nothing
	uint32_t round_chunk = 0xffffffffu; // frames cut into rounds: the early-exit chunk of this launch = its round; all ones = hier_chunk_of(segment)
	uint32_t cap;
	uint32_t* seg_counts;          // [num_segs][cluster_count]   pass 1 out, pass 2 in
	uint32_t* seg_prefix;          // [num_segs][cluster_count]   finalize out, pass 2 in
	uint32_t* cum;                 // [num_chunks][cluster_count] zeroed per frame
	const uint32_t* offsets;       // [cluster_count] shard-local exclusive offsets
	const uint32_t* counts;        // [cluster_count] capped counts
	const uint32_t* index_base;    // device scalar or NULL
	const uint32_t* shard_total;
	hvcf_light_cell* cells;        // whole grid
	uint32_t* indices;
	uint32_t* frame_total;         // globalIndexCountBuffers_[frame] or NULL
	DeviceStats* stats;
	unsigned long long* trace;     // debug: 4 words per CTA of pass 1, then of pass 2 (or NULL)
	// second generation (hvcf_bin.cu): survivors of level 0, one slot per (group, tile)
	const uint32_t* bin_count;     // [ncb][tiles_cap] lights that passed the group's union box (may exceed kBinCap)
	const float4* bin_rec;         // [ncb][tiles_cap][kBinCap] {x, y, z, r*r} in light order
	const uint16_t* bin_idx;       // [ncb][tiles_cap][kBinCap] light index inside the tile
	uint32_t tiles_cap;
	uint32_t* unit_any;            // [num_segs][ncb * 8] count pass: did this warp program count any hit?  The write pass leaves at once if not
	// routed frames (hvcf_shard.cu): the tiles are this rank's INBOX -- plain {x, y, z, r*r} records that the light owners
	// compacted block by block -- instead of the packed view cache of all lights.  A light's index is then
	//   (tile & tile_mask) * 1024 + tile_ids[tile][position]      (first light of the owner's block + index inside the block).
	uint32_t tile_mask = 0xffffffffu; // ~(tiles per route block - 1); all ones = every tile starts at tile * 1024
	const uint16_t* tile_ids = nullptr; // [tiles][1024] or NULL: lights_view is the packed cache and position = index inside the tile
	const uint32_t* tile_cnt = nullptr; // [tiles] valid records of an inbox tile (written by the routed binning kernel)
};

#ifndef HVCF_BIN_CAP
#define HVCF_BIN_CAP 512
#endif
constexpr uint32_t kBinCap = HVCF_BIN_CAP;  // records per (group, tile) slot; more survivors -> consumers read the raw tile

struct BinParams
{
	const float4* lights_view;
	const float* cb_union;         // [ncb][8]
	uint32_t ncb;
	uint32_t tiles_cap;
	uint32_t* bin_count;
	float4* bin_rec;
	uint16_t* bin_idx;
	DeviceStats* stats;
	uint32_t tile_begin;           // first tile of this launch (grid.y counts from it)
	uint32_t tile_end = 0;         // set by the launcher: one past the last tile of this launch
	uint32_t list_supers = 0;      // set by the launcher: super-group survivor lists held in shared memory at a time
	uint32_t macro = 1;            // two-level kernel: tiles per macro tile (power of two <= 64; hvcf_bin.cu), 1 = every tile has its own slots
	const uint32_t* group_alive = nullptr; // [ncb] or NULL: 0 = every cluster of the group already holds its cap (rounds, hier_alive_kernel): its slots stay empty
	// two-level variant (hier_bin2_kernel): ns > 0 selects it
	const float* sb_union;         // [ns][8] union boxes of the super-groups
	uint32_t sg;                   // groups per super-group
	uint32_t ns;                   // super-groups, at most 32
	// routed variant (hier_bin2_kernel<true>): the tile is an inbox tile of plain records, partly filled
	const uint32_t* in_mask = nullptr; // [tiles][32] survivor bitmasks of the SOURCE tiles of each route block (the j-th record of the block's region is its j-th set bit)
	const uint32_t* in_cnt = nullptr; // [tiles / block_tiles] records of each route block
	uint32_t block_tiles = 1;         // tiles per route block (power of two)
	uint32_t* tile_cnt = nullptr;     // [tiles] out: valid records of the tile
	uint16_t* tile_ids = nullptr;     // [tiles][1024] out: index of every record's light inside its route block (for the raw-tile path of the passes)
};

struct FinalizeParams
{
	uint32_t* seg_counts;
	uint32_t* seg_prefix;
	uint32_t num_segs;
	uint32_t cluster_count;
	uint32_t cap;
	unsigned long long* scan_state; // [ceil(cluster_count / 32)] of this frame slot; tagged with stats->epoch
	uint32_t* counts;
	uint32_t* offsets;
	uint32_t* shard_total;
	DeviceStats* stats;
	// rounds: the warp programs of a group that was full before a round write nothing, so a cluster's counts are read only up to
	// the round before which it held its cap (cum[cluster][round] = its hits of that round, exact once the count passes are done)
	const uint32_t* cum = nullptr;            // [cluster_count][kHMaxChunks]
	uint32_t num_rounds = 0;                  // < 2: every (segment, cluster) count is valid
	uint32_t round_seg_begin[kHMaxChunks + 1]{};
};

// concatenates the (segment, cluster) hit lists of the count pass into the light-index buffer and writes the light cells

void hier_plan_grid(uint32_t X, uint32_t Y, uint32_t Zs, int k_override, const uint32_t shape_override[3], uint32_t max_lights, HierPlan* out);
// cluster_lo / cluster_hi: the shard's clusters inside the X x Y x Zs box (box-local indices); clusters outside are masked, blocks
// without any cluster of the shard are dropped and h->nblocks / h->ncb shrink accordingly
void hier_build_lane_map(HierPlan* h, uint32_t X, uint32_t Y, uint32_t Zs, uint32_t cluster_lo, uint32_t cluster_hi, std::vector<uint32_t>* out);
void hier_plan_lights(const HierPlan& h, uint32_t tiles_total, uint32_t max_segs, int tps_override, uint32_t* tiles_per_seg,
	uint32_t* num_segs);
uint32_t hier_num_chunks(uint32_t num_segs);
cudaError_t launch_hier_finalize(const FinalizeParams& p, cudaStream_t stream);
cudaError_t launch_hier_unions(const uint32_t* lane_cluster, const hvcf_aabb* aabbs, uint32_t cluster_begin, uint32_t K, uint32_t ncb,
	float* cb_union, cudaStream_t stream);
cudaError_t launch_hier_bin(const BinParams& p, uint32_t tiles, cudaStream_t stream);
// rounds: which cluster groups still have a cluster below its cap after the rounds counted so far (cum is exact between launches)
cudaError_t launch_hier_alive(const uint32_t* lane_cluster, const hvcf_aabb* aabbs, uint32_t cluster_begin, const uint32_t* cum, uint32_t K, uint32_t ncb,
	uint32_t cap, uint32_t* group_alive, float* alive_union, cudaStream_t stream);
// transform + one-level binning fused (hier_xbin_kernel): the first kernel of a frame with few lights
struct XBinParams
{
	BinParams bin;
	const hvcf_cluster_forward_ubo* ubo;
	const hvcf_light_data* lights;
	uint32_t num_lights;
	uint32_t tiles;                // set by the launcher
	uint32_t* zero;                // early-exit counters to clear
	uint32_t zero_words;
};
cudaError_t launch_hier_xbin(const XBinParams& q, uint32_t tiles, cudaStream_t stream);
cudaError_t launch_hier_super_unions(const float* cb_union, uint32_t ncb, uint32_t sg, uint32_t ns, float* sb_union, cudaStream_t stream);
size_t hier_bin2_smem_bytes(uint32_t ns, uint32_t ncb, bool routed, uint32_t macro, uint32_t list_supers);
cudaError_t launch_hier2(int K, bool write, const HierParams& p, cudaStream_t stream);

// ---- exchange step of the z-slab sharded grid over peer memory (hvcf_peer.cu) ----------------------------
constexpr uint32_t kMaxPeers = 16;
constexpr uint32_t kSlabBoxes = 2;  // live boxes per slab in the second phase of a routed frame: view-space x < 0 and x >= 0
struct PeerMailbox
{
	unsigned long long totals[2][kMaxPeers]; // [frame parity][rank] = epoch << 32 | index total of that rank's slab
	unsigned long long arrived[kMaxPeers];   // [rank] = epoch << 32 once that rank's cells + indices are in OUR buffers
	unsigned int done;                       // CTA counter of our own push kernel
	unsigned int error;                      // 1 = a wait timed out
	unsigned long long routed[2][kMaxPeers]; // [frame parity][rank] = epoch << 32 once that rank's share of the lights is in OUR inbox
	unsigned int route_done;                 // (unused)
	unsigned int pad_;
	// two-phase routed frames (hvcf_shard.cu): the frame's first lights are routed and counted, every rank then publishes the
	// boxes of its clusters that are still below their cap, and the rest of the lights is routed against THOSE
	unsigned long long routed2[2][kMaxPeers];    // [frame parity][rank] = epoch << 32 once that rank's second share is in OUR inbox
	unsigned long long alive_flag[2][kMaxPeers]; // [frame parity][rank] = epoch << 32 once that rank's live boxes are in alive_box
	float alive_box[2][kMaxPeers][kSlabBoxes][8]; // [frame parity][rank][box] { lo.xyz, 0, hi.xyz, 0 }; (+inf, -inf) = nothing alive there
};
struct PeerParams
{
	hvcf_light_cell* cells[kMaxPeers];       // every rank's lightCells buffer (own pointer at [rank])
	uint32_t* indices[kMaxPeers];            // every rank's lightIndices buffer
	PeerMailbox* mail[kMaxPeers];
	uint32_t world, rank;
	uint32_t epoch;                          // sharded-frame serial, > 0, the same on every rank
};
// ---- lights owned in shards: transform + slab test + route into the owning ranks' inboxes (hvcf_shard.cu) ----
constexpr uint32_t kMaxRouteBlockTiles = 16;
struct RouteParams
{
	const hvcf_cluster_forward_ubo* ubo;
	const hvcf_light_data* lights;    // the rank's light buffer, indexed by GLOBAL light index
	uint32_t light_end;               // owned lights end here (<= the total); they begin at block0 * block_tiles * 1024
	uint32_t block_tiles;             // B: tiles per route block, power of two <= kMaxRouteBlockTiles
	uint32_t block0, blocks;          // first route block of this rank (global numbering), own blocks
	uint32_t world, rank, epoch, parity;
	const float* slab_box;            // [world][boxes][8] boxes of every rank's cluster slab { lo.xyz, 0, hi.xyz, 0 }: a light goes where it hits any
	uint32_t boxes;                   // 1 = the slabs' union boxes; kSlabBoxes = their live boxes (second phase)
	uint32_t phase;                   // 0 = first (or only) route launch of the frame: clears the counters, raises `routed`; 1 = second: adds, raises `routed2`
	float4* rec[kMaxPeers];           // every rank's inbox of this frame parity: [tiles][1024] records (a route block's region is dense from its start),
	uint32_t* mask[kMaxPeers];        //   [tiles][32] survivor bitmask of every source tile: the j-th record of a block's region is its j-th set bit,
	uint32_t* cnt[kMaxPeers];         //   [blocks] records per route block
	PeerMailbox* mail[kMaxPeers];
	uint32_t* zero;                   // early-exit counters of the own frame to clear (the route kernel is the frame's first)
	uint32_t zero_words;
	DeviceStats* stats;
};
cudaError_t launch_shard_route(const RouteParams& p, cudaStream_t stream);
// what: 0 / 2 = raise this rank's flag of the first / second share on every rank, then wait for everybody's; 1 = wait for the live boxes
cudaError_t launch_shard_wait_routed(const PeerParams& p, uint32_t parity, cudaStream_t stream, uint32_t what = 0);
// reduces the live boxes of this rank's groups (hier_alive_kernel) to kSlabBoxes boxes and publishes them on every rank
cudaError_t launch_slab_alive(const PeerParams& p, uint32_t parity, const uint32_t* group_alive, const float* alive_union, uint32_t ncb, cudaStream_t stream);
// union box of each rank's cluster range [begin[r], end[r]) of the whole-grid AABB buffer
cudaError_t launch_slab_unions(const hvcf_aabb* aabbs, const uint32_t* d_ranges, uint32_t world, float* slab_box, cudaStream_t stream);
cudaError_t preload_shard_kernels();

cudaError_t launch_peer_totals(const PeerParams& p, const uint32_t* shard_total, uint32_t* index_base, uint32_t* totals_out,
	cudaStream_t stream, uint32_t what = 3u);
cudaError_t launch_peer_push_relocate(const PeerParams& p, const uint32_t* shard_total, const uint32_t* index_base, const uint32_t* seg,
	uint32_t cluster_begin, uint32_t cluster_count, uint32_t blocks, cudaStream_t stream);
cudaError_t launch_peer_push(const PeerParams& p, const uint32_t* shard_total, const uint32_t* index_base, uint32_t cluster_begin,
	uint32_t cluster_count, uint32_t blocks, cudaStream_t stream);
cudaError_t launch_peer_wait(const PeerParams& p, const uint32_t* totals, uint32_t* frame_total, cudaStream_t stream);
// force-load every kernel a frame can launch (see hvcf_kernels.cu: lazy loading vs kernels that spin on another shard)
cudaError_t launch_grid_download(const void* cells, size_t cells_bytes, const uint32_t* indices, const uint32_t* total_ptr,
	uint32_t index_capacity, uint32_t max_words, void* cells_dst, uint32_t* indices_dst, uint32_t* total_dst, uint32_t blocks, cudaStream_t stream);
cudaError_t preload_basic_kernels();
cudaError_t preload_hier_kernels();
cudaError_t preload_bin_kernels();
cudaError_t preload_peer_kernels();

} // namespace hvcf
