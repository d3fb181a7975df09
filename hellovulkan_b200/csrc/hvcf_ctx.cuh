// hvcf_ctx.cuh -- the context behind the opaque hvcf_ctx* of include/hvcf.h and the error helpers, shared by the translation
// units that implement the C ABI (hvcf_api.cu: context, lights, UBOs, the frame; hvcf_api_peer.cu: the peer exchange; hvcf_api_io.cu:
// downloads, counters, raw buffers; hvcf_api_next.cu: consumer + frustum culling; hvcf_shard_api.cu: the routed frame; hvcf_noalloc.cu).
#pragma once

#include "../../include/hvcf.h"

#include <cstdlib>
#include <functional>

#include <nvtx3/nvToolsExt.h>

#include "hvcf_kernels.cuh"

namespace hvcf
{
// sets the calling thread's hvcf_last_error_string() and returns `code`
int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));

inline int env_int(const char* name, int dflt)
{
	const char* v = getenv(name);
	return (v && *v) ? atoi(v) : dflt;
}
// profiler zone of the calling thread for the duration of a scope: the reference wraps the pass in TracyVkZoneC(..., "Light_Culling", ...)
// (Source/Pipelines/PipelineLightCulling.cpp:53); here the same names as NVTX ranges (nsys / ncu --nvtx pick them up; a no-op otherwise)
struct NvtxZone
{
	explicit NvtxZone(const char* name) { nvtxRangePushA(name); }
	~NvtxZone() { nvtxRangePop(); }
	NvtxZone(const NvtxZone&) = delete;
	NvtxZone& operator=(const NvtxZone&) = delete;
};
} // namespace hvcf

struct hvcf_ctx;
void peer_unmap(hvcf_ctx* c); // hvcf_api_peer.cu: closes the IPC mappings of the peers (also on hvcf_destroy)
// pieces of the frame shared by hvcf_api.cu (hvcf_cull*) and hvcf_shard_api.cu (hvcf_cull_routed)
extern "C" {
int hvcf_cull_open(hvcf_ctx* ctx, uint32_t frame, void* stream, const char* who, bool routed, bool allow_rounds);
int hvcf_hier_close(hvcf_ctx* ctx, uint32_t frame, cudaStream_t s);
void hvcf_shard_free(hvcf_ctx* ctx);
cudaError_t hvcf_kmark(hvcf_ctx* ctx, uint32_t frame, int i, cudaStream_t s);
int hvcf_peer_error_after_sync(hvcf_ctx* ctx, const char* what);
// rounds: level-0 binning (bin_round(r) launches round r's, against what hier_alive_kernel left alive) + count pass, round by round
int hvcf_hier_rounds(hvcf_ctx* ctx, uint32_t frame, cudaStream_t s, uint32_t* launches, const std::function<int(uint32_t)>& bin_round);
int hvcf_peer_totals_step(hvcf_ctx* ctx, void* stream, uint32_t what); // 1 = post, 2 = wait + index base, 3 = both
}

#define HVCF_CUDA(expr)                                                                                      \
	do                                                                                                       \
	{                                                                                                        \
		cudaError_t e_ = (expr);                                                                             \
		if (e_ != cudaSuccess)                                                                               \
			return hvcf::fail(e_ == cudaErrorMemoryAllocation ? HVCF_ERR_OUT_OF_MEMORY : HVCF_ERR_CUDA,            \
				"%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_));                  \
	} while (0)

using namespace hvcf; // private header of the C ABI translation units only

constexpr uint32_t kBin2MinGroups = 48;

// lights owned in shards + clusters sharded by z-slab (hvcf_shard.cu, hvcf_shard_api.cu)
struct ShardState
{
	bool on = false;
	uint32_t rank = 0, world = 0, total_lights = 0;
	uint32_t block_tiles = 0;                  // B: tiles per route block
	uint32_t blocks_per_rank = 0, blocks_phys = 0, tiles_phys = 0; // blocks_phys = blocks_per_rank * world, tiles_phys = blocks_phys * B
	uint32_t row_cuts[HVCF_MAX_PEERS + 1]{};   // rank r owns the cluster rows [row_cuts[r], row_cuts[r + 1]), row = y + z * Y
	// inbox: ONE allocation (one IPC handle), per frame parity { rec[tiles_phys][1024] float4, mask[tiles_phys][32] u32, cnt[blocks_phys] u32 }
	unsigned char* d_inbox = nullptr;
	size_t parity_bytes = 0, off_mask = 0, off_cnt = 0;
	unsigned char* inbox_peer[HVCF_MAX_PEERS]{}; // every rank's inbox (own pointer at [rank]); valid while peers are attached
	uint32_t* d_ranges = nullptr;              // [world][2] cluster range of every rank
	float* d_slab_box = nullptr;               // [world][8]
	uint32_t* d_tile_cnt = nullptr;            // [tiles_phys] valid records of every inbox tile of the open frame
	uint16_t* d_tile_ids = nullptr;            // [tiles_phys][1024] in-block light index of every inbox record of the open frame (rebuilt by the binning kernel)
	uint32_t epoch = 0;                        // routed-frame serial, the same on every rank
	uint32_t* d_cnt_last = nullptr;            // [blocks_phys] the last routed frame's records per route block (the inbox's own counts are cleared after use)
	uint32_t* d_seg = nullptr;                 // [cluster_count * cap] this rank's index segment at LOCAL positions (the write pass runs before the totals are known)
	bool two_phase = true;                     // HVCF_ROUTE_TWO_PHASE=0: every light is routed against the slabs' union boxes, in one launch
	bool seg_local = false;                    // the write pass of the open frame writes d_seg / local offsets (hvcf_cull_routed)
};

struct hvcf_ctx
{
	ShardState shard{};
	bool routed_frame = false;               // the open frame is a routed one (hvcf_cull_routed): tiles = inbox tiles
	hvcf_config cfg{};
	int device = 0;
	uint32_t X = 0, Y = 0, Z = 0;
	uint32_t cluster_total = 0;  // whole grid
	uint32_t cluster_begin = 0;  // shard [begin, end)
	uint32_t cluster_end = 0;
	uint32_t cluster_count = 0;  // shard size
	uint32_t cap = 0;
	uint32_t max_lights = 0;
	uint32_t padded_capacity = 0; // max_lights rounded up to a tile
	uint32_t frame_count = 0;
	uint32_t num_lights = 0;
	const hvcf_light_data* bound_lights = nullptr; // hvcf_bind_lights: external buffer, not owned
	uint32_t max_segs = 1;
	uint32_t count_blocks = 0;
	// hierarchical mode
	bool hier = false;
	HierPlan hp{};
	std::vector<uint32_t> lane_map;          // host copy between planning and upload (hvcf_create)
	uint32_t h_max_segs = 0, h_max_chunks = 0;
	uint32_t h_plan_max_segs = 0;            // most whole-tile segments the plan may use
	uint32_t h_parts_cfg = 0, h_parts = 1;   // HVCF_HIER_PARTS (0 = auto); parts per tile of this frame (1 unless every segment is one tile)
	uint32_t h_pipe = 0;                     // this frame walks its slots with the pipelined kernels
	int h_tps_override = 0;
	uint32_t h_split_head = 0;
	uint32_t h_tiles_total = 0, h_tiles_per_seg = 0, h_num_segs = 0; // of the frame between cull_begin and cull_finish
	// rounds (hier_alive_kernel, hvcf_bin.cu): a frame of many tiles is binned and counted in a few rounds of growing length, each
	// knowing exactly which cluster groups the earlier rounds have filled.  h_num_rounds <= 1: one launch over all segments.
	struct Round
	{
		uint32_t seg_begin, seg_end, tile_begin, tile_end, tiles_per_seg, pipe, macro_shift;
	};
	Round h_rounds[kHMaxChunks]{};
	uint32_t h_num_rounds = 0;
	int rounds_mode = -1;                    // HVCF_ROUNDS: -1 = by the number of tiles, 0 = never, n > 0 = at most n rounds whenever the frame has two tiles or more
	uint32_t rounds_min_tiles = 2048, rounds_head = 64, rounds_growth = 8; // HVCF_ROUNDS_MIN_TILES / _HEAD / _GROWTH
	bool small_mode = true;                  // HVCF_SMALL: frames of at most 2,048 lights on an unsharded context run as one kernel (hvcf_small.cu)
	unsigned long long* d_small_state = nullptr; // its look-back words, per-CTA overflow counts, {serial, ticket}
	uint32_t* d_small_over = nullptr;
	uint32_t* d_small_ctl = nullptr;
	bool write_fork = true;                  // HVCF_WRITE_FORK: the write passes of the rounds after the first run on side streams
	cudaStream_t write_streams[3]{};
	cudaEvent_t write_ev[4]{};               // [0] fork (after finalize), [1..3] join
	uint32_t* d_group_alive = nullptr;       // [kHMaxChunks][ncb]: per round, which groups still had a cluster below its cap when it began
	float* d_alive_union = nullptr;          // [ncb][8] union box of each group's live clusters (rounds after the first)
	const uint32_t* bin_alive = nullptr;     // what the next binning launch takes as BinParams::group_alive
	uint32_t rounds_early_segs = 16;         // HVCF_ROUNDS_EARLY_SEGS: light segments of each of the first two rounds (at most one per tile)
	uint32_t rounds_late_segs = 32;          // HVCF_ROUNDS_LATE_SEGS: light segments of every round after the first two
	uint32_t bin_macro = 1;                  // ... and as BinParams::macro
	int rounds_macro = -1;                   // HVCF_ROUNDS_MACRO: -1 = by the round's size, else tiles per macro tile of the rounds after the first two (power of two <= 64)
	cudaEvent_t kev_phase2[8][2]{};          // HVCF_KERNEL_TIMES, two-phase routed frames: around live boxes + second route + waits (counted as "transform")
	cudaEvent_t kev_round[8][kHMaxChunks][3]{}; // HVCF_KERNEL_TIMES with rounds: [frame][round][before, after alive + bin, after count]
	uint32_t* d_hseg_counts = nullptr;
	uint32_t* d_hseg_prefix = nullptr;
	uint32_t* d_hunit_any = nullptr;         // [h_max_segs][ncb * 8]
	uint32_t* d_hscratch = nullptr;          // cum[cluster_count][kHMaxChunks]
	uint32_t* d_hoffsets = nullptr;
	uint32_t* d_hlane_cluster = nullptr;
	float* d_cb_union = nullptr;             // [ncb][8]
	float* d_sb_union = nullptr;             // [bin_ns][8] super-group boxes of the two-level bin kernel
	uint32_t bin_sg = 0, bin_ns = 0;         // groups per super-group, super-groups (0 = one-level bin kernel only)
	int bin2_mode = -1;                      // HVCF_BIN2: -1 = per frame by size, 0 = never, 1 = always
	bool download_kernel = true;             // HVCF_DOWNLOAD_KERNEL: hvcf_download_grid into pinned memory with one kernel
	uint32_t download_blocks = 64;           // HVCF_DOWNLOAD_BLOCKS
	uint32_t download_max_words = 1u << 20;  // HVCF_DOWNLOAD_KERNEL_MAX_MB (4): indices beyond this go through the copy engine
	int pipe_mode = -1;                      // HVCF_BATCH_PIPE: -1 = by tiles per segment, 0 = never, 1 = always
	uint32_t pipe_min_tps = 16;              // HVCF_PIPE_MIN_TPS
	uint32_t xbin_max_lights = 0;            // HVCF_XBIN_MAX: frames with at most this many lights fuse transform + binning (0 = never)
	uint32_t* d_bin_count = nullptr;         // [ncb][tiles_cap]
	float4* d_bin_rec = nullptr;             // [ncb][tiles_cap][kBinCap]
	uint16_t* d_bin_idx = nullptr;
	uint32_t tiles_cap = 0;
	unsigned long long* d_scan_state = nullptr;
	unsigned long long* d_trace = nullptr;   // HVCF_HIER_TRACE=1: per-CTA timestamps (debug)

	// reference-layout buffers
	hvcf_light_data* d_lights = nullptr;   // ResourcesLight::storageBuffer_
	hvcf_aabb* d_aabb = nullptr;           // ResourcesClusterForward::aabbBuffer_
	hvcf_light_cell* d_cells = nullptr;    // lightCellsBuffer_
	uint32_t* d_indices = nullptr;         // lightIndicesBuffer_
	uint32_t* d_global_index = nullptr;    // globalIndexCountBuffers_[frame]
	hvcf_cluster_forward_ubo* d_ubo = nullptr; // cfUBOBuffers_[frame]
	// library-internal
	float4* d_lights_view = nullptr;
	float* d_zplanes = nullptr;            // [frame][Z+1]
	uint32_t* d_seg_counts = nullptr;
	uint32_t* d_seg_lists = nullptr;
	uint32_t* d_counts = nullptr;
	uint32_t* d_block_sums = nullptr;
	uint32_t* d_block_offsets = nullptr;
	uint32_t* d_shard_total = nullptr;
	DeviceStats* d_stats = nullptr;        // [frame]
	// pinned staging
	hvcf_cluster_forward_ubo* h_ubo = nullptr; // [frame]
	float* h_zplanes = nullptr;                // [frame][Z+1]
	uint32_t* h_total = nullptr;               // [8] pinned: hvcf_download_grid
	hvcf_light_data* h_one_light = nullptr;    // hvcf_update_light staging ring
	uint32_t one_light_cursor = 0;
	cudaEvent_t one_light_events[16]{};

	// device-side light animation (hvcf_set_light_animation) and batched position updates (hvcf_update_lights)
	float* d_anim_table = nullptr;
	float* d_base_y = nullptr;                 // [max_lights]
	uint32_t base_y_count = 0;                 // lights whose rest height is in d_base_y
	uint32_t anim_len = 0;                     // 0 = off
	float anim_amplitude = 0.0f;
	uint32_t anim_step[8]{};                   // per frame slot
	uint32_t* d_upd_idx = nullptr;             // device staging of hvcf_update_lights
	float* d_upd_xyz = nullptr;
	uint32_t* h_upd_idx[2]{};                  // pinned, double-buffered
	float* h_upd_xyz[2]{};
	uint32_t upd_capacity = 0;
	uint32_t upd_cursor = 0;
	cudaEvent_t upd_events[2]{};

	// consumer (hvcf_shade): z-slice thresholds in linear depth, per frame slot, built lazily
	float* d_slice_thr = nullptr;              // [frame][Z+1]
	float* h_slice_thr = nullptr;              // pinned
	float slice_key[8][2]{};                   // sliceScaling, sliceBias of the table in the slot
	bool slice_valid[8]{};

	// exchange step of the sharded grid over peer memory (hvcf_peer.cu)
	PeerParams peer{};                         // world == 0: not attached
	PeerMailbox* d_mail = nullptr;
	uint32_t* d_index_base = nullptr;
	uint32_t* d_peer_totals = nullptr;         // [kMaxPeers]
	bool peer_ipc = false;                     // peers were mapped with cudaIpcOpenMemHandle
	uint32_t peer_epoch = 0;
	uint32_t peer_frame = 0;                   // frame slot of the last hvcf_cull_begin
	// hvcf_cull_begin_range progress of the open frame
	bool range_open = false, range_cleared = false;
	uint32_t range_frame = 0, range_transformed = 0, range_binned = 0; // lights (padded at the end) done so far
	uint32_t open_launches = 0;

	cudaEvent_t ubo_events[8]{};               // per frame slot: the staged UBO has been consumed
	bool* ubo_set = nullptr;
	bool aabb_dirty = true;
	bool have_aabb_key = false;
	float aabb_key[16 + 2]{}; // inverse projection + near/far of the last generated AABBs
	uint64_t device_bytes = 0;           // every cudaMalloc of the context (hvcf_stats)
	uint32_t last_plan_flags = 0;        // HVCF_PLAN_* of the last frame
	uint32_t last_num_segs = 0;
	uint32_t last_launches[8]{}; // per frame
	bool timing = false;
	bool ktimes = false;                 // HVCF_KERNEL_TIMES=1 (debug): an event after every kernel of the hierarchical frame
	bool kev_exchange[8]{};              // kev[frame][6] (after the totals exchange of a sharded frame) was recorded in this frame
	cudaEvent_t kev[8][8]{};             // [frame][begin, transform, bin, count, finalize, write]
	cudaEvent_t ev[8][4]{}; // per frame: call begin, kernel begin, kernel end, call end
	bool ev_valid[8]{};
};
