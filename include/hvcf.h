/*
 * hvcf.h -- C ABI of libhvcf.so: the B200-native replacement for HelloVulkan's
 * clustered-forward light-culling pass pair (AABBGenerator.comp -> LightCulling.comp).
 *
 * The reference is a monolithic Vulkan executable with no FFI of its own
 * (HelloVulkan/CMakeLists.txt:7-11); each entry point below therefore cites the C++
 * method or shader it replaces.  Paths are relative to /root/reference/HelloVulkan.
 * INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns HVCF_OK (0) or a negative hvcf_status; nothing throws
 *     across this boundary; hvcf_last_error_string() describes the last failure on the
 *     calling thread.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream);
 *     all work is stream-ordered on it -- stream order replaces the reference's
 *     vkCmdPipelineBarrier2 buffer barriers (Source/Pipelines/PipelineAABBGenerator.cpp:70-82,
 *     Source/Pipelines/PipelineLightCulling.cpp:85-118).
 *   - there is NO CPU fallback: every call fails with HVCF_ERR_CUDA when no sm_100 device
 *     is usable.
 *   - a context is owned by one host thread at a time (the reference records and submits
 *     from a single thread, Source/Apps/AppBase.cpp:94-196).
 */
#ifndef HVCF_H
#define HVCF_H

#include "hvcf_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define HVCF_API __attribute__((visibility("default")))
#else
#define HVCF_API
#endif

typedef enum hvcf_status
{
	HVCF_OK = 0,
	HVCF_ERR_INVALID_ARGUMENT = -1,
	HVCF_ERR_CUDA = -2,
	HVCF_ERR_OUT_OF_MEMORY = -3,
	HVCF_ERR_CAPACITY = -4,   /* more lights than hvcf_config.max_lights */
	HVCF_ERR_NOT_READY = -5,  /* e.g. cull before set_lights / set_ubo */
	HVCF_ERR_UNSUPPORTED = -6
} hvcf_status;

/* Culling strategy.  All of them produce byte-identical light cells / index lists. */
typedef enum hvcf_mode
{
	HVCF_MODE_AUTO = 0,        /* library default (hierarchical); the environment variable HVCF_MODE=brute|hier overrides it */
	HVCF_MODE_BRUTE = 1,       /* every cluster x light pair is tested until the cluster holds max_lights_per_cluster hits */
	HVCF_MODE_HIERARCHICAL = 2 /* exactly conservative union-box pre-filters (CTA, 32-cluster block), exact test on survivors */
} hvcf_mode;

typedef struct hvcf_ctx hvcf_ctx;

/*
 * Replaces the compile-time ClusterForwardConfig / AppConfig constants
 * (Header/Configs.h:8-14, :46-56) with run-time fields; zero means "reference default".
 */
typedef struct hvcf_config
{
	uint32_t struct_size;            /* = sizeof(hvcf_config) */
	int32_t device;                  /* CUDA ordinal; -1 = current device */
	uint32_t slice_count_x;          /* ClusterForwardConfig::SliceCountX (16) */
	uint32_t slice_count_y;          /* ClusterForwardConfig::SliceCountY (9) */
	uint32_t slice_count_z;          /* ClusterForwardConfig::SliceCountZ (24) */
	uint32_t max_lights_per_cluster; /* ClusterForwardConfig::maxLightPerCluster (150) */
	uint32_t max_lights;             /* capacity of the light buffer (lights) */
	uint32_t frame_count;            /* AppConfig::FrameCount (2) */
	uint32_t mode;                   /* hvcf_mode */
	/* z-slab shard owned by this context: slices [z_begin, z_end); 0,0 = whole grid. */
	uint32_t z_begin;
	uint32_t z_end;
	/* or, finer, a range of cluster ROWS: row = y + z * slice_count_y, clusters [row_begin * X, row_end * X) -- still a contiguous
	 * cluster-index range (tileIndex = x + y*X + z*X*Y, LightCulling.comp:61-64), so shard segments concatenate exactly as z-slabs
	 * do.  row_end != 0 overrides z_begin / z_end.  Cuts on multiples of the cluster-block height (4 rows for X >= 8) keep every
	 * 32-cluster block whole. */
	uint32_t row_begin;
	uint32_t row_end;
	uint32_t reserved[3];
} hvcf_config;

/* Device pointers in the reference's exact layouts (ResourcesClusterForward.h:24-30). */
typedef struct hvcf_buffers
{
	hvcf_aabb* aabb;                    /* aabbBuffer_          : cluster_count x 32 B */
	hvcf_light_cell* light_cells;       /* lightCellsBuffer_    : cluster_count x 8 B */
	uint32_t* light_indices;            /* lightIndicesBuffer_  : max_lights_per_cluster x cluster_count x 4 B */
	uint32_t* global_index_count;       /* globalIndexCountBuffers_[frame]: frame_count x 4 B (total indices of that frame) */
	hvcf_light_data* lights;            /* ResourcesLight::storageBuffer_: max_lights x 48 B */
	float* lights_view;                 /* view-space SoA float4 {x,y,z,r*r} of the last culled frame (library-internal cache) */
	uint64_t light_indices_capacity;    /* elements */
	uint32_t cluster_count;
	uint32_t light_count;
	uint32_t cluster_begin;             /* shard, in cluster indices */
	uint32_t cluster_end;
} hvcf_buffers;

/* Work counters of the most recent hvcf_cull on `frame` (read back on request only). */
typedef struct hvcf_stats
{
	uint64_t nominal_tests;   /* clusters x lights of the shard */
	uint64_t executed_tests;  /* exact sphere-vs-AABB lane tests (brute: the cull kernel; hierarchical: the count pass) */
	uint64_t prefilter_tests; /* union-box tests of the hierarchical mode (level 0 = binning, level 1 = count pass), else 0 */
	uint32_t total_indices;   /* sum of counts */
	uint32_t overflow_clusters; /* clusters whose count hit max_lights_per_cluster */
	uint32_t kernel_launches; /* kernels launched by that hvcf_cull call */
	uint32_t reserved;
	uint64_t bin_tests;        /* the level-0 (binning kernel) part of prefilter_tests */
	uint64_t write_pass_tests; /* union + exact tests the write pass repeats; in neither counter above */
	/* what the context holds on the device, and the plan the frame ran with (nothing here depends on the frame's lights) */
	uint64_t device_bytes;     /* every cudaMalloc of the context */
	uint64_t scratch_bytes;    /* the library-internal part: everything but lights, AABBs, cells, indices, counters and UBOs.  Mostly the
	                              level-0 survivor slots: groups x light tiles x 512 records x 18 B, committed whatever the filter lets through */
	uint32_t plan_groups;      /* cluster groups (level 0) of the shard */
	uint32_t plan_blocks_per_program; /* K: 32-cluster blocks per warp program */
	uint32_t plan_segments;    /* light segments of the frame */
	uint32_t plan_rounds;      /* rounds the frame was binned and counted in (1 = a single launch each) */
	uint32_t plan_flags;       /* HVCF_PLAN_* */
	uint32_t reserved2;
} hvcf_stats;
#define HVCF_PLAN_TWO_LEVEL_BIN 1u  /* level 0 through super-groups (hier_bin2_kernel) */
#define HVCF_PLAN_FUSED_BIN 2u      /* transform + level 0 in one kernel (hier_xbin_kernel) */
#define HVCF_PLAN_PIPELINED_WALK 4u /* some segment walks its slots with the cp.async double buffer */
#define HVCF_PLAN_SUBTILE_SEGMENTS 8u /* a tile's slot is cut into several light segments */
#define HVCF_PLAN_MACRO_TILES 16u   /* late rounds keep one slot per group and several tiles (or route block) */
#define HVCF_PLAN_TWO_PHASE_ROUTE 32u /* routed frame: second share routed against the slabs' live boxes */
#define HVCF_PLAN_SINGLE_KERNEL 64u   /* at most 2,048 lights, unsharded: transform + test + scan + write in one kernel (hvcf_small.cu) */

HVCF_API const char* hvcf_last_error_string(void);
HVCF_API const char* hvcf_version(void);

/* ResourcesClusterForward::CreateBuffers (Source/Resources/ResourcesClusterForward.cpp:5-34)
 * + pipeline construction (PipelineAABBGenerator.cpp:8-21, PipelineLightCulling.cpp:9-28). */
HVCF_API int hvcf_create(const hvcf_config* config, hvcf_ctx** out_ctx);
/* ResourcesClusterForward::Destroy (ResourcesClusterForward.cpp:36-46) + pipeline destructors. */
HVCF_API int hvcf_destroy(hvcf_ctx* ctx);

/* ResourcesLight::AddLights (Source/Resources/ResourcesLight.cpp:8-18).  `lights` is n x 48 B
 * (hvcf_light_data); src_is_device != 0 means a device pointer (device-to-device copy). */
HVCF_API int hvcf_set_lights(hvcf_ctx* ctx, const void* lights, uint32_t n, int src_is_device, void* stream);
/* ResourcesLight::UpdateLightPosition (ResourcesLight.cpp:20-29): writes xyz and w = 1;
 * an out-of-range index is ignored and returns HVCF_OK, as the reference does. */
HVCF_API int hvcf_update_light(hvcf_ctx* ctx, uint32_t index, const float xyz[3], void* stream);

/* Pipeline{AABBGenerator,LightCulling}::SetClusterForwardUBO
 * (Header/Pipelines/PipelineAABBGenerator.h:24-28, PipelineLightCulling.cpp:45-49): one call
 * feeds both passes.  Marks the AABBs dirty when the inverse projection, near/far or the
 * grid changed (the reference forgets to on zoom). */
HVCF_API int hvcf_set_ubo(hvcf_ctx* ctx, uint32_t frame, const hvcf_cluster_forward_ubo* ubo, void* stream);
/* PipelineAABBGenerator::OnWindowResized (PipelineAABBGenerator.cpp:31-34). */
HVCF_API int hvcf_mark_aabb_dirty(hvcf_ctx* ctx);
HVCF_API int hvcf_is_aabb_dirty(hvcf_ctx* ctx, int* out_dirty);

/* PipelineAABBGenerator::Execute (PipelineAABBGenerator.cpp:49-83) running
 * AABBGenerator.comp:39-90; unconditional -- the dirty gate lives in the C++ class
 * (FillCommandBuffer, PipelineAABBGenerator.cpp:36-47) and in hvcf_cull. */
HVCF_API int hvcf_generate_aabbs(hvcf_ctx* ctx, uint32_t frame, void* stream);

/* PipelineLightCulling::ResetGlobalIndex + Execute (PipelineLightCulling.cpp:38-43, :58-119)
 * running LightCulling.comp:59-88.  Regenerates the AABBs first if they are dirty. */
HVCF_API int hvcf_cull(hvcf_ctx* ctx, uint32_t frame, void* stream);

/* Sharded variant of hvcf_cull, split around the one exchange step of the path:
 * begin  = transform + test + per-cluster counts; leaves this shard's total in
 *          *hvcf_shard_total_ptr() on the device;
 * finish = offsets (+ *d_index_base, a device uint32 holding the sum of the totals of the
 *          lower-ranked shards; NULL = 0) + cells + indices written at their GLOBAL
 *          positions, so shard segments concatenate without a fix-up pass. */
HVCF_API int hvcf_cull_begin(hvcf_ctx* ctx, uint32_t frame, void* stream);
HVCF_API int hvcf_cull_finish(hvcf_ctx* ctx, uint32_t frame, const uint32_t* d_index_base, void* stream);
/* hvcf_cull_begin in pieces, so that the light exchange of the sharded grid (the ncclBroadcast of the light buffer)
 * overlaps the first kernels of the frame: lights [light_begin, light_end) are transformed to view space and binned
 * against the cluster-group boxes while later ranges are still in flight.  Ranges are whole 1,024-light tiles, issued
 * in ascending order from 0; the call that reaches the light count also runs the count pass and the scan, i.e.
 * completes hvcf_cull_begin.  flags = 0 means TRANSFORM | BIN.  With BIN alone the view-space records of the range
 * (16 B per light, hvcf_lights_view_ptr) must already be in place -- e.g. broadcast from the one rank that ran
 * TRANSFORM on them, 16 B per light on the wire instead of 48.  Hierarchical mode only. */
#define HVCF_RANGE_TRANSFORM 1u
#define HVCF_RANGE_BIN 2u
HVCF_API int hvcf_cull_begin_range(hvcf_ctx* ctx, uint32_t frame, uint32_t light_begin, uint32_t light_end, uint32_t flags, void* stream);
HVCF_API int hvcf_lights_view_ptr(hvcf_ctx* ctx, void** out_device_ptr, size_t* out_bytes);
/* The TRANSFORM half alone, for any whole-tile range in any order and without touching the frame's progress: lights
 * [light_begin, light_end) of the bound light buffer -> view-space records with the UBO of `frame` (LightCulling.comp:51-53
 * hoisted).  For lights OWNED in shards: every rank transforms its own slice, the 16 B records are all-gathered, and
 * hvcf_cull_begin_range(0, n, HVCF_RANGE_BIN) takes it from there.  The range that reaches the light count also writes the
 * padding records of the last tile. */
HVCF_API int hvcf_transform_lights(hvcf_ctx* ctx, uint32_t frame, uint32_t light_begin, uint32_t light_end, void* stream);
HVCF_API int hvcf_shard_total_ptr(hvcf_ctx* ctx, uint32_t** out_device_ptr);

/* hvcf_cull for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  It runs the single-kernel frame (the kernel
 * hvcf_cull picks for at most 2,048 lights: LightCulling.comp:49-88 for the whole grid in one launch) with the kernel's scratch --
 * scan state words, overflow counts, ticket counter, work counters -- in a file-scope __device__ array of the library instead of
 * a context's cudaMalloc'ed buffers: 49,664 B, held by every GPU that loads the library until the process exits.  There is no
 * context: every pointer is a DEVICE pointer of the current device, inputs and outputs in the reference's layouts
 * (UBO 176 B, lights n x 48 B, AABBs clusters x 32 B as hvcf_generate_aabbs writes them, light cells clusters x 8 B, light indices,
 * *d_total = globalIndex after the frame); UBO, lights and AABBs 16-byte aligned, light cells 8.  Results are byte-identical to
 * hvcf_cull's.  The call allocates nothing: one cudaMemsetAsync of the scratch in use and one kernel launch on `stream` (the first
 * call on a device also asks for the array's device address, cudaGetSymbolAddress -- make that call outside a capture).
 * Limits: cluster_count <= 131,072, num_lights <= 2,048, light_indices_capacity (elements) >= cluster_count x min(num_lights,
 * max_lights_per_cluster); beyond them HVCF_ERR_CAPACITY and nothing is launched.
 * Every call on a device shares the one array: calls that overlap on the GPU (two streams, two threads) RACE -- serialise them. */
HVCF_API int hvcf_cull_small_noalloc(const hvcf_cluster_forward_ubo* d_ubo, const hvcf_light_data* d_lights, uint32_t num_lights,
	const hvcf_aabb* d_aabbs, uint32_t cluster_count, uint32_t max_lights_per_cluster, hvcf_light_cell* d_light_cells,
	uint32_t* d_light_indices, uint64_t light_indices_capacity, uint32_t* d_total, void* stream);

/* The exchange step of the z-slab sharded grid done by the GPUs themselves over NVLink peer memory (new: the
 * reference is single-GPU).  Every rank exports three allocations (light cells, light indices,
 * a small mailbox) and attaches every other rank's; after that one call per frame,
 *   hvcf_cull_sharded = hvcf_cull_begin                       (local: transform, test, counts, scan)
 *                     + hvcf_peer_exchange_totals             (kernel: post total to all peers, wait, index base)
 *                     + hvcf_cull_finish(own index base)      (cells + indices at GLOBAL positions)
 *                     + hvcf_peer_push_grid                   (kernel: own cells + index segment into every peer)
 *                     + hvcf_peer_wait_grid                   (kernel: until every peer's segment has arrived)
 * leaves the FULL light grid (lightCells[], lightIndices[] in the reference layouts) on every rank, with no host
 * round trip and no collective-library call.  All ranks must call it the same number of times.  Waits are bounded:
 * a missing peer raises the error that hvcf_peer_status reports (after synchronising `stream`).  A rank whose wait timed
 * out skips its remote stores for that frame and marks every peer's error word, and every hvcf_download_* of a context
 * with peers attached returns HVCF_ERR_NOT_READY while the error stands (cleared by hvcf_peer_detach / re-attach).
 *   hvcf_peer_export        handle = 4 x cudaIpcMemHandle_t (64 B each), to be sent to the other processes
 *   hvcf_peer_attach        handles[world], own entry ignored; peers are mapped with cudaIpcOpenMemHandle
 *   hvcf_peer_attach_local  the same for contexts living in THIS process (same or peer-accessible devices)
 *   hvcf_peer_detach        unmaps; call on every rank (after a barrier) before hvcf_destroy */
#define HVCF_MAX_PEERS 16
typedef struct hvcf_peer_handle
{
	unsigned char cells[64];
	unsigned char indices[64];
	unsigned char mailbox[64];
	unsigned char inbox[64];   /* routed frames (hvcf_create_sharded); zero otherwise */
} hvcf_peer_handle;
HVCF_API int hvcf_peer_export(hvcf_ctx* ctx, hvcf_peer_handle* out);
HVCF_API int hvcf_peer_attach(hvcf_ctx* ctx, uint32_t rank, uint32_t world, const hvcf_peer_handle* handles);
HVCF_API int hvcf_peer_attach_local(hvcf_ctx* ctx, uint32_t rank, uint32_t world, hvcf_ctx* const* contexts);
HVCF_API int hvcf_peer_detach(hvcf_ctx* ctx);
HVCF_API int hvcf_peer_exchange_totals(hvcf_ctx* ctx, void* stream);
HVCF_API int hvcf_peer_push_grid(hvcf_ctx* ctx, void* stream);
HVCF_API int hvcf_peer_wait_grid(hvcf_ctx* ctx, void* stream);
HVCF_API int hvcf_cull_sharded(hvcf_ctx* ctx, uint32_t frame, void* stream);
/* device uint32[world]: every rank's index total of the last exchange (valid after hvcf_peer_exchange_totals) */
HVCF_API int hvcf_peer_totals_ptr(hvcf_ctx* ctx, uint32_t** out_device_ptr);
/* device uint32: this rank's index base of the last exchange (what hvcf_cull_sharded passes to hvcf_cull_finish) */
HVCF_API int hvcf_peer_index_base_ptr(hvcf_ctx* ctx, uint32_t** out_device_ptr);
HVCF_API int hvcf_peer_status(hvcf_ctx* ctx, void* stream, int* out_error);

/* ---------------------------------------------------------------------------------------------------
 * The sharded frame with the LIGHTS sharded too (new: the reference is single-GPU).
 * Rank r owns the z-slab [z_cuts[r], z_cuts[r + 1]) of the cluster grid AND the LightData of a contiguous range
 * of lights (hvcf_shard_light_range; whole route blocks of block_tiles x 1,024 lights).  One call per frame,
 *   hvcf_cull_routed = route    own lights -> view space, tested against every rank's slab box, survivors stored
 *                               straight into the owning rank's inbox over NVLink peer memory (in light order)
 *                    + wait     until every rank's share has arrived
 *                    + level-0 binning of the inbox, count pass, scan          (local)
 *                    + totals exchange, write pass at GLOBAL offsets, grid push, wait   (as hvcf_cull_sharded)
 * leaves the FULL light grid on every rank, byte-identical to the single-GPU frame over all the lights.  No rank ever
 * holds, transforms or bins all the lights; there is no collective-library call and no host round trip in the frame.
 * Setup: hvcf_create_sharded on every rank, hvcf_peer_export / exchange the handles / hvcf_peer_attach (or
 * hvcf_peer_attach_local for contexts of one process), hvcf_set_lights_range for the owned lights, hvcf_set_ubo.
 * world = 1 needs no attach.  Replaces, at scale, the single dispatch of PipelineLightCulling::Execute
 * (Source/Pipelines/PipelineLightCulling.cpp:58-119) over LightCulling.comp:59-88.
 * ------------------------------------------------------------------------------------------------- */
typedef struct hvcf_shard_desc
{
	uint32_t struct_size;                  /* = sizeof(hvcf_shard_desc) */
	uint32_t rank, world;                  /* world <= HVCF_MAX_PEERS */
	uint32_t total_lights;                 /* lights of the whole scene, the same on every rank */
	uint32_t block_tiles;                  /* tiles per route block, a power of two <= 16; 0 = library default */
	uint32_t z_cuts[HVCF_MAX_PEERS + 1];   /* ascending, z_cuts[0] = 0, z_cuts[world] = slice_count_z */
	/* or, finer (used when row_cuts[world] != 0): cuts in cluster rows (row = y + z * slice_count_y), row_cuts[0] = 0,
	 * row_cuts[world] = slice_count_y * slice_count_z */
	uint32_t row_cuts[HVCF_MAX_PEERS + 1];
	uint32_t reserved[3];
} hvcf_shard_desc;
/* hvcf_create with z_begin / z_end = this rank's slab and max_lights raised to the padded light count of the
 * decomposition, plus the inbox, the slab boxes and the ownership plan. */
HVCF_API int hvcf_create_sharded(const hvcf_config* config, const hvcf_shard_desc* desc, hvcf_ctx** out_ctx);
/* the lights this rank owns: [*out_begin, *out_end) in GLOBAL light indices (may be empty) */
HVCF_API int hvcf_shard_light_range(hvcf_ctx* ctx, uint32_t* out_begin, uint32_t* out_end);
/* ResourcesLight::AddLights for a part of the light buffer (VulkanBuffer::UploadOffsetBufferData,
 * Source/Vulkan/VulkanBuffer.cpp:158-167): lights [begin, begin + count) <- `lights` (count x 48 B). */
HVCF_API int hvcf_set_lights_range(hvcf_ctx* ctx, const void* lights, uint32_t begin, uint32_t count, int src_is_device, void* stream);
HVCF_API int hvcf_cull_routed(hvcf_ctx* ctx, uint32_t frame, void* stream);
/* This rank's OWN part of the result -- its slab's light cells [cluster_begin, cluster_end) and its segment
 * lightIndices[*out_index_base, *out_index_base + *out_total) -- into host memory (synchronises `stream`).  Over all ranks
 * the parts add up to the full light grid exactly once: what a host-side consumer of a sharded frame reads back.  Works
 * after hvcf_cull_routed, hvcf_cull_sharded and, with base 0, hvcf_cull. */
HVCF_API int hvcf_download_shard(hvcf_ctx* ctx, void* cells_dst, size_t cells_bytes, void* indices_dst, size_t indices_capacity_bytes,
	uint32_t* out_index_base, uint32_t* out_total, void* stream);
/* records that arrived in this rank's inbox in the last routed frame (synchronises `stream`); the wire carried
 * 16 B per record that came from another rank, plus one bit per (light, destination) */
HVCF_API int hvcf_shard_inbox_records(hvcf_ctx* ctx, void* stream, uint64_t* out_records);

HVCF_API int hvcf_get_device_buffers(hvcf_ctx* ctx, hvcf_buffers* out);

/* VulkanBuffer::DownloadBufferData (Source/Vulkan/VulkanBuffer.cpp:169-178): synchronous
 * (they synchronise `stream` before returning). `bytes` must not exceed the buffer. */
HVCF_API int hvcf_download_aabbs(hvcf_ctx* ctx, void* dst, size_t bytes, void* stream);
HVCF_API int hvcf_download_light_cells(hvcf_ctx* ctx, void* dst, size_t bytes, void* stream);
HVCF_API int hvcf_download_light_indices(hvcf_ctx* ctx, void* dst, size_t bytes, void* stream);
HVCF_API int hvcf_download_total(hvcf_ctx* ctx, uint32_t frame, uint32_t* out_total, void* stream);
/* The whole result of a frame in one call: total + light cells, one synchronisation, then exactly `total` light indices
 * (what a host-side consumer of lightCells / lightIndices reads back; two synchronisations instead of three calls' three).
 * cells_bytes <= 8 x clusters; at most indices_capacity_bytes of indices are copied (HVCF_ERR_CAPACITY if total does not fit).
 * When both destinations are pinned host memory (hvcf_host_alloc, cudaMallocHost), 16-byte aligned, one kernel reads the count on
 * the device and stores count, cells and the first 4 MB of indices across PCIe itself: ONE synchronisation, no host round trip
 * between "how many" and "copy them" (longer lists: the copy engine takes the rest).  HVCF_DOWNLOAD_KERNEL=0 switches it off. */
HVCF_API int hvcf_download_grid(hvcf_ctx* ctx, uint32_t frame, void* cells_dst, size_t cells_bytes, void* indices_dst,
	size_t indices_capacity_bytes, uint32_t* out_total, void* stream);
HVCF_API int hvcf_get_stats(hvcf_ctx* ctx, uint32_t frame, hvcf_stats* out, void* stream);

/* Binds an EXTERNAL device light buffer (n x 48 B) instead of copying into the context's own:
 * the descriptor write of binding 1 in PipelineLightCulling::CreateDescriptor
 * (PipelineLightCulling.cpp:121-141).  The caller keeps ownership; hvcf_cull reads it in place,
 * so ResourcesLight::UpdateLightPosition writes are seen by the next frame without a copy.
 * NULL rebinds the context's own buffer.  The buffer is read-only to the library EXCEPT with device-side light animation on
 * (hvcf_set_light_animation): the moved positions are then written back into it, as UpdateLightPosition writes them into the
 * SSBO, and binding a new buffer (or more lights) re-captures the rest heights on the default stream.
 * A context of hvcf_create_sharded binds exactly the lights it OWNS: device_lights[0] is light `begin` of hvcf_shard_light_range and
 * n = end - begin.  hvcf_cull_routed reads it in place (e.g. two buffers bound in turn, the next frame's lights uploading into one
 * while the other is culled); the other light calls of such a context return HVCF_ERR_UNSUPPORTED while a buffer is bound. */
HVCF_API int hvcf_bind_lights(hvcf_ctx* ctx, const void* device_lights, uint32_t n);

/* Raw device buffers for the C++ mirror of VulkanBuffer (include/hv/Context.h):
 * VulkanBuffer::CreateBuffer / Destroy / UploadOffsetBufferData / DownloadBufferData
 * (Source/Vulkan/VulkanBuffer.cpp:6-50, :148-178).  upload is stream-ordered and stages through
 * pageable memory (synchronous w.r.t. the host buffer); download synchronises `stream`. */
HVCF_API int hvcf_device_alloc(int device, size_t bytes, void** out_device_ptr);
HVCF_API int hvcf_device_free(void* device_ptr);
HVCF_API int hvcf_device_upload(void* device_dst, size_t offset, const void* host_src, size_t bytes, void* stream);
HVCF_API int hvcf_device_download(void* host_dst, const void* device_src, size_t offset, size_t bytes, void* stream);
HVCF_API int hvcf_stream_synchronize(void* stream);

/* Host-side producers of the path's inputs, compiled from include/hv/Camera.h and
 * hellovulkan_b200/host/hvcf_host.cpp (C++, CPU -- as in the reference):
 *   Camera::SetPositionAndTarget + GetClusterForwardUBO (Source/Camera.cpp:81-93,113-139,178-197)
 *   AppPBRClusterForward::InitLights with an explicit seed (Source/Apps/AppPBRClusterForward.cpp:81-118) */
HVCF_API int hvcf_host_camera_ubo(const float position[3], const float target[3], float screen_width, float screen_height,
	uint32_t slice_count_x, uint32_t slice_count_y, uint32_t slice_count_z, hvcf_cluster_forward_ubo* out);
HVCF_API int hvcf_host_init_lights(uint32_t seed, uint32_t count, hvcf_light_data* out);

/* Measurement hooks (TracyVkZone "Light_Culling", PipelineLightCulling.cpp:53, becomes CUDA events):
 * with timing enabled hvcf_cull brackets the culling kernel and the whole call with events on `stream`;
 * hvcf_get_timing synchronises the end event and returns milliseconds. */
/* debug: per-CTA {start ns, end ns, SM id, level-2 lane tests} of the last hierarchical frame, pass 1 then pass 2;
 * needs HVCF_HIER_TRACE=1 in the environment at hvcf_create. */
HVCF_API int hvcf_debug_trace(hvcf_ctx* ctx, unsigned long long* out_words, size_t max_words, size_t* out_ctas_per_pass);
/* debug: milliseconds of {transform, bin, count pass, finalize, write pass} of the last hierarchical hvcf_cull on `frame`
 * (an event after every kernel, which also switches their programmatic overlap off); needs HVCF_KERNEL_TIMES=1 at hvcf_create. */
HVCF_API int hvcf_debug_kernel_times(hvcf_ctx* ctx, uint32_t frame, float out_ms[5]);
/* debug: milliseconds of the route kernel alone in the last hvcf_cull_routed on `frame` (out_ms[0] of hvcf_debug_kernel_times is
 * route + the wait for the other ranks' shares); needs HVCF_KERNEL_TIMES=1 at hvcf_create. */
HVCF_API int hvcf_debug_route_ms(hvcf_ctx* ctx, uint32_t frame, float* out_route_ms);
/* debug: the rounds plan of the last hierarchical frame and the level-0 state it left behind.
 *   plan[0..3] = groups (ncb), blocks per warp program (K), slot rows per group (tiles_cap), rounds (0 or 1 = a single launch);
 *   plan[4 + 8 r ..] = round r: seg_begin, seg_end, tile_begin, tile_end, tiles_per_seg, pipe, log2(tiles per macro tile), 0
 *   what = 0: uint32 alive[rounds][ncb] (row 0 is not used);  1: uint32 bin_count[ncb][tiles_cap];  2: float alive_union[ncb][8]
 * Synchronises the device.  plan holds 4 + 8 * 8 words; dst may be NULL to fetch the plan only. */
HVCF_API int hvcf_debug_rounds(hvcf_ctx* ctx, uint32_t plan[68], uint32_t what, void* dst, size_t bytes);
HVCF_API int hvcf_set_timing(hvcf_ctx* ctx, int enabled);
HVCF_API int hvcf_get_timing(hvcf_ctx* ctx, uint32_t frame, float* out_cull_kernel_ms, float* out_total_ms);
/* Measured FP32 issue rate of the device (independent FADD/FMUL chains, no FMA), in lane-operations per
 * second: the denominator of the FP32-issue roofline; MEASURED_PEAKS.json has none. */
HVCF_API int hvcf_measure_fp32_issue_rate(int device, double* out_lane_ops_per_second);

/* ---------------------------------------------------------------------------------------------------
 * Light upload / animation path: what precedes the culling pass when lights move.
 * ------------------------------------------------------------------------------------------------- */
/* A batch of ResourcesLight::UpdateLightPosition (Source/Resources/ResourcesLight.cpp:21-29): for k < count,
 * lights[indices[k]].position = (xyz[3k], xyz[3k+1], xyz[3k+2], 1).  Host arrays; staged through double-buffered
 * pinned memory (free again on return), one scatter kernel on `stream`.  Out-of-range indices are ignored as in the
 * reference.  Indices of one batch should be unique (of duplicates closer than 64 entries the last wins). */
HVCF_API int hvcf_update_lights(hvcf_ctx* ctx, const uint32_t* indices, const float* xyz, uint32_t count, void* stream);
/* Device-side animation for the fly-through of BASELINE.json configs[4] (synthetic; the reference only moves light 0
 * from the UI, ResourcesLight.h:68-72): from now on every hvcf_cull(frame) first sets
 *     lights[i].position.y = y0[i] + amplitude * table[(step[frame] + i) & (table_len - 1)]
 * (one rounded multiply, one rounded add: a host evaluation of the same expression is bit-identical), y0 being
 * position.y at the time of this call, and writes it back into the light buffer for the fragment pass.  The update is
 * fused into the per-frame light transform: no extra pass over the lights, no host->device light traffic.
 * table: host floats, table_len a power of two; NULL / 0 switches the animation off. */
HVCF_API int hvcf_set_light_animation(hvcf_ctx* ctx, const float* table, uint32_t table_len, float amplitude, void* stream);
HVCF_API int hvcf_set_animation_step(hvcf_ctx* ctx, uint32_t frame, uint32_t step);

/* ---------------------------------------------------------------------------------------------------
 * Consumer: the light loop of the PBR cluster-forward fragment pass as a CUDA kernel.
 * Replaces, headless: PipelinePBRClusterForward::FillCommandBuffer (Source/Pipelines/PipelinePBRClusterForward.cpp
 * :88-143) for the part of Shaders/ClusteredForward/Scene.frag that reads this library's outputs -- lines 55-78
 * (gl_FragCoord -> cluster -> (count, offset)) and 103-134 (Lo = albedo*albedoMultipler + sum Radiance(light)),
 * with Shaders/ClusteredForward/Radiance.glsl:2-57 and Shaders/PBR/PBRHeader.glsl:16-64.  What rasterisation and
 * the material textures supply in the reference is a G-buffer in device memory here; ambient (IBL cubemaps) and
 * emissive need textures and are out of scope.
 * ------------------------------------------------------------------------------------------------- */
typedef struct hvcf_gbuffer
{
	uint32_t width, height;        /* pixels; p = x + y * width, gl_FragCoord.xy = (x + 0.5, y + 0.5) */
	const float* frag_depth;       /* [P]     gl_FragCoord.z in [0, 1) */
	const float* world_pos;        /* [P][4]  worldPos.xyz (Scene.frag location 0), w ignored */
	const float* normal_rough;     /* [P][4]  N.xyz after NormalTBN (:106), roughness (:101) */
	const float* albedo_metal;     /* [P][4]  linear albedo.rgb (:97), metallic (:100) */
} hvcf_gbuffer;

/* out_color: [P][4] floats (Lo.rgb, 1); out_cluster_index: [P] uint32 clusterIdx (Scene.frag:73-76) or NULL.
 * Reads ubo[frame], the light cells, the index list and the light buffer of the context (all device-resident;
 * stream order after hvcf_cull replaces the compute->fragment barriers of PipelineLightCulling.cpp:85-118).
 * All pointers are DEVICE pointers.  out_cluster_index is exact; out_color matches the shader to float
 * tolerance (documented in tests/test_gpu_shade.py). */
HVCF_API int hvcf_shade(hvcf_ctx* ctx, uint32_t frame, const hvcf_gbuffer* gbuffer, const hvcf_push_const_pbr* pc,
	const float camera_position[3], float* out_color, uint32_t* out_cluster_index, void* stream);

/* Synthetic G-buffer for the camera of hvcf_host_camera_ubo (host arrays, sizes as in hvcf_gbuffer): rays through
 * the pixel centres hit the ground plane y = 0 or a cylinder of radius 12 around the light disc; materials are hashed
 * from the pixel index and `seed`.  Pure input generation: there is no reference counterpart (the reference
 * rasterises Sponza). */
HVCF_API int hvcf_host_init_gbuffer(uint32_t seed, const float position[3], const float target[3], uint32_t width,
	uint32_t height, float* frag_depth, float* world_pos, float* normal_rough, float* albedo_metal);

/* ---------------------------------------------------------------------------------------------------
 * Compute frustum culling: Shaders/FrustumCulling.comp:14-57 over `count` world-space boxes,
 * writing instanceCount = 0/1 of each 16-byte indirect command (the other three words are left untouched).
 * Replaces PipelineFrustumCulling::Execute (Source/Pipelines/PipelineFrustumCulling.cpp:32-69).  Context-free:
 * boxes/commands are DEVICE pointers (Scene::transformedBoundingBoxBuffer_ / indirectBuffer_), `frustum` is a
 * HOST pointer (the per-frame FrustumUBO upload of PipelineFrustumCulling.h SetFrustumUBO), passed by value to
 * the kernel.
 * ------------------------------------------------------------------------------------------------- */
HVCF_API int hvcf_frustum_cull(int device, const hvcf_frustum_ubo* frustum, const hvcf_aabb* boxes, uint32_t count,
	hvcf_draw_indirect_command* commands, void* stream);
/* Camera::GetFrustumUBO (Source/Camera.cpp:199-234) for the camera of hvcf_host_camera_ubo. */
HVCF_API int hvcf_host_frustum_ubo(const float position[3], const float target[3], float screen_width, float screen_height,
	hvcf_frustum_ubo* out);
/* Synthetic instance boxes (host array): centres uniform in [-extent, extent]^3, half sizes in [0.05, 1.5]. */
HVCF_API int hvcf_host_init_boxes(uint32_t seed, uint32_t count, float extent, hvcf_aabb* out);

/* Pinned host staging for callers that want asynchronous host<->device copies
 * (VMA_MEMORY_USAGE_CPU_TO_GPU in the reference, ResourcesLight.cpp:14-16). */
HVCF_API int hvcf_host_alloc(size_t bytes, void** out_ptr);
HVCF_API int hvcf_host_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* HVCF_H */
