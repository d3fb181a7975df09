// hvcf_kernels.cuh -- launch interface between the C ABI (hvcf_api.cu) and the kernels.
#pragma once

#include "hvcf_device.cuh"

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

cudaError_t launch_aabb_generate(const hvcf_cluster_forward_ubo* ubo, const float* zplanes, hvcf_aabb* out,
	uint32_t cluster_count, cudaStream_t stream);
cudaError_t launch_light_transform(const hvcf_cluster_forward_ubo* ubo, const hvcf_light_data* lights,
	uint32_t num_lights, uint32_t padded_lights, float4* out, cudaStream_t stream);
// variant: 0..2 = packed FP32x2 with 1/2/4 clusters per warp, 3..5 = scalar FP32 with 1/2/4
constexpr int kNumCullVariants = 6;
uint32_t cull_variant_clusters_per_cta(int variant);
cudaError_t launch_cull(int variant, const CullParams& p, uint32_t num_segs, cudaStream_t stream);
cudaError_t launch_count(const uint32_t* seg_counts, uint32_t num_segs, uint32_t cluster_count, uint32_t cap,
	uint32_t* counts, uint32_t* block_sums, DeviceStats* stats, cudaStream_t stream);
cudaError_t launch_scan(const uint32_t* block_sums, uint32_t num_blocks, uint32_t* block_offsets,
	uint32_t* shard_total, DeviceStats* stats, cudaStream_t stream);
cudaError_t launch_scatter(const ScatterParams& p, cudaStream_t stream);
// 16 lane-ops per thread per iteration (8 FMUL + 8 FADD)
cudaError_t launch_fp32_issue_rate(float* sink, uint32_t blocks, uint32_t iters, cudaStream_t stream);

} // namespace hvcf
