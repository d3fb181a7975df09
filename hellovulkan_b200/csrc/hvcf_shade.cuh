// hvcf_shade.cuh -- launch interface of the rows next to the culling path:
//   cluster_shade_kernel   the light loop of Shaders/ClusteredForward/Scene.frag over a device G-buffer
//   frustum_cull_kernel    Shaders/FrustumCulling.comp
//   light_animate_kernel   device-side light animation of the fly-through (BASELINE.json configs[4])
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/hvcf_types.h"

namespace hvcf
{

struct ShadeParams
{
	const hvcf_light_cell* cells;   // whole grid
	const uint32_t* indices;
	const hvcf_light_data* lights;  // 48 B AoS, as the fragment shader reads them (Scene.frag:122)
	const float* slice_thresholds;  // [Z + 1]; [k] = smallest linear depth whose zIndex is >= k ([0] unused)
	hvcf_push_const_pbr pc;
	float cam[3];
	float near_plane, far_plane;    // cfUBO.cameraNear / cameraFar
	float slice_scaling, slice_bias;
	float tile_w, tile_h;           // cfUBO.screenSize / float(sliceCount), Scene.frag:66-68 (host float division)
	uint32_t X, Y, Z;
	uint32_t width, height;
	const float* frag_depth;
	const float4* world_pos;
	const float4* normal_rough;
	const float4* albedo_metal;
	float4* out_color;
	uint32_t* out_cluster;          // may be null
};

// FrustumCulling.comp: planes as uploaded; the eight corners reduced on the host to per-axis min/max
// (all 8 corners > box.max  <=>  min corner > box.max; NaN corners disable the test exactly as in the shader)
struct FrustumParams
{
	float planes[6][4];
	float corner_min[3];
	float corner_max[3];
};

cudaError_t launch_cluster_shade(const ShadeParams& p, cudaStream_t stream);
cudaError_t launch_frustum_cull(const FrustumParams& f, const hvcf_aabb* boxes, uint32_t count,
	hvcf_draw_indirect_command* commands, cudaStream_t stream);

} // namespace hvcf
