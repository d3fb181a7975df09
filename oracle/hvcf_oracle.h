/*
 * hvcf_oracle.h -- CPU oracle for the clustered-forward light-culling path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under hellovulkan_b200/ or include/ may include, link or
 * dlopen this; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs use it, and only as the checker / the timed CPU baseline.
 *
 * It is a plain-C restatement of the reference's algorithm (paths relative to
 * /root/reference/HelloVulkan unless noted):
 *   camera -> UBO   Source/Camera.cpp:81-93,113-139,178-197, Header/Configs.h:33-56, glm 0.9.9.9
 *   AABB generation Shaders/ClusteredForward/AABBGenerator.comp:18-90
 *   light culling   Shaders/ClusteredForward/LightCulling.comp:26-88
 *   scene           Source/Apps/AppPBRClusterForward.cpp:81-118, Header/Utility.h:9-20
 *
 * Semantics fixed by the oracle: IEEE-754 binary32, round-to-nearest-even, NO
 * FMA contraction (build with -ffp-contract=off), mat4*vec4 associated as glm does
 * ((m0*x + m1*y) + (m2*z + m3*w), External/glm/glm/detail/type_mat4x4.inl:561-572), true division,
 * pow() = this libm's powf.
 *
 * PARITY PIN: the reference has no tests, golden vectors or CPU implementation of this path.  The
 * oracle is pinned against the reference's OWN sources compiled here as C++ (oracle/refbuild ->
 * oracle/_ref/libhvcf_ref.so: Camera.cpp + the two .comp shaders against the vendored glm) and
 * against fixtures generated from that build (tests/golden/, scripts alongside).
 */
#ifndef HVCF_ORACLE_H
#define HVCF_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#include "../include/hvcf_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Seeded restatement of AppPBRClusterForward::InitLights: mt19937(seed) +
 * uniform_real_distribution<float>(0,1) as libstdc++ implements it; draw order
 * yPos, radius, rad, colour r,g,b, light radius. */
void hvo_init_lights(uint32_t seed, uint32_t count, hvcf_light_data* out);

/* Camera((0,0,3)) -> SetScreenSize -> SetPositionAndTarget -> GetClusterForwardUBO. */
void hvo_camera_ubo(const float pos[3], const float target[3], float width, float height,
	uint32_t X, uint32_t Y, uint32_t Z, hvcf_cluster_forward_ubo* out);

/* AABBGenerator.comp over the whole grid of ubo->slice_count_{x,y,z}. */
void hvo_generate_aabbs(const hvcf_cluster_forward_ubo* ubo, hvcf_aabb* out);

/* posView = cameraView * position; out[i] = {x, y, z, radius*radius} (LightCulling.comp:51-56). */
void hvo_transform_lights(const hvcf_cluster_forward_ubo* ubo, const hvcf_light_data* lights,
	uint32_t count, float* out_xyzr2);

/*
 * LightCulling.comp for clusters [cluster_begin, cluster_end) of `aabbs` (indexed by global cluster
 * id).  Offsets are the exclusive scan of counts in cluster order starting at 0 for cluster_begin
 * (the reference's atomicAdd order is non-deterministic).
 *   cells            (cluster_end - cluster_begin) entries
 *   indices          capacity max_per_cluster * (cluster_end - cluster_begin)
 *   out_total        sum of counts (may be NULL)
 *   out_executed     sum over clusters of the number of loop iterations the reference shader
 *                    executes (index of the max_per_cluster-th hit + 1, else num_lights) (may be NULL)
 *   threads          OpenMP threads (<= 1: serial)
 * use_fma != 0 evaluates sqDist with contraction (fmaf) -- ONLY for the boundary report.
 */
void hvo_cull(const hvcf_cluster_forward_ubo* ubo, const hvcf_aabb* aabbs, const hvcf_light_data* lights,
	uint32_t num_lights, uint32_t max_per_cluster, uint32_t cluster_begin, uint32_t cluster_end,
	hvcf_light_cell* cells, uint32_t* indices, uint64_t* out_total, uint64_t* out_executed,
	int threads, int use_fma);

/* One cluster only: writes up to max_per_cluster ascending indices, returns the count. */
uint32_t hvo_cull_one(const hvcf_cluster_forward_ubo* ubo, const hvcf_aabb* aabb, const hvcf_light_data* lights,
	uint32_t num_lights, uint32_t max_per_cluster, uint32_t* out_indices, uint64_t* out_executed);

/* Boundary report over clusters [begin,end) x all lights (uncapped): pairs whose sqDist is within
 * 1 ulp of r*r, and pairs whose hit/miss decision flips under FMA contraction. */
void hvo_boundary_report(const hvcf_cluster_forward_ubo* ubo, const hvcf_aabb* aabbs,
	const hvcf_light_data* lights, uint32_t num_lights, uint32_t cluster_begin, uint32_t cluster_end,
	uint64_t* out_within_1ulp, uint64_t* out_fma_flips, int threads);

/* Consumer contract, Shaders/ClusteredForward/Scene.frag:55-78: fragment (x+0.5, y+0.5, depth) ->
 * cluster index. */
uint32_t hvo_cluster_index(const hvcf_cluster_forward_ubo* ubo, float frag_x, float frag_y, float frag_depth);

/*
 * Consumer: the light loop of the PBR cluster-forward fragment shader.
 *   Shaders/ClusteredForward/Scene.frag:55-78   LinearDepth, zIndex, tile -> clusterIdx, (count, offset)
 *   Scene.frag:103,107-108,112-113,117,119-134   alphaRoughness, V, NoV, F0, Lo = albedo*albedoMultipler + sum Radiance
 *   Shaders/ClusteredForward/Radiance.glsl:2-57, Shaders/PBR/PBRHeader.glsl:16-64
 * What the rasteriser and the material textures supply in the reference is a device-side G-buffer here
 * (one record per pixel, row-major, gl_FragCoord = (x + 0.5, y + 0.5, frag_depth)):
 *   frag_depth[p]      gl_FragCoord.z in [0, 1)
 *   world_pos[p][4]    worldPos.xyz, w ignored
 *   normal_rough[p][4] N.xyz (after NormalTBN), roughness
 *   albedo_metal[p][4] albedo.rgb (after the sRGB->linear pow), metallic
 * Outputs for pixels [pixel_begin, pixel_end): out_color[p] = (Lo.rgb, 1), out_cluster[p] = clusterIdx.
 * Ambient (IBL cubemaps) and emissive need textures and are out of scope.  A zIndex beyond the grid (only
 * reachable with frag_depth == 1) is clamped to Z-1 so that nothing is read out of bounds.
 * Float parity of out_color is by tolerance (pow/sqrt/divide orders), out_cluster is exact.
 */
void hvo_shade(const hvcf_cluster_forward_ubo* ubo, const hvcf_light_cell* cells, const uint32_t* indices,
	const hvcf_light_data* lights, const hvcf_push_const_pbr* pc, const float camera_position[3],
	uint32_t width, uint32_t height, uint64_t pixel_begin, uint64_t pixel_end, uint64_t pixel_stride,
	const float* frag_depth, const float* world_pos, const float* normal_rough, const float* albedo_metal,
	float* out_color, uint32_t* out_cluster, int threads);

/* Camera::GetFrustumUBO (Source/Camera.cpp:199-234) for the camera of hvo_camera_ubo. */
void hvo_frustum_ubo(const float pos[3], const float target[3], float width, float height, hvcf_frustum_ubo* out);

/* FrustumCulling.comp:14-57: instance_count = IsBoxInFrustum ? 1 : 0 for boxes [0, count). */
void hvo_frustum_cull(const hvcf_frustum_ubo* frustum, const hvcf_aabb* boxes, uint32_t count,
	hvcf_draw_indirect_command* commands, int threads);

int hvo_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
