/*
 * hvcf_types.h -- plain-C POD layouts of the clustered-forward light-culling path.
 *
 * These structs ARE the ABI: every byte offset below is the one the reference's GLSL
 * (std430 SSBOs / std140 UBO) and its C++ mirror structs use, so a buffer produced by
 * this library can be bound to the reference's fragment pass unchanged.
 *
 * Reference layouts (paths relative to /root/reference/HelloVulkan):
 *   LightData          Header/Resources/ResourcesLight.h:13-21, Shaders/LightData.glsl:1-6
 *   AABB               Header/Resources/ResourcesLight.h:24-30, Shaders/AABB/AABB.glsl:1-5
 *   LightCell          Header/Resources/ResourcesLight.h:33-39, Shaders/ClusteredForward/Header.glsl:1-5
 *   ClusterForwardUBO  Header/UBOs.h:67-89,                     Shaders/ClusteredForward/Header.glsl:7-19
 */
#ifndef HVCF_TYPES_H
#define HVCF_TYPES_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
#define HVCF_ALIGNAS(n) alignas(n)
#define HVCF_STATIC_ASSERT(c, m) static_assert(c, m)
extern "C" {
#else
#define HVCF_ALIGNAS(n) _Alignas(n)
#define HVCF_STATIC_ASSERT(c, m) _Static_assert(c, m)
#endif

/* One point light: 48-byte stride; culling reads position (16 B) and radius (4 B). */
typedef struct hvcf_light_data
{
	HVCF_ALIGNAS(16) float position[4]; /* world space, w = 1 */
	HVCF_ALIGNAS(16) float color[4];
	HVCF_ALIGNAS(4) float radius;
	float pad_[3];
} hvcf_light_data;

/* One cluster box in VIEW space (camera looks down -z); .w of both points is 0. */
typedef struct hvcf_aabb
{
	HVCF_ALIGNAS(16) float min_point[4];
	HVCF_ALIGNAS(16) float max_point[4];
} hvcf_aabb;

/* Per-cluster (offset, count) into the light-index list; cluster = x + y*X + z*X*Y. */
typedef struct hvcf_light_cell
{
	uint32_t offset;
	uint32_t count;
} hvcf_light_cell;

/* Per-frame constants; matrices are column-major (glm / GLSL convention). */
typedef struct hvcf_cluster_forward_ubo
{
	HVCF_ALIGNAS(16) float camera_inverse_projection[16];
	HVCF_ALIGNAS(16) float camera_view[16];
	HVCF_ALIGNAS(8) float screen_size[2];
	float slice_scaling;
	float slice_bias;
	float camera_near;
	float camera_far;
	uint32_t slice_count_x;
	uint32_t slice_count_y;
	uint32_t slice_count_z;
	uint32_t pad_[3];
} hvcf_cluster_forward_ubo;

HVCF_STATIC_ASSERT(sizeof(hvcf_light_data) == 48, "LightData must be 48 bytes");
HVCF_STATIC_ASSERT(offsetof(hvcf_light_data, color) == 16, "LightData.color @16");
HVCF_STATIC_ASSERT(offsetof(hvcf_light_data, radius) == 32, "LightData.radius @32");
HVCF_STATIC_ASSERT(sizeof(hvcf_aabb) == 32, "AABB must be 32 bytes");
HVCF_STATIC_ASSERT(offsetof(hvcf_aabb, max_point) == 16, "AABB.maxPoint @16");
HVCF_STATIC_ASSERT(sizeof(hvcf_light_cell) == 8, "LightCell must be 8 bytes");
HVCF_STATIC_ASSERT(sizeof(hvcf_cluster_forward_ubo) == 176, "ClusterForwardUBO must be 176 bytes");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, camera_view) == 64, "ubo.cameraView @64");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, screen_size) == 128, "ubo.screenSize @128");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, slice_scaling) == 136, "ubo.sliceScaling @136");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, slice_bias) == 140, "ubo.sliceBias @140");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, camera_near) == 144, "ubo.cameraNear @144");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, camera_far) == 148, "ubo.cameraFar @148");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, slice_count_x) == 152, "ubo.sliceCountX @152");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, slice_count_y) == 156, "ubo.sliceCountY @156");
HVCF_STATIC_ASSERT(offsetof(hvcf_cluster_forward_ubo, slice_count_z) == 160, "ubo.sliceCountZ @160");

/* PushConstPBR (Header/PushConstants.h:24-31; GLSL Shaders/PBR/PBRPushConstants.glsl): consumer-side constants. */
typedef struct hvcf_push_const_pbr
{
	float light_intensity;   /* default 1 */
	float base_reflectivity; /* default 0.04 */
	float max_reflection_lod; /* default 4 (unused by the light loop) */
	float light_falloff;     /* default 1 */
	float albedo_multipler;  /* default 0 (sic: the reference spells it this way) */
} hvcf_push_const_pbr;

/* FrustumUBO (Header/UBOs.h:92-99; GLSL Shaders/Frustum.glsl): six planes + eight corners, std140. */
typedef struct hvcf_frustum_ubo
{
	HVCF_ALIGNAS(16) float planes[6][4];
	HVCF_ALIGNAS(16) float corners[8][4];
} hvcf_frustum_ubo;

/* VkDrawIndirectCommand as Shaders/DrawIndirectCommand.glsl declares it (16 B). */
typedef struct hvcf_draw_indirect_command
{
	uint32_t vertex_count;
	uint32_t instance_count;
	uint32_t first_vertex;
	uint32_t first_instance;
} hvcf_draw_indirect_command;

HVCF_STATIC_ASSERT(sizeof(hvcf_push_const_pbr) == 20, "PushConstPBR must be 20 bytes");
HVCF_STATIC_ASSERT(sizeof(hvcf_frustum_ubo) == 224, "FrustumUBO must be 224 bytes");
HVCF_STATIC_ASSERT(sizeof(hvcf_draw_indirect_command) == 16, "DrawIndirectCommand must be 16 bytes");

/* Defaults of the reference (Header/Configs.h:33-56, :8-14). */
#define HVCF_DEFAULT_SLICE_COUNT_X 16u
#define HVCF_DEFAULT_SLICE_COUNT_Y 9u
#define HVCF_DEFAULT_SLICE_COUNT_Z 24u
#define HVCF_DEFAULT_MAX_LIGHTS_PER_CLUSTER 150u
#define HVCF_DEFAULT_FRAME_COUNT 2u

#ifdef __cplusplus
}
#endif
#endif /* HVCF_TYPES_H */
