// hv/Types.h -- C++ names of the reference's PODs and compile-time configs for the light-culling path.
// Layout-identical to include/hvcf_types.h (the ABI) and to the reference structs:
//   LightData / AABB / LightCell   HelloVulkan/Header/Resources/ResourcesLight.h:13-39
//   ClusterForwardUBO, FrustumUBO  HelloVulkan/Header/UBOs.h:67-99
//   PushConstPBR                   HelloVulkan/Header/PushConstants.h:24-31
//   AppConfig / CameraConfig / ClusterForwardConfig   HelloVulkan/Header/Configs.h:8-56
// glm is not shipped with this library; hv::vec/mat below are the minimal column-major stand-ins.
#ifndef HV_TYPES_H
#define HV_TYPES_H

#include <cstddef>
#include <array>
#include <cstdint>

#include "../hvcf_types.h"

namespace hv
{
struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4
{
	float x, y, z, w;
	float& operator[](int i) { return (&x)[i]; }
	const float& operator[](int i) const { return (&x)[i]; }
};
struct mat4 // column-major: m[col][row], as glm
{
	vec4 c[4];
	vec4& operator[](int i) { return c[i]; }
	const vec4& operator[](int i) const { return c[i]; }
};
} // namespace hv

namespace AppConfig
{
	constexpr uint32_t FrameCount = 2;             // Configs.h:14
	constexpr uint32_t InitialScreenWidth = 1600;  // Configs.h:10
	constexpr uint32_t InitialScreenHeight = 1200; // Configs.h:11
}

namespace CameraConfig
{
	constexpr float Yaw = -1.5708f;
	constexpr float Pitch = 0.0f;
	constexpr float Zoom = 0.785398f;
	constexpr float Speed = 2.5f;
	constexpr float Sensitivity = 0.1f;
	constexpr float Near = 0.1f;
	constexpr float Far = 100.0f;
}

namespace ClusterForwardConfig
{
	constexpr uint32_t SliceCountX = 16;
	constexpr uint32_t SliceCountY = 9;
	constexpr uint32_t SliceCountZ = 24;
	constexpr uint32_t ClusterCount = SliceCountX * SliceCountY * SliceCountZ;
	constexpr uint32_t maxLightPerCluster = 150;
}

struct LightData
{
	alignas(16) hv::vec4 position_;
	alignas(16) hv::vec4 color_ = { 1.0f, 1.0f, 1.0f, 1.0f };
	alignas(4) float radius_ = 1.0f;
};

struct AABB
{
	alignas(16) hv::vec4 minPoint;
	alignas(16) hv::vec4 maxPoint;
};

struct LightCell
{
	alignas(4) uint32_t offset;
	alignas(4) uint32_t count;
};

struct ClusterForwardUBO
{
	alignas(16) hv::mat4 cameraInverseProjection;
	alignas(16) hv::mat4 cameraView;
	alignas(8) hv::vec2 screenSize;
	alignas(4) float sliceScaling;
	alignas(4) float sliceBias;
	alignas(4) float cameraNear;
	alignas(4) float cameraFar;
	alignas(4) uint32_t sliceCountX;
	alignas(4) uint32_t sliceCountY;
	alignas(4) uint32_t sliceCountZ;
};

// The fragment-pass camera block; the two compute pipelines ignore it (PipelineAABBGenerator.h:20).
struct CameraUBO
{
	alignas(16) hv::mat4 projection;
	alignas(16) hv::mat4 view;
	alignas(16) hv::vec4 position;
	alignas(4) float cameraNear;
	alignas(4) float cameraFar;
};

// For compute-based frustum culling (HelloVulkan/Header/UBOs.h:92-99)
struct FrustumUBO
{
	alignas(16) hv::vec4 planes[6];
	alignas(16) hv::vec4 corners[8];
};

// Additional customization for PBR (HelloVulkan/Header/PushConstants.h:24-31)
struct PushConstPBR
{
	float lightIntensity = 1.f;
	float baseReflectivity = 0.04f;
	float maxReflectionLod = 4.f;
	float lightFalloff = 1.0f;
	float albedoMultipler = 0.0f;
};

// VkDrawIndirectCommand as the culling shader sees it (HelloVulkan/Shaders/DrawIndirectCommand.glsl)
struct DrawIndirectCommand
{
	uint32_t vertexCount;
	uint32_t instanceCount;
	uint32_t firstVertex;
	uint32_t firstInstance;
};

// The members of the UI state that this path reads (HelloVulkan/Header/UIData.h:19-36): the PBR push constants consumed by
// PipelinePBRClusterForward and the shadow-caster position that ResourcesLight::UpdateFromUIData writes into light 0.
// Mouse / gizmo / ImGui state is out of scope (no window on this path).
struct UIData
{
public:
	PushConstPBR pbrPC_ =
	{
		.lightIntensity = 1.75f,
		.baseReflectivity = 0.01f,
		.maxReflectionLod = 1.f,
		.lightFalloff = 0.1f,
		.albedoMultipler = 0.1f
	};

	std::array<float, 3> shadowCasterPosition_ = { 0.f, 0.f, 0.f };
};

static_assert(sizeof(LightData) == sizeof(hvcf_light_data) && offsetof(LightData, radius_) == offsetof(hvcf_light_data, radius));
static_assert(sizeof(AABB) == sizeof(hvcf_aabb) && sizeof(LightCell) == sizeof(hvcf_light_cell));
static_assert(sizeof(ClusterForwardUBO) == sizeof(hvcf_cluster_forward_ubo));
static_assert(sizeof(FrustumUBO) == sizeof(hvcf_frustum_ubo) && sizeof(PushConstPBR) == sizeof(hvcf_push_const_pbr));
static_assert(sizeof(DrawIndirectCommand) == sizeof(hvcf_draw_indirect_command));
static_assert(offsetof(ClusterForwardUBO, screenSize) == 128 && offsetof(ClusterForwardUBO, sliceCountZ) == 160);

#endif
