// hv/ClusterForward.h -- drop-in C++ surface of the reference's clustered-forward compute path, backed
// by libhvcf.so.  Same class names, method names, arity and call order as the reference:
//   ResourcesBase              Header/Resources/ResourcesBase.h:6-20
//   PipelineBase (virtuals)    Header/Pipelines/PipelineBase.h:33-48
//   ResourcesLight             Header/Resources/ResourcesLight.h:42-81, Source/Resources/ResourcesLight.cpp:3-29
//   ResourcesClusterForward    Header/Resources/ResourcesClusterForward.h:11-31, Source/.../ResourcesClusterForward.cpp:5-46
//   PipelineAABBGenerator      Header/Pipelines/PipelineAABBGenerator.h:14-39, Source/.../PipelineAABBGenerator.cpp:8-102
//   PipelineLightCulling       Header/Pipelines/PipelineLightCulling.h:15-36, Source/.../PipelineLightCulling.cpp:9-141
//   PipelinePBRClusterForward  Header/Pipelines/PipelinePBRClusterForward.h:17-67 (the light loop of Scene.frag, headless)
//   PipelineFrustumCulling     Header/Pipelines/PipelineFrustumCulling.h:9-32, Source/.../PipelineFrustumCulling.cpp:5-84
// Type substitutions only (hv/Context.h): VulkanContext -> hv::Context, VkCommandBuffer -> stream,
// VulkanBuffer -> hv::DeviceBuffer.  Grid size / cap / light capacity are run-time members that default
// to ClusterForwardConfig.  Ownership as in the reference: the app owns resources and pipelines,
// pipelines hold raw non-owning pointers, resources may be destroyed before pipelines.
#ifndef HV_CLUSTER_FORWARD_H
#define HV_CLUSTER_FORWARD_H

#include <array>
#include <span>
#include <vector>

#include "Context.h"

class ResourcesBase
{
public:
	virtual ~ResourcesBase() = default;
	virtual void Destroy() = 0;
	virtual void OnWindowResized(VulkanContext&) {}
	virtual void UpdateFromUIData(VulkanContext&, UIData&) {}
};

class PipelineBase
{
public:
	virtual ~PipelineBase() = default;
	virtual void OnWindowResized(VulkanContext&) {}
	virtual void FillCommandBuffer(VulkanContext& ctx, VkCommandBuffer commandBuffer) = 0;
	virtual void UpdateFromUIData(VulkanContext&, UIData&) {}
	virtual void SetCameraUBO(VulkanContext&, CameraUBO&) {}
};

// A collection of lights, including the device storage buffer (binding 1 of LightCulling.comp)
struct ResourcesLight : ResourcesBase
{
public:
	ResourcesLight() = default;
	~ResourcesLight() override { Destroy(); }

	void Destroy() override { storageBuffer_.Destroy(); }

	void AddLights(VulkanContext& ctx, const std::vector<LightData>& lights)
	{
		lights_ = lights;
		const size_t storageBufferSize = sizeof(LightData) * lights.size();
		lightCount_ = static_cast<uint32_t>(lights.size());
		storageBuffer_.CreateBuffer(ctx, storageBufferSize);
		storageBuffer_.UploadBufferData(ctx, lights.data(), storageBufferSize);
		ctx.WaitIdle(); // `lights` may be pageable and go away; the reference's mapped copy is synchronous too
	}

	void UpdateLightPosition(VulkanContext& ctx, size_t index, const std::span<float> position)
	{
		if (index >= lights_.size())
			return;
		lights_[index].position_ = { position[0], position[1], position[2], 1.0f };
		storageBuffer_.UploadOffsetBufferData(ctx, &(lights_[index]), sizeof(LightData) * index, sizeof(LightData));
	}

	VulkanBuffer* GetVulkanBufferPtr() { return &storageBuffer_; }
	uint32_t GetLightCount() const { return lightCount_; }

	// Header/Resources/ResourcesLight.h:68-72: light 0 is the shadow caster and follows the UI every frame
	void UpdateFromUIData(VulkanContext& ctx, UIData& uiData) override
	{
		UpdateLightPosition(ctx, 0, uiData.shadowCasterPosition_);
	}

public:
	std::vector<LightData> lights_{};

private:
	uint32_t lightCount_ = 0;
	VulkanBuffer storageBuffer_{};
};

struct ResourcesClusterForward : ResourcesBase
{
public:
	ResourcesClusterForward() = default;
	~ResourcesClusterForward() override { Destroy(); }

	// run-time replacements of the ClusterForwardConfig constants; set before CreateBuffers
	uint32_t sliceCountX_ = ClusterForwardConfig::SliceCountX;
	uint32_t sliceCountY_ = ClusterForwardConfig::SliceCountY;
	uint32_t sliceCountZ_ = ClusterForwardConfig::SliceCountZ;
	uint32_t maxLightPerCluster_ = ClusterForwardConfig::maxLightPerCluster;

	void Destroy() override
	{
		aabbBuffer_.Destroy();
		lightCellsBuffer_.Destroy();
		lightIndicesBuffer_.Destroy();
		for (VulkanBuffer& buffer : globalIndexCountBuffers_)
			buffer.Destroy();
		if (handle_)
			hvcf_destroy(handle_);
		handle_ = nullptr;
	}

	// The reference ignores lightCount; here it sizes the light capacity of the culling context.
	void CreateBuffers(VulkanContext& ctx, uint32_t lightCount)
	{
		Destroy();
		aabbDirty_ = true;
		hvcf_config cfg{};
		cfg.struct_size = sizeof(cfg);
		cfg.device = ctx.GetDeviceOrdinal();
		cfg.slice_count_x = sliceCountX_;
		cfg.slice_count_y = sliceCountY_;
		cfg.slice_count_z = sliceCountZ_;
		cfg.max_lights_per_cluster = maxLightPerCluster_;
		cfg.max_lights = lightCount ? lightCount : 1;
		cfg.frame_count = AppConfig::FrameCount;
		cfg.mode = HVCF_MODE_AUTO;
		hv::Check(hvcf_create(&cfg, &handle_), "hvcf_create");
		hvcf_buffers b{};
		hv::Check(hvcf_get_device_buffers(handle_, &b), "hvcf_get_device_buffers");
		const size_t clusterCount = b.cluster_count;
		aabbBuffer_.Wrap(b.aabb, clusterCount * sizeof(AABB));
		lightCellsBuffer_.Wrap(b.light_cells, clusterCount * sizeof(LightCell));
		lightIndicesBuffer_.Wrap(b.light_indices, static_cast<size_t>(b.light_indices_capacity) * sizeof(uint32_t));
		for (uint32_t i = 0; i < AppConfig::FrameCount; ++i)
			globalIndexCountBuffers_[i].Wrap(b.global_index_count + i, sizeof(uint32_t));
	}

	uint32_t ClusterCount() const { return sliceCountX_ * sliceCountY_ * sliceCountZ_; }
	hvcf_ctx* Handle() const { return handle_; }

public:
	bool aabbDirty_ = true;
	VulkanBuffer aabbBuffer_;
	VulkanBuffer lightCellsBuffer_;
	VulkanBuffer lightIndicesBuffer_;
	// per frame in flight; holds the frame's total index count after culling (the reference's atomic cursor)
	std::array<VulkanBuffer, AppConfig::FrameCount> globalIndexCountBuffers_;

private:
	hvcf_ctx* handle_ = nullptr;
};

// Generate voxelized frustum for clustered forward
class PipelineAABBGenerator final : public PipelineBase
{
public:
	PipelineAABBGenerator(VulkanContext&, ResourcesClusterForward* resourcesCF) : resourcesCF_(resourcesCF) {}
	~PipelineAABBGenerator() override = default;

	void SetCameraUBO(VulkanContext&, CameraUBO&) override {}

	void FillCommandBuffer(VulkanContext& ctx, VkCommandBuffer commandBuffer) override
	{
		const uint32_t frameIndex = ctx.GetFrameIndex();
		int libDirty = 0; // the library also notices projection changes the reference misses (zoom)
		hv::Check(hvcf_is_aabb_dirty(resourcesCF_->Handle(), &libDirty), "hvcf_is_aabb_dirty");
		if (!resourcesCF_->aabbDirty_ && !libDirty)
			return;
		Execute(ctx, commandBuffer, frameIndex);
		resourcesCF_->aabbDirty_ = false;
	}

	void OnWindowResized(VulkanContext&) override
	{
		resourcesCF_->aabbDirty_ = true;
		hv::Check(hvcf_mark_aabb_dirty(resourcesCF_->Handle()), "hvcf_mark_aabb_dirty");
	}

	void SetClusterForwardUBO(VulkanContext& ctx, ClusterForwardUBO& ubo)
	{
		const uint32_t frameIndex = ctx.GetFrameIndex();
		hv::Check(hvcf_set_ubo(resourcesCF_->Handle(), frameIndex, reinterpret_cast<const hvcf_cluster_forward_ubo*>(&ubo),
					  ctx.GetStream()),
			"hvcf_set_ubo");
	}

private:
	ResourcesClusterForward* resourcesCF_{};

	void Execute(VulkanContext&, VkCommandBuffer commandBuffer, uint32_t frameIndex)
	{
		hv::Check(hvcf_generate_aabbs(resourcesCF_->Handle(), frameIndex, commandBuffer), "hvcf_generate_aabbs");
	}
};

// Clustered Forward light culling
class PipelineLightCulling final : public PipelineBase
{
public:
	PipelineLightCulling(VulkanContext&, ResourcesLight* resourcesLight, ResourcesClusterForward* resourcesCF)
		: resourcesLight_(resourcesLight), resourcesCF_(resourcesCF)
	{
		// CreateDescriptor (PipelineLightCulling.cpp:121-141): bind the light SSBO in place
		hv::Check(hvcf_bind_lights(resourcesCF_->Handle(), resourcesLight_->GetVulkanBufferPtr()->buffer_,
					  resourcesLight_->GetLightCount()),
			"hvcf_bind_lights");
	}
	~PipelineLightCulling() override = default;

	void SetCameraUBO(VulkanContext&, CameraUBO&) override {}

	void FillCommandBuffer(VulkanContext& ctx, VkCommandBuffer commandBuffer) override
	{
		const uint32_t frameIndex = ctx.GetFrameIndex();
		Execute(ctx, commandBuffer, frameIndex);
	}

	// The counter is reset on the device at the head of hvcf_cull (stream-ordered); kept for call-order parity
	// with AppPBRClusterForward::UpdateUBOs (AppPBRClusterForward.cpp:130-133).
	void ResetGlobalIndex(VulkanContext&) {}

	void SetClusterForwardUBO(VulkanContext& ctx, ClusterForwardUBO& ubo)
	{
		const uint32_t frameIndex = ctx.GetFrameIndex();
		hv::Check(hvcf_set_ubo(resourcesCF_->Handle(), frameIndex, reinterpret_cast<const hvcf_cluster_forward_ubo*>(&ubo),
					  ctx.GetStream()),
			"hvcf_set_ubo");
	}

private:
	ResourcesLight* resourcesLight_;
	ResourcesClusterForward* resourcesCF_;

	void Execute(VulkanContext&, VkCommandBuffer commandBuffer, uint32_t frameIndex)
	{
		// re-bind every frame (a pointer store): AddLights may have re-created the storage buffer
		hv::Check(hvcf_bind_lights(resourcesCF_->Handle(), resourcesLight_->GetVulkanBufferPtr()->buffer_,
					  resourcesLight_->GetLightCount()),
			"hvcf_bind_lights");
		hv::Check(hvcf_cull(resourcesCF_->Handle(), frameIndex, commandBuffer), "hvcf_cull");
	}
};

// ---- rows next to the culling path ---------------------------------------------------------

// What the rasteriser and the material textures hand the fragment shader in the reference, as device arrays
// (include/hvcf.h hvcf_gbuffer); plus the colour attachment the pass writes (ResourcesShared::multiSampledColorImage_).
struct ResourcesGBufferCF : ResourcesBase
{
	~ResourcesGBufferCF() override { Destroy(); }
	void Destroy() override
	{
		fragDepth_.Destroy(), worldPos_.Destroy(), normalRough_.Destroy(), albedoMetal_.Destroy(), color_.Destroy(), clusterIndex_.Destroy();
	}
	// host arrays as hvcf_host_init_gbuffer fills them
	void Create(VulkanContext& ctx, uint32_t width, uint32_t height, const float* fragDepth, const float* worldPos,
		const float* normalRough, const float* albedoMetal)
	{
		width_ = width, height_ = height;
		const size_t P = static_cast<size_t>(width) * height;
		fragDepth_.CreateBuffer(ctx, 4 * P), worldPos_.CreateBuffer(ctx, 16 * P), normalRough_.CreateBuffer(ctx, 16 * P);
		albedoMetal_.CreateBuffer(ctx, 16 * P), color_.CreateBuffer(ctx, 16 * P), clusterIndex_.CreateBuffer(ctx, 4 * P);
		fragDepth_.UploadBufferData(ctx, fragDepth, 4 * P), worldPos_.UploadBufferData(ctx, worldPos, 16 * P);
		normalRough_.UploadBufferData(ctx, normalRough, 16 * P), albedoMetal_.UploadBufferData(ctx, albedoMetal, 16 * P);
		ctx.WaitIdle();
	}
	uint32_t width_ = 0, height_ = 0;
	VulkanBuffer fragDepth_, worldPos_, normalRough_, albedoMetal_, color_, clusterIndex_;
};

// Render using PBR materials, clustered forward renderer: the light loop of Scene.frag over the G-buffer
// (Header/Pipelines/PipelinePBRClusterForward.h:17-67; Scene*, ResourcesIBL*, ResourcesShared* and the material pass
// arguments of the reference constructor are the G-buffer here)
class PipelinePBRClusterForward final : public PipelineBase
{
public:
	PipelinePBRClusterForward(VulkanContext&, ResourcesGBufferCF* gbuffer, ResourcesLight* lights, ResourcesClusterForward* resourcesCF)
		: resourcesCF_(resourcesCF), resourcesLight_(lights), gbuffer_(gbuffer)
	{
	}
	~PipelinePBRClusterForward() override = default;

	void SetCameraUBO(VulkanContext&, CameraUBO& ubo) override { cameraPosition_ = ubo.position; }
	void SetPBRPushConstants(const PushConstPBR& pbrPC) { pc_ = pbrPC; }
	// the culling pipelines already uploaded this frame's UBO into the shared context (one copy instead of the
	// reference's three, PipelinePBRClusterForward.h:33-37); kept for call-order parity
	void SetClusterForwardUBO(VulkanContext&, ClusterForwardUBO&) {}

	void FillCommandBuffer(VulkanContext& ctx, VkCommandBuffer commandBuffer) override
	{
		hvcf_gbuffer g{};
		g.width = gbuffer_->width_, g.height = gbuffer_->height_;
		g.frag_depth = static_cast<const float*>(gbuffer_->fragDepth_.buffer_);
		g.world_pos = static_cast<const float*>(gbuffer_->worldPos_.buffer_);
		g.normal_rough = static_cast<const float*>(gbuffer_->normalRough_.buffer_);
		g.albedo_metal = static_cast<const float*>(gbuffer_->albedoMetal_.buffer_);
		const float cam[3] = { cameraPosition_.x, cameraPosition_.y, cameraPosition_.z };
		hv::Check(hvcf_shade(resourcesCF_->Handle(), ctx.GetFrameIndex(), &g, reinterpret_cast<const hvcf_push_const_pbr*>(&pc_), cam,
					  static_cast<float*>(gbuffer_->color_.buffer_), static_cast<uint32_t*>(gbuffer_->clusterIndex_.buffer_), commandBuffer),
			"hvcf_shade");
	}

private:
	PushConstPBR pc_{};
	ResourcesClusterForward* resourcesCF_;
	ResourcesLight* resourcesLight_;
	ResourcesGBufferCF* gbuffer_;
	hv::vec4 cameraPosition_{};
};

// The three members of the reference Scene that frustum culling touches (Header/Scene/Scene.h: transformedBoundingBoxBuffer_,
// indirectBuffer_, GetInstanceCount())
struct Scene
{
	VulkanBuffer transformedBoundingBoxBuffer_; // AABB[instanceCount], world space
	VulkanBuffer indirectBuffer_;               // DrawIndirectCommand[instanceCount]
	uint32_t instanceCount_ = 0;
	uint32_t GetInstanceCount() const { return instanceCount_; }
	void Create(VulkanContext& ctx, const std::vector<AABB>& boxes, const std::vector<DrawIndirectCommand>& commands)
	{
		instanceCount_ = static_cast<uint32_t>(boxes.size());
		transformedBoundingBoxBuffer_.CreateBuffer(ctx, sizeof(AABB) * boxes.size());
		indirectBuffer_.CreateBuffer(ctx, sizeof(DrawIndirectCommand) * commands.size());
		transformedBoundingBoxBuffer_.UploadBufferData(ctx, boxes.data(), sizeof(AABB) * boxes.size());
		indirectBuffer_.UploadBufferData(ctx, commands.data(), sizeof(DrawIndirectCommand) * commands.size());
		ctx.WaitIdle();
	}
	void Destroy() { transformedBoundingBoxBuffer_.Destroy(), indirectBuffer_.Destroy(); }
};

// Compute-based frustum culling (Header/Pipelines/PipelineFrustumCulling.h:9-32, Source/.../PipelineFrustumCulling.cpp:5-84)
class PipelineFrustumCulling final : public PipelineBase
{
public:
	PipelineFrustumCulling(VulkanContext&, Scene* scene) : scene_(scene) {}
	~PipelineFrustumCulling() override = default;

	void SetCameraUBO(VulkanContext&, CameraUBO&) override {}
	void SetFrustumUBO(VulkanContext& ctx, FrustumUBO& ubo) { frustums_[ctx.GetFrameIndex()] = ubo; }

	void FillCommandBuffer(VulkanContext& ctx, VkCommandBuffer commandBuffer) override { Execute(ctx, commandBuffer, ctx.GetFrameIndex()); }

private:
	void Execute(VulkanContext& ctx, VkCommandBuffer commandBuffer, uint32_t frameIndex)
	{
		hv::Check(hvcf_frustum_cull(ctx.GetDeviceOrdinal(), reinterpret_cast<const hvcf_frustum_ubo*>(&frustums_[frameIndex]),
					  static_cast<const hvcf_aabb*>(scene_->transformedBoundingBoxBuffer_.buffer_), scene_->GetInstanceCount(),
					  static_cast<hvcf_draw_indirect_command*>(scene_->indirectBuffer_.buffer_), commandBuffer),
			"hvcf_frustum_cull");
	}

	Scene* scene_;
	std::array<FrustumUBO, AppConfig::FrameCount> frustums_{};
};

#endif
