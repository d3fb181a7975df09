// hv/ClusterForward.h -- drop-in C++ surface of the reference's clustered-forward compute path, backed
// by libhvcf.so.  Same class names, method names, arity and call order as the reference:
//   ResourcesBase              Header/Resources/ResourcesBase.h:6-20
//   PipelineBase (virtuals)    Header/Pipelines/PipelineBase.h:33-48
//   ResourcesLight             Header/Resources/ResourcesLight.h:42-81, Source/Resources/ResourcesLight.cpp:3-29
//   ResourcesClusterForward    Header/Resources/ResourcesClusterForward.h:11-31, Source/.../ResourcesClusterForward.cpp:5-46
//   PipelineAABBGenerator      Header/Pipelines/PipelineAABBGenerator.h:14-39, Source/.../PipelineAABBGenerator.cpp:8-102
//   PipelineLightCulling       Header/Pipelines/PipelineLightCulling.h:15-36, Source/.../PipelineLightCulling.cpp:9-141
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

#endif
