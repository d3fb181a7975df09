// app_cluster_forward.cpp -- headless driver written in the shape of AppPBRClusterForward
// (HelloVulkan/Source/Apps/AppPBRClusterForward.cpp:19-79 Init, :81-118 InitLights, :120-136 UpdateUBOs,
// :160-182 MainLoop; AppBase::DrawFrame / FillCommandBuffer, Source/Apps/AppBase.cpp:94-229), reduced to the
// two compute pipelines of the light-culling path.  No window, swapchain or UI: frames are a counted loop
// and the results are written to files so a test can compare them with the oracle.
//
// usage: hv_cluster_forward_app <X> <Y> <Z> <width> <height> <lights> <seed> <frames> <out_prefix>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "../../include/hv/Camera.h"
#include "../../include/hv/ClusterForward.h"


class AppPBRClusterForward
{
public:
	AppPBRClusterForward(uint32_t X, uint32_t Y, uint32_t Z, float width, float height, uint32_t lightCount, uint32_t seed)
		: vulkanContext_(-1, nullptr), X_(X), Y_(Y), Z_(Z), width_(width), height_(height), lightCount_(lightCount), seed_(seed)
	{
		camera_ = std::make_unique<Camera>(hv::vec3{ 0.0f, 0.0f, 3.0f }); // AppBase::InitCamera
		camera_->SetSliceCounts(X, Y, Z);
		camera_->SetScreenSize(width, height);
	}

	void Init()
	{
		camera_->SetPositionAndTarget({ 0.0f, 1.0f, 6.0f }, { 0.0f, 2.5f, 0.0f });
		InitLights();
		resourcesCF_ = std::make_unique<ResourcesClusterForward>();
		resourcesCF_->sliceCountX_ = X_;
		resourcesCF_->sliceCountY_ = Y_;
		resourcesCF_->sliceCountZ_ = Z_;
		resourcesCF_->CreateBuffers(vulkanContext_, resourcesLight_->GetLightCount());
		auto aabb = std::make_unique<PipelineAABBGenerator>(vulkanContext_, resourcesCF_.get());
		auto cull = std::make_unique<PipelineLightCulling>(vulkanContext_, resourcesLight_.get(), resourcesCF_.get());
		aabbPtr_ = aabb.get();
		lightCullPtr_ = cull.get();
		pipelines_.push_back(std::move(aabb));
		pipelines_.push_back(std::move(cull));
	}

	void InitLights()
	{
		std::vector<LightData> lights(lightCount_);
		hv::Check(hvcf_host_init_lights(seed_, lightCount_, reinterpret_cast<hvcf_light_data*>(lights.data())), "hvcf_host_init_lights");
		resourcesLight_ = std::make_unique<ResourcesLight>();
		resourcesLight_->AddLights(vulkanContext_, lights);
	}

	void UpdateUBOs()
	{
		CameraUBO ubo = camera_->GetCameraUBO();
		for (auto& pipeline : pipelines_)
			pipeline->SetCameraUBO(vulkanContext_, ubo);
		ClusterForwardUBO cfUBO = camera_->GetClusterForwardUBO();
		aabbPtr_->SetClusterForwardUBO(vulkanContext_, cfUBO);
		lightCullPtr_->ResetGlobalIndex(vulkanContext_);
		lightCullPtr_->SetClusterForwardUBO(vulkanContext_, cfUBO);
	}

	void DrawFrame()
	{
		UpdateUBOs();
		for (auto& pipeline : pipelines_) // AppBase::FillCommandBuffer, insertion order
			pipeline->FillCommandBuffer(vulkanContext_, vulkanContext_.GetStream());
		vulkanContext_.IncrementFrameIndex();
	}

	// frame f > 0 moves light 0 (the UI's shadow-caster slider, ResourcesLight.h:68-72) and orbits the camera
	void MainLoop(uint32_t frames)
	{
		Init();
		for (uint32_t f = 0; f < frames; ++f)
		{
			if (f > 0)
			{
				float p[3] = { 0.25f * static_cast<float>(f), 2.0f, 1.0f };
				resourcesLight_->UpdateLightPosition(vulkanContext_, 0, std::span<float>(p, 3));
				camera_->SetPositionAndTarget({ 0.5f * static_cast<float>(f), 1.0f, 6.0f }, { 0.0f, 2.5f, 0.0f });
			}
			lastFrameIndex_ = vulkanContext_.GetFrameIndex();
			DrawFrame();
		}
		vulkanContext_.WaitIdle();
	}

	void Dump(const std::string& prefix)
	{
		const size_t C = resourcesCF_->ClusterCount();
		std::vector<AABB> aabbs(C);
		std::vector<LightCell> cells(C);
		uint32_t total = 0;
		resourcesCF_->aabbBuffer_.DownloadBufferData(vulkanContext_, aabbs.data(), C * sizeof(AABB));
		resourcesCF_->lightCellsBuffer_.DownloadBufferData(vulkanContext_, cells.data(), C * sizeof(LightCell));
		resourcesCF_->globalIndexCountBuffers_[lastFrameIndex_].DownloadBufferData(vulkanContext_, &total, sizeof(total));
		std::vector<uint32_t> indices(total ? total : 1);
		resourcesCF_->lightIndicesBuffer_.DownloadBufferData(vulkanContext_, indices.data(), total * sizeof(uint32_t));
		const ClusterForwardUBO ubo = camera_->GetClusterForwardUBO();
		Write(prefix + ".ubo", &ubo, sizeof(ubo));
		Write(prefix + ".aabb", aabbs.data(), C * sizeof(AABB));
		Write(prefix + ".cells", cells.data(), C * sizeof(LightCell));
		Write(prefix + ".indices", indices.data(), total * sizeof(uint32_t));
		Write(prefix + ".lights", resourcesLight_->lights_.data(), resourcesLight_->lights_.size() * sizeof(LightData));
		std::printf("clusters %zu lights %u total_indices %u\n", C, resourcesLight_->GetLightCount(), total);
	}

	// AppBase::DestroyResources order: resources first, then pipelines (AppBase.cpp:298-299)
	~AppPBRClusterForward()
	{
		resourcesLight_.reset();
		resourcesCF_.reset();
		pipelines_.clear();
	}

private:
	static void Write(const std::string& path, const void* data, size_t bytes)
	{
		FILE* f = std::fopen(path.c_str(), "wb");
		if (!f)
			throw std::runtime_error("cannot open " + path);
		std::fwrite(data, 1, bytes, f);
		std::fclose(f);
	}

	VulkanContext vulkanContext_;
	std::unique_ptr<Camera> camera_;
	std::unique_ptr<ResourcesLight> resourcesLight_;
	std::unique_ptr<ResourcesClusterForward> resourcesCF_;
	std::vector<std::unique_ptr<PipelineBase>> pipelines_;
	PipelineAABBGenerator* aabbPtr_ = nullptr;
	PipelineLightCulling* lightCullPtr_ = nullptr;
	uint32_t X_, Y_, Z_;
	float width_, height_;
	uint32_t lightCount_, seed_;
	uint32_t lastFrameIndex_ = 0;
};

int main(int argc, char** argv)
{
	if (argc != 10)
	{
		std::fprintf(stderr, "usage: %s X Y Z width height lights seed frames out_prefix\n", argv[0]);
		return 2;
	}
	try
	{
		AppPBRClusterForward app(std::atoi(argv[1]), std::atoi(argv[2]), std::atoi(argv[3]), static_cast<float>(std::atof(argv[4])),
			static_cast<float>(std::atof(argv[5])), static_cast<uint32_t>(std::atoll(argv[6])), static_cast<uint32_t>(std::atoll(argv[7])));
		app.MainLoop(static_cast<uint32_t>(std::atoi(argv[8])));
		app.Dump(argv[9]);
	}
	catch (const std::exception& e)
	{
		std::fprintf(stderr, "fatal: %s\n", e.what());
		return 1;
	}
	return 0;
}
