// app_cluster_forward.cpp -- headless driver written in the shape of AppPBRClusterForward
// (HelloVulkan/Source/Apps/AppPBRClusterForward.cpp:19-79 Init, :81-118 InitLights, :120-136 UpdateUBOs,
// :160-182 MainLoop; AppBase::DrawFrame / FillCommandBuffer, Source/Apps/AppBase.cpp:94-229), reduced to the
// compute pipelines of the light-culling path plus its consumer (the PBR cluster-forward light loop over a synthetic
// G-buffer) and, as in AppPBRBindless-style apps, compute frustum culling ahead of them (AppFrustumCulling.cpp).
// No window, swapchain or UI: frames are a counted loop and the results are written to files so a test can compare
// them with the oracle.
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
		InitScene();
		// pipeline order of AppPBRClusterForward::Init (AppPBRClusterForward.cpp:46-78), compute + consumer only
		auto frustum = std::make_unique<PipelineFrustumCulling>(vulkanContext_, &scene_);
		auto aabb = std::make_unique<PipelineAABBGenerator>(vulkanContext_, resourcesCF_.get());
		auto cull = std::make_unique<PipelineLightCulling>(vulkanContext_, resourcesLight_.get(), resourcesCF_.get());
		auto pbr = std::make_unique<PipelinePBRClusterForward>(vulkanContext_, resourcesGBuffer_.get(), resourcesLight_.get(), resourcesCF_.get());
		frustumPtr_ = frustum.get();
		aabbPtr_ = aabb.get();
		lightCullPtr_ = cull.get();
		pbrPtr_ = pbr.get();
		pipelines_.push_back(std::move(frustum));
		pipelines_.push_back(std::move(aabb));
		pipelines_.push_back(std::move(cull));
		pipelines_.push_back(std::move(pbr));
	}

	// synthetic stand-ins of what Scene / rasterisation provide in the reference
	void InitScene()
	{
		const uint32_t w = static_cast<uint32_t>(width_), h = static_cast<uint32_t>(height_);
		const size_t P = static_cast<size_t>(w) * h;
		std::vector<float> depth(P), wp(4 * P), nr(4 * P), am(4 * P);
		const float pos[3] = { 0.0f, 1.0f, 6.0f }, target[3] = { 0.0f, 2.5f, 0.0f };
		hv::Check(hvcf_host_init_gbuffer(7, pos, target, w, h, depth.data(), wp.data(), nr.data(), am.data()), "hvcf_host_init_gbuffer");
		resourcesGBuffer_ = std::make_unique<ResourcesGBufferCF>();
		resourcesGBuffer_->Create(vulkanContext_, w, h, depth.data(), wp.data(), nr.data(), am.data());
		std::vector<AABB> boxes(kBoxes);
		hv::Check(hvcf_host_init_boxes(4242, kBoxes, 60.0f, reinterpret_cast<hvcf_aabb*>(boxes.data())), "hvcf_host_init_boxes");
		std::vector<DrawIndirectCommand> commands(kBoxes);
		for (uint32_t i = 0; i < kBoxes; ++i)
			commands[i] = { 36u, 1u, 0u, i };
		scene_.Create(vulkanContext_, boxes, commands);
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
		pbrPtr_->SetClusterForwardUBO(vulkanContext_, cfUBO);
		FrustumUBO frustumUBO = camera_->GetFrustumUBO(); // AppFrustumCulling::UpdateUBOs
		frustumPtr_->SetFrustumUBO(vulkanContext_, frustumUBO);
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
				// AppBase hands the UI state to every resource each frame; ResourcesLight moves light 0 to the shadow-caster position
				uiData_.shadowCasterPosition_ = { 0.25f * static_cast<float>(f), 2.0f, 1.0f };
				static_cast<ResourcesBase*>(resourcesLight_.get())->UpdateFromUIData(vulkanContext_, uiData_);
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
		const size_t P = static_cast<size_t>(resourcesGBuffer_->width_) * resourcesGBuffer_->height_;
		std::vector<float> color(4 * P);
		std::vector<uint32_t> cluster(P);
		resourcesGBuffer_->color_.DownloadBufferData(vulkanContext_, color.data(), 16 * P);
		resourcesGBuffer_->clusterIndex_.DownloadBufferData(vulkanContext_, cluster.data(), 4 * P);
		Write(prefix + ".color", color.data(), 16 * P);
		Write(prefix + ".cluster", cluster.data(), 4 * P);
		std::vector<DrawIndirectCommand> commands(kBoxes);
		scene_.indirectBuffer_.DownloadBufferData(vulkanContext_, commands.data(), sizeof(DrawIndirectCommand) * kBoxes);
		Write(prefix + ".commands", commands.data(), sizeof(DrawIndirectCommand) * kBoxes);
		const hv::vec3 cp = camera_->Position();
		const float camera[3] = { cp.x, cp.y, cp.z };
		Write(prefix + ".camera", camera, sizeof(camera));
		const FrustumUBO frustum = camera_->GetFrustumUBO();
		Write(prefix + ".frustum", &frustum, sizeof(frustum));
		std::printf("clusters %zu lights %u total_indices %u\n", C, resourcesLight_->GetLightCount(), total);
	}

	// AppBase::DestroyResources order: resources first, then pipelines (AppBase.cpp:298-299)
	~AppPBRClusterForward()
	{
		resourcesLight_.reset();
		resourcesCF_.reset();
		resourcesGBuffer_.reset();
		scene_.Destroy();
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
	PipelinePBRClusterForward* pbrPtr_ = nullptr;
	PipelineFrustumCulling* frustumPtr_ = nullptr;
	std::unique_ptr<ResourcesGBufferCF> resourcesGBuffer_;
	Scene scene_;
	static constexpr uint32_t kBoxes = 10000; // the instance count of the README's frustum-culling figure
	uint32_t X_, Y_, Z_;
	float width_, height_;
	uint32_t lightCount_, seed_;
	uint32_t lastFrameIndex_ = 0;
	UIData uiData_{};
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
