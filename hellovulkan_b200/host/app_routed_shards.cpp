// app_routed_shards.cpp -- the routed sharded frame (hvcf_cull_routed, include/hvcf.h) driven from C++ alone: no Python, no
// collective library, nothing but the C ABI plus one CUDA stream per shard.  `world` shard contexts live in THIS process on the
// current device (hvcf_peer_attach_local); each owns a slab of cluster rows and a contiguous range of the lights, as the ranks
// of a multi-GPU job do.  After every frame EVERY shard must hold the full light grid, byte-identical to the unsharded frame
// (hvcf_cull; PipelineLightCulling::Execute, Source/Pipelines/PipelineLightCulling.cpp:58-119) over all the lights.
// The scene is the one of AppPBRClusterForward (AppPBRClusterForward.cpp:81-118 InitLights, :120-136 UpdateUBOs); frame f > 0
// orbits the camera and moves light 0, as app_cluster_forward.cpp does.
//
// usage: hv_routed_shards_app <X> <Y> <Z> <width> <height> <lights> <seed> <world> <frames>
// prints "OK ..." and exits 0 when every shard of every frame agrees with the unsharded frame.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/hv/Camera.h"
#include "../../include/hv/Context.h"

namespace
{
struct Grid
{
	std::vector<hvcf_light_cell> cells;
	std::vector<uint32_t> indices;
	uint32_t total = 0;
};

void Cuda(cudaError_t e, const char* what)
{
	if (e != cudaSuccess)
		throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}

Grid Download(hvcf_ctx* ctx, uint32_t frame, size_t clusters, uint32_t cap, void* stream)
{
	Grid g;
	g.cells.resize(clusters);
	g.indices.resize(clusters * cap);
	hv::Check(hvcf_download_grid(ctx, frame, g.cells.data(), clusters * sizeof(hvcf_light_cell), g.indices.data(),
		g.indices.size() * sizeof(uint32_t), &g.total, stream), "hvcf_download_grid");
	g.indices.resize(g.total);
	return g;
}

bool Same(const Grid& a, const Grid& b)
{
	return a.total == b.total && std::memcmp(a.cells.data(), b.cells.data(), a.cells.size() * sizeof(hvcf_light_cell)) == 0 &&
		std::memcmp(a.indices.data(), b.indices.data(), size_t(a.total) * sizeof(uint32_t)) == 0;
}
} // namespace

int main(int argc, char** argv)
{
	if (argc != 10)
	{
		std::fprintf(stderr, "usage: %s X Y Z width height lights seed world frames\n", argv[0]);
		return 2;
	}
	const uint32_t X = std::atoi(argv[1]), Y = std::atoi(argv[2]), Z = std::atoi(argv[3]);
	const float width = static_cast<float>(std::atof(argv[4])), height = static_cast<float>(std::atof(argv[5]));
	const uint32_t n = static_cast<uint32_t>(std::atoll(argv[6])), seed = static_cast<uint32_t>(std::atoll(argv[7]));
	const uint32_t world = std::atoi(argv[8]), frames = std::atoi(argv[9]);
	if (world < 1 || world > HVCF_MAX_PEERS || world > Z || frames < 1)
	{
		std::fprintf(stderr, "world must be 1..min(%d, Z), frames >= 1\n", HVCF_MAX_PEERS);
		return 2;
	}
	std::vector<hvcf_ctx*> shards(world, nullptr);
	std::vector<cudaStream_t> streams(world, nullptr);
	hvcf_ctx* whole = nullptr;
	int rc = 0;
	try
	{
		const size_t clusters = size_t(X) * Y * Z;
		const uint32_t cap = ClusterForwardConfig::maxLightPerCluster;
		std::vector<LightData> lights(n ? n : 1);
		hv::Check(hvcf_host_init_lights(seed, n, reinterpret_cast<hvcf_light_data*>(lights.data())), "hvcf_host_init_lights");
		Camera camera(hv::vec3{ 0.0f, 0.0f, 3.0f });
		camera.SetSliceCounts(X, Y, Z);
		camera.SetScreenSize(width, height);

		hvcf_config cfg{};
		cfg.struct_size = sizeof(cfg);
		cfg.device = -1;
		cfg.slice_count_x = X, cfg.slice_count_y = Y, cfg.slice_count_z = Z;
		cfg.max_lights_per_cluster = cap;
		cfg.max_lights = n ? n : 1;
		cfg.frame_count = AppConfig::FrameCount;
		hv::Check(hvcf_create(&cfg, &whole), "hvcf_create");
		hv::Check(hvcf_set_lights(whole, lights.data(), n, 0, nullptr), "hvcf_set_lights");

		// uneven slabs on purpose: the near slices hold most of the lights' hits
		hvcf_shard_desc desc{};
		desc.struct_size = sizeof(desc);
		desc.world = world;
		desc.total_lights = n;
		for (uint32_t r = 0; r <= world; ++r)
		{
			uint32_t cut = r == world ? Z : (r * Z * 3u) / (world * 4u) + (r ? 1u : 0u);
			if (r && cut <= desc.z_cuts[r - 1])
				cut = desc.z_cuts[r - 1] + 1;
			if (r < world && cut > Z - (world - r)) // leave a slice for every shard behind this one
				cut = Z - (world - r);
			desc.z_cuts[r] = cut;
		}
		uint32_t covered = 0;
		for (uint32_t r = 0; r < world; ++r)
		{
			desc.rank = r;
			hv::Check(hvcf_create_sharded(&cfg, &desc, &shards[r]), "hvcf_create_sharded");
			Cuda(cudaStreamCreateWithFlags(&streams[r], cudaStreamNonBlocking), "cudaStreamCreateWithFlags");
			uint32_t b = 0, e = 0;
			hv::Check(hvcf_shard_light_range(shards[r], &b, &e), "hvcf_shard_light_range");
			if (b != (covered < n ? covered : n) || e < b || e > n)
				throw std::runtime_error("light ranges of the shards do not tile the light buffer");
			covered = e > covered ? e : covered;
			// every shard holds ONLY the lights it owns
			hv::Check(hvcf_set_lights_range(shards[r], lights.data() + b, b, e - b, 0, streams[r]), "hvcf_set_lights_range");
		}
		if (covered != n)
			throw std::runtime_error("light ranges of the shards do not cover the light buffer");
		Cuda(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
		if (world > 1)
			for (uint32_t r = 0; r < world; ++r)
				hv::Check(hvcf_peer_attach_local(shards[r], r, world, shards.data()), "hvcf_peer_attach_local");

		unsigned long long indices_seen = 0, inbox_seen = 0;
		for (uint32_t f = 0; f < frames; ++f)
		{
			const uint32_t slot = f % AppConfig::FrameCount;
			if (f > 0 && n > 0)
			{
				// ResourcesLight::UpdateLightPosition (ResourcesLight.cpp:20-29) on light 0: on the unsharded context and on its owner
				const float xyz[3] = { 0.25f * static_cast<float>(f), 2.0f, 1.0f };
				hv::Check(hvcf_update_light(whole, 0, xyz, nullptr), "hvcf_update_light");
				lights[0].position_ = { xyz[0], xyz[1], xyz[2], 1.0f };
				hv::Check(hvcf_set_lights_range(shards[0], lights.data(), 0, 1, 0, streams[0]), "hvcf_set_lights_range");
			}
			camera.SetPositionAndTarget({ 0.5f * static_cast<float>(f), 1.0f, 6.0f }, { 0.0f, 2.5f, 0.0f });
			const ClusterForwardUBO ubo = camera.GetClusterForwardUBO();
			const auto* cubo = reinterpret_cast<const hvcf_cluster_forward_ubo*>(&ubo);
			hv::Check(hvcf_set_ubo(whole, slot, cubo, nullptr), "hvcf_set_ubo");
			hv::Check(hvcf_cull(whole, slot, nullptr), "hvcf_cull");
			const Grid want = Download(whole, slot, clusters, cap, nullptr);

			for (uint32_t r = 0; r < world; ++r)
				hv::Check(hvcf_set_ubo(shards[r], slot, cubo, streams[r]), "hvcf_set_ubo");
			// one call per shard and frame; the waits inside overlap the other shards' work because every shard has its own stream
			for (uint32_t r = 0; r < world; ++r)
				hv::Check(hvcf_cull_routed(shards[r], slot, streams[r]), "hvcf_cull_routed");
			for (uint32_t r = 0; r < world; ++r)
			{
				int err = 0;
				hv::Check(hvcf_peer_status(shards[r], streams[r], &err), "hvcf_peer_status");
				if (err)
					throw std::runtime_error("peer wait timed out on shard " + std::to_string(r));
			}
			for (uint32_t r = 0; r < world; ++r)
			{
				const Grid got = Download(shards[r], slot, clusters, cap, streams[r]);
				if (!Same(got, want))
				{
					std::fprintf(stderr, "frame %u shard %u: light grid differs from the unsharded frame (total %u vs %u)\n", f, r, got.total,
						want.total);
					rc = 1;
				}
				uint64_t records = 0;
				hv::Check(hvcf_shard_inbox_records(shards[r], streams[r], &records), "hvcf_shard_inbox_records");
				inbox_seen += records;
			}
			indices_seen += want.total;
		}
		if (world > 1)
			for (uint32_t r = 0; r < world; ++r)
				hv::Check(hvcf_peer_detach(shards[r]), "hvcf_peer_detach");
		if (rc == 0)
			std::printf("OK grid %ux%ux%u lights %u world %u frames %u indices %llu inbox_records %llu\n", X, Y, Z, n, world, frames,
				indices_seen, inbox_seen);
	}
	catch (const std::exception& e)
	{
		std::fprintf(stderr, "fatal: %s\n", e.what());
		rc = 1;
	}
	for (uint32_t r = 0; r < world; ++r)
	{
		if (shards[r])
			hvcf_destroy(shards[r]);
		if (streams[r])
			cudaStreamDestroy(streams[r]);
	}
	if (whole)
		hvcf_destroy(whole);
	return rc;
}
