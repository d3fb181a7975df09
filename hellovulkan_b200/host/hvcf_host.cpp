// hvcf_host.cpp -- host-side (CPU) producers of the path's inputs, exported through the C ABI:
//   hvcf_host_camera_ubo   Camera -> ClusterForwardUBO (include/hv/Camera.h)
//   hvcf_host_init_lights  AppPBRClusterForward::InitLights (Source/Apps/AppPBRClusterForward.cpp:81-118)
//                          with Utility::RandomNumber (Header/Utility.h:9-20) on an explicit seed
// Compiled -ffp-contract=off.  These run on the CPU in the reference too; the device path starts at
// hvcf_set_ubo / hvcf_set_lights.
#include <cmath>
#include <cstring>
#include <random>

#include "../../include/hvcf.h"
#include "../../include/hv/Camera.h"

namespace hv
{
// Utility::RandomNumber(min, max) over one seeded engine (the reference seeds from std::random_device)
class SeededRandom
{
public:
	explicit SeededRandom(uint32_t seed) : generator_(seed) {}
	float operator()(float min, float max) { return min + (max - min) * distribution_(generator_); }

private:
	std::uniform_real_distribution<float> distribution_{ 0.0, 1.0 };
	std::mt19937 generator_;
};

void InitLights(uint32_t seed, uint32_t count, LightData* out)
{
	SeededRandom RandomNumber(seed);
	const float pi2 = 6.28318530717958647692528676655900576f; // glm::two_pi<float>()
	for (uint32_t i = 0; i < count; ++i)
	{
		const float yPos = RandomNumber(-2.f, 10.0f);
		const float radius = RandomNumber(0.0f, 10.0f);
		const float rad = RandomNumber(0.0f, pi2);
		LightData l{};
		l.position_ = { std::cos(rad) * radius, yPos, std::sin(rad) * radius, 1.f };
		const float r = RandomNumber(0.0f, 1.0f);
		const float g = RandomNumber(0.0f, 1.0f);
		const float b = RandomNumber(0.0f, 1.0f);
		l.color_ = { r, g, b, 1.f };
		l.radius_ = RandomNumber(0.5f, 2.0f);
		std::memset(static_cast<void*>(&out[i]), 0, sizeof(LightData)); // deterministic tail padding
		out[i].position_ = l.position_;
		out[i].color_ = l.color_;
		out[i].radius_ = l.radius_;
	}
}
} // namespace hv

extern "C" {

int hvcf_host_camera_ubo(const float position[3], const float target[3], float screen_width, float screen_height,
	uint32_t slice_count_x, uint32_t slice_count_y, uint32_t slice_count_z, hvcf_cluster_forward_ubo* out)
{
	if (!position || !target || !out || !(screen_width > 0.0f) || !(screen_height > 0.0f))
		return HVCF_ERR_INVALID_ARGUMENT;
	// AppBase::InitCamera places the camera at (0,0,3); AppPBRClusterForward::Init then re-targets it
	Camera camera({ 0.0f, 0.0f, 3.0f });
	camera.SetSliceCounts(slice_count_x ? slice_count_x : ClusterForwardConfig::SliceCountX,
		slice_count_y ? slice_count_y : ClusterForwardConfig::SliceCountY,
		slice_count_z ? slice_count_z : ClusterForwardConfig::SliceCountZ);
	camera.SetScreenSize(screen_width, screen_height);
	camera.SetPositionAndTarget({ position[0], position[1], position[2] }, { target[0], target[1], target[2] });
	const ClusterForwardUBO ubo = camera.GetClusterForwardUBO();
	std::memset(out, 0, sizeof(*out));
	std::memcpy(out, &ubo, offsetof(ClusterForwardUBO, sliceCountZ) + sizeof(uint32_t));
	return HVCF_OK;
}

int hvcf_host_init_lights(uint32_t seed, uint32_t count, hvcf_light_data* out)
{
	if (count && !out)
		return HVCF_ERR_INVALID_ARGUMENT;
	hv::InitLights(seed, count, reinterpret_cast<LightData*>(out));
	return HVCF_OK;
}

} // extern "C"
