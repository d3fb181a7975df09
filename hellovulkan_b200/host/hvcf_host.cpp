// hvcf_host.cpp -- host-side (CPU) producers of the path's inputs, exported through the C ABI:
//   hvcf_host_camera_ubo   Camera -> ClusterForwardUBO (include/hv/Camera.h)
//   hvcf_host_frustum_ubo  Camera::GetFrustumUBO (Source/Camera.cpp:199-234)
//   hvcf_host_init_gbuffer / hvcf_host_init_boxes   synthetic inputs of the consumer and of frustum culling
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
// ---- synthetic inputs of the rows next to the path (no reference counterpart: the reference rasterises meshes) ----
static inline uint32_t hash32(uint32_t x)
{
	x ^= x >> 16, x *= 0x7feb352du, x ^= x >> 15, x *= 0x846ca68bu, x ^= x >> 16;
	return x;
}
static inline float hash01(uint32_t a, uint32_t b) { return static_cast<float>(hash32(a * 0x9e3779b9u + b) >> 8) * (1.0f / 16777216.0f); }

// Rays through the pixel centres of `camera` against the ground plane y = 0 and the inside of the cylinder
// x^2 + z^2 = 12^2 (the light disc of InitLights has radius 10); gl_FragCoord.z from the projection (depth 0..1).
void InitGBuffer(uint32_t seed, const Camera& camera, uint32_t width, uint32_t height, float* depth, float* worldPos,
	float* normalRough, float* albedoMetal)
{
	const mat4 invProj = camera.GetInverseProjectionMatrix(), view = camera.GetViewMatrix(), proj = camera.GetProjectionMatrix();
	const vec3 eye = camera.Position();
	const double R = 12.0;
	for (uint32_t y = 0; y < height; ++y)
		for (uint32_t x = 0; x < width; ++x)
		{
			const size_t p = static_cast<size_t>(y) * width + x;
			// clip -> view ray (point on the near plane), view -> world by the transpose of the rotation
			const double cx = (x + 0.5) / width * 2.0 - 1.0, cy = (y + 0.5) / height * 2.0 - 1.0;
			double v[4];
			for (int k = 0; k < 4; ++k)
				v[k] = invProj[0][k] * cx + invProj[1][k] * cy + invProj[3][k];
			const double vx = v[0] / v[3], vy = v[1] / v[3], vz = v[2] / v[3];
			double d[3];
			for (int k = 0; k < 3; ++k)
				d[k] = view[k][0] * vx + view[k][1] * vy + view[k][2] * vz;
			const double dl = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
			d[0] /= dl, d[1] /= dl, d[2] /= dl;
			double t = 1e30, n[3] = { 0, 1, 0 };
			if (d[1] < -1e-9)
				t = -eye.y / d[1];
			{
				const double a = d[0] * d[0] + d[2] * d[2], b = 2.0 * (eye.x * d[0] + eye.z * d[2]), c = eye.x * eye.x + eye.z * eye.z - R * R;
				const double disc = b * b - 4.0 * a * c;
				if (a > 1e-12 && disc > 0.0)
				{
					const double tc = (-b + std::sqrt(disc)) / (2.0 * a);
					if (tc > 0.0 && tc < t)
					{
						t = tc;
						n[0] = -(eye.x + t * d[0]) / R, n[1] = 0.0, n[2] = -(eye.z + t * d[2]) / R;
					}
				}
			}
			if (!(t < 1e29))
				t = 90.0; // nothing hit (cannot happen inside the cylinder): park the fragment far away
			const double w[3] = { eye.x + t * d[0], eye.y + t * d[1], eye.z + t * d[2] };
			// view-space depth -> gl_FragCoord.z with the reference projection (RH, depth 0..1)
			const double zv = view[0][2] * w[0] + view[1][2] * w[1] + view[2][2] * w[2] + view[3][2];
			double zn = (proj[2][2] * zv + proj[3][2]) / (-zv);
			zn = std::fmin(std::fmax(zn, 0.0), 0.99999);
			depth[p] = static_cast<float>(zn);
			worldPos[4 * p] = static_cast<float>(w[0]), worldPos[4 * p + 1] = static_cast<float>(w[1]);
			worldPos[4 * p + 2] = static_cast<float>(w[2]), worldPos[4 * p + 3] = 1.0f;
			// materials and a bumped normal, hashed per 4x4 pixel patch so that neighbours are coherent
			const uint32_t key = (y >> 2) * 0x1f1f1f1fu + (x >> 2) + seed * 0x68bc21ebu;
			double bn[3] = { n[0] + 0.25 * (hash01(key, 1) - 0.5), n[1] + 0.25 * (hash01(key, 2) - 0.5), n[2] + 0.25 * (hash01(key, 3) - 0.5) };
			const double bl = std::sqrt(bn[0] * bn[0] + bn[1] * bn[1] + bn[2] * bn[2]);
			normalRough[4 * p] = static_cast<float>(bn[0] / bl), normalRough[4 * p + 1] = static_cast<float>(bn[1] / bl);
			normalRough[4 * p + 2] = static_cast<float>(bn[2] / bl);
			normalRough[4 * p + 3] = 0.05f + 0.95f * hash01(key, 4); // roughness
			albedoMetal[4 * p] = 0.2f + 0.7f * hash01(key, 5), albedoMetal[4 * p + 1] = 0.2f + 0.7f * hash01(key, 6);
			albedoMetal[4 * p + 2] = 0.2f + 0.7f * hash01(key, 7);
			albedoMetal[4 * p + 3] = hash01(key, 8) < 0.3f ? hash01(key, 9) : 0.0f; // metallic
		}
}

void InitBoxes(uint32_t seed, uint32_t count, float extent, AABB* out)
{
	SeededRandom RandomNumber(seed);
	for (uint32_t i = 0; i < count; ++i)
	{
		const float cx = RandomNumber(-extent, extent), cy = RandomNumber(-extent, extent), cz = RandomNumber(-extent, extent);
		const float hx = RandomNumber(0.05f, 1.5f), hy = RandomNumber(0.05f, 1.5f), hz = RandomNumber(0.05f, 1.5f);
		out[i].minPoint = { cx - hx, cy - hy, cz - hz, 1.0f };
		out[i].maxPoint = { cx + hx, cy + hy, cz + hz, 1.0f };
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

int hvcf_host_frustum_ubo(const float position[3], const float target[3], float screen_width, float screen_height, hvcf_frustum_ubo* out)
{
	if (!position || !target || !out || !(screen_width > 0.0f) || !(screen_height > 0.0f))
		return HVCF_ERR_INVALID_ARGUMENT;
	Camera camera({ 0.0f, 0.0f, 3.0f });
	camera.SetScreenSize(screen_width, screen_height);
	camera.SetPositionAndTarget({ position[0], position[1], position[2] }, { target[0], target[1], target[2] });
	const FrustumUBO ubo = camera.GetFrustumUBO();
	std::memcpy(out, &ubo, sizeof(ubo));
	return HVCF_OK;
}

int hvcf_host_init_gbuffer(uint32_t seed, const float position[3], const float target[3], uint32_t width, uint32_t height,
	float* frag_depth, float* world_pos, float* normal_rough, float* albedo_metal)
{
	if (!position || !target || !width || !height || !frag_depth || !world_pos || !normal_rough || !albedo_metal)
		return HVCF_ERR_INVALID_ARGUMENT;
	Camera camera({ 0.0f, 0.0f, 3.0f });
	camera.SetScreenSize(static_cast<float>(width), static_cast<float>(height));
	camera.SetPositionAndTarget({ position[0], position[1], position[2] }, { target[0], target[1], target[2] });
	hv::InitGBuffer(seed, camera, width, height, frag_depth, world_pos, normal_rough, albedo_metal);
	return HVCF_OK;
}

int hvcf_host_init_boxes(uint32_t seed, uint32_t count, float extent, hvcf_aabb* out)
{
	if ((count && !out) || !(extent > 0.0f))
		return HVCF_ERR_INVALID_ARGUMENT;
	hv::InitBoxes(seed, count, extent, reinterpret_cast<AABB*>(out));
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
