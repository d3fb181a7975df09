// hv/Camera.h -- headless restatement of the reference camera's ClusterForwardUBO producer
// (HelloVulkan/Header/Camera.h, Source/Camera.cpp:32-39,81-111,113-139,178-234).  Host CPU code, as in
// the reference.  Same method names; window/input handling is left out.  The arithmetic follows
// glm 0.9.9.9 operation by operation (perspectiveRH_ZO, inverse, lookAtRH) so the UBO bytes are the
// ones the reference feeds its shaders; build with -ffp-contract=off.
#ifndef HV_CAMERA_H
#define HV_CAMERA_H

#include <cmath>

#include "Types.h"

namespace hv
{
namespace detail
{
	inline vec3 sub(vec3 a, vec3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
	inline vec3 add(vec3 a, vec3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
	inline float dot(vec3 a, vec3 b)
	{
		const vec3 t = { a.x * b.x, a.y * b.y, a.z * b.z };
		return t.x + t.y + t.z;
	}
	inline vec3 cross(vec3 a, vec3 b) { return { a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y }; }
	inline vec3 normalize(vec3 v)
	{
		const float inv = 1.0f / std::sqrt(dot(v, v));
		return { v.x * inv, v.y * inv, v.z * inv };
	}
	inline float clampf(float x, float lo, float hi) { return std::fmin(std::fmax(x, lo), hi); }

	// glm::perspective under GLM_FORCE_DEPTH_ZERO_TO_ONE (right-handed, depth 0..1)
	inline mat4 perspective(float fovy, float aspect, float zNear, float zFar)
	{
		const float t = std::tan(fovy / 2.0f);
		mat4 r{};
		r[0][0] = 1.0f / (aspect * t);
		r[1][1] = 1.0f / t;
		r[2][2] = zFar / (zNear - zFar);
		r[2][3] = -1.0f;
		r[3][2] = -(zFar * zNear) / (zFar - zNear);
		return r;
	}

	// glm::inverse(mat4): cofactors from 2x2 minors, then one reciprocal of the determinant
	inline mat4 inverse(const mat4& m)
	{
		// minor(a, b, c, d) = m[a][b] * m[c][d] - m[c][b] * m[a][d]
		auto minor = [&m](int a, int b, int c, int d) { return m[a][b] * m[c][d] - m[c][b] * m[a][d]; };
		// the six Fac vectors of glm, each {k, k, l, n} built from the column triple (2,3),(1,3),(1,2)
		auto fac = [&minor](int r0, int r1) -> vec4 {
			const float a = minor(2, r0, 3, r1), b = minor(1, r0, 3, r1), c = minor(1, r0, 2, r1);
			return { a, a, b, c };
		};
		const vec4 F0 = fac(2, 3), F1 = fac(1, 3), F2 = fac(1, 2);
		// Fac3..5 use column pairs over rows (0,3), (0,2), (0,1)
		const vec4 F3 = fac(0, 3), F4 = fac(0, 2), F5 = fac(0, 1);
		auto col = [&m](int row) -> vec4 { return { m[1][row], m[0][row], m[0][row], m[0][row] }; };
		const vec4 V0 = col(0), V1 = col(1), V2 = col(2), V3 = col(3);
		const float sa[4] = { +1.0f, -1.0f, +1.0f, -1.0f };
		const float sb[4] = { -1.0f, +1.0f, -1.0f, +1.0f };
		mat4 inv{};
		for (int i = 0; i < 4; ++i)
		{
			inv[0][i] = ((V1[i] * F0[i] - V2[i] * F1[i]) + V3[i] * F2[i]) * sa[i];
			inv[1][i] = ((V0[i] * F0[i] - V2[i] * F3[i]) + V3[i] * F4[i]) * sb[i];
			inv[2][i] = ((V0[i] * F1[i] - V1[i] * F3[i]) + V3[i] * F5[i]) * sa[i];
			inv[3][i] = ((V0[i] * F2[i] - V1[i] * F4[i]) + V2[i] * F5[i]) * sb[i];
		}
		const float d0 = m[0][0] * inv[0][0], d1 = m[0][1] * inv[1][0], d2 = m[0][2] * inv[2][0], d3 = m[0][3] * inv[3][0];
		const float oneOverDet = 1.0f / ((d0 + d1) + (d2 + d3));
		for (int c = 0; c < 4; ++c)
			for (int r = 0; r < 4; ++r)
				inv[c][r] = inv[c][r] * oneOverDet;
		return inv;
	}

	// glm mat4 * mat4: Result[c] = ((A0*b0 + A1*b1) + A2*b2) + A3*b3 (left to right)
	inline mat4 mul(const mat4& a, const mat4& b)
	{
		mat4 r{};
		for (int c = 0; c < 4; ++c)
			for (int k = 0; k < 4; ++k)
				r[c][k] = ((a[0][k] * b[c][0] + a[1][k] * b[c][1]) + a[2][k] * b[c][2]) + a[3][k] * b[c][3];
		return r;
	}

	// glm mat4 * vec4: (m0*x + m1*y) + (m2*z + m3*w)
	inline vec4 mul(const mat4& m, const vec4& v)
	{
		vec4 r{};
		for (int k = 0; k < 4; ++k)
			r[k] = (m[0][k] * v.x + m[1][k] * v.y) + (m[2][k] * v.z + m[3][k] * v.w);
		return r;
	}

	// glm::lookAt (right-handed)
	inline mat4 lookAt(vec3 eye, vec3 center, vec3 up)
	{
		const vec3 f = normalize(sub(center, eye));
		const vec3 s = normalize(cross(f, up));
		const vec3 u = cross(s, f);
		mat4 r{};
		r[3][3] = 1.0f;
		r[0][0] = s.x, r[1][0] = s.y, r[2][0] = s.z;
		r[0][1] = u.x, r[1][1] = u.y, r[2][1] = u.z;
		r[0][2] = -f.x, r[1][2] = -f.y, r[2][2] = -f.z;
		r[3][0] = -dot(s, eye);
		r[3][1] = -dot(u, eye);
		r[3][2] = dot(f, eye);
		return r;
	}
} // namespace detail
} // namespace hv

class Camera
{
public:
	explicit Camera(hv::vec3 position, hv::vec3 worldUp = { 0.0f, 1.0f, 0.0f },
		float screenWidth = static_cast<float>(AppConfig::InitialScreenWidth),
		float screenHeight = static_cast<float>(AppConfig::InitialScreenHeight), float yaw = CameraConfig::Yaw,
		float pitch = CameraConfig::Pitch)
		: position_(position), worldUp_(worldUp), yaw_(yaw), pitch_(pitch), zoom_(CameraConfig::Zoom),
		  screenWidth_(screenWidth), screenHeight_(screenHeight)
	{
		UpdateInternal();
	}

	// run-time grid (the reference bakes ClusterForwardConfig into GetClusterForwardUBO)
	void SetSliceCounts(uint32_t x, uint32_t y, uint32_t z) { sliceX_ = x, sliceY_ = y, sliceZ_ = z; }

	void SetPositionAndTarget(hv::vec3 cameraPosition, hv::vec3 cameraTarget)
	{
		position_ = cameraPosition;
		const hv::vec3 direction = hv::detail::normalize(hv::detail::sub(cameraTarget, position_));
		// Camera.cpp:86-87 calls the C library's double asin/atan2 and narrows
		pitch_ = static_cast<float>(std::asin(static_cast<double>(direction.y)));
		yaw_ = static_cast<float>(std::atan2(static_cast<double>(direction.z), static_cast<double>(direction.x)));
		ConstrainPitch();
		UpdateInternal();
	}

	void SetScreenSize(float width, float height)
	{
		screenWidth_ = width;
		screenHeight_ = height;
		UpdateInternal();
	}

	void ProcessMouseScroll(float yOffset)
	{
		zoom_ -= yOffset;
		zoom_ = hv::detail::clampf(zoom_, 1.0f, CameraConfig::Zoom);
		UpdateInternal();
	}

	hv::mat4 GetProjectionMatrix() const { return projectionMatrix_; }
	hv::mat4 GetInverseProjectionMatrix() const { return inverseProjectionMatrix_; }
	hv::mat4 GetViewMatrix() const { return viewMatrix_; }
	hv::vec3 Position() const { return position_; }

	CameraUBO GetCameraUBO() const
	{
		CameraUBO u{};
		u.projection = projectionMatrix_;
		u.view = viewMatrix_;
		u.position = { position_.x, position_.y, position_.z, 1.0f };
		u.cameraNear = CameraConfig::Near;
		u.cameraFar = CameraConfig::Far;
		return u;
	}

	ClusterForwardUBO GetClusterForwardUBO() const
	{
		const float zFloat = static_cast<float>(sliceZ_);
		const float log2FarDivNear = std::log2(CameraConfig::Far / CameraConfig::Near);
		const float log2Near = std::log2(CameraConfig::Near);
		ClusterForwardUBO u{};
		u.cameraInverseProjection = inverseProjectionMatrix_;
		u.cameraView = viewMatrix_;
		u.screenSize = { screenWidth_, screenHeight_ };
		u.sliceScaling = zFloat / log2FarDivNear;
		u.sliceBias = -(zFloat * log2Near / log2FarDivNear);
		u.cameraNear = CameraConfig::Near;
		u.cameraFar = CameraConfig::Far;
		u.sliceCountX = sliceX_;
		u.sliceCountY = sliceY_;
		u.sliceCountZ = sliceZ_;
		return u;
	}

	// Camera.cpp:199-234
	FrustumUBO GetFrustumUBO() const
	{
		const hv::mat4 projView = hv::detail::mul(projectionMatrix_, viewMatrix_);
		hv::mat4 t{}; // glm::transpose
		for (int c = 0; c < 4; ++c)
			for (int r = 0; r < 4; ++r)
				t[c][r] = projView[r][c];
		auto add4 = [](const hv::vec4& a, const hv::vec4& b) -> hv::vec4 { return { a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w }; };
		auto sub4 = [](const hv::vec4& a, const hv::vec4& b) -> hv::vec4 { return { a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w }; };
		FrustumUBO ubo{};
		ubo.planes[0] = add4(t[3], t[0]); // left
		ubo.planes[1] = sub4(t[3], t[0]); // right
		ubo.planes[2] = add4(t[3], t[1]); // bottom
		ubo.planes[3] = sub4(t[3], t[1]); // top
		ubo.planes[4] = add4(t[3], t[2]); // near
		ubo.planes[5] = sub4(t[3], t[2]); // far
		const hv::mat4 invProjView = hv::detail::inverse(projView);
		const hv::vec4 corners[8] = { { -1, -1, -1, 1 }, { 1, -1, -1, 1 }, { 1, 1, -1, 1 }, { -1, 1, -1, 1 },
			{ -1, -1, 1, 1 }, { 1, -1, 1, 1 }, { 1, 1, 1, 1 }, { -1, 1, 1, 1 } };
		for (int i = 0; i < 8; ++i)
		{
			const hv::vec4 q = hv::detail::mul(invProjView, corners[i]);
			ubo.corners[i] = { q.x / q.w, q.y / q.w, q.z / q.w, q.w / q.w };
		}
		return ubo;
	}

private:
	void ConstrainPitch()
	{
		constexpr float pitchLimit = 1.55334f;
		pitch_ = hv::detail::clampf(pitch_, -pitchLimit, pitchLimit);
	}

	void UpdateInternal()
	{
		// Camera.cpp:116-119: double cos/sin, product narrowed to float
		const double cy = std::cos(static_cast<double>(yaw_)), sy = std::sin(static_cast<double>(yaw_));
		const double cp = std::cos(static_cast<double>(pitch_)), sp = std::sin(static_cast<double>(pitch_));
		hv::vec3 front = { static_cast<float>(cy * cp), static_cast<float>(sp), static_cast<float>(sy * cp) };
		front_ = hv::detail::normalize(front);
		right_ = hv::detail::normalize(hv::detail::cross(front_, worldUp_));
		up_ = hv::detail::normalize(hv::detail::cross(right_, front_));
		projectionMatrix_ = hv::detail::perspective(zoom_, screenWidth_ / screenHeight_, CameraConfig::Near, CameraConfig::Far);
		projectionMatrix_[1][1] *= -1;
		inverseProjectionMatrix_ = hv::detail::inverse(projectionMatrix_);
		viewMatrix_ = hv::detail::lookAt(position_, hv::detail::add(position_, front_), up_);
	}

	hv::mat4 projectionMatrix_{}, inverseProjectionMatrix_{}, viewMatrix_{};
	hv::vec3 position_{}, worldUp_{}, front_{}, up_{}, right_{};
	float yaw_ = 0, pitch_ = 0, zoom_ = 0, screenWidth_ = 0, screenHeight_ = 0;
	uint32_t sliceX_ = ClusterForwardConfig::SliceCountX, sliceY_ = ClusterForwardConfig::SliceCountY,
			 sliceZ_ = ClusterForwardConfig::SliceCountZ;
};

#endif
