/*
 * hvcf_oracle.c -- CPU oracle (plain C) for the clustered-forward light-culling path.
 * TEST INFRASTRUCTURE ONLY -- see hvcf_oracle.h for the rules and the parity pin.
 * Build: gcc -O2 -ffp-contract=off -fopenmp (oracle/Makefile).
 */
#include "hvcf_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------------------------
 * std::mt19937 (32-bit Mersenne twister, the standard's parameters) and libstdc++'s
 * uniform_real_distribution<float>(0,1) == generate_canonical<float, 24>: one 32-bit draw,
 * float(draw) / 2^32 (float arithmetic), clamped below 1.  Header/Utility.h:9-20.
 * ------------------------------------------------------------------------------------------ */
typedef struct
{
	uint32_t mt[624];
	int idx;
} hvo_mt19937;

static void mt_seed(hvo_mt19937* g, uint32_t seed)
{
	g->mt[0] = seed;
	for (int i = 1; i < 624; ++i)
		g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
	g->idx = 624;
}

static uint32_t mt_next(hvo_mt19937* g)
{
	if (g->idx >= 624)
	{
		for (int i = 0; i < 624; ++i)
		{
			uint32_t y = (g->mt[i] & 0x80000000u) | (g->mt[(i + 1) % 624] & 0x7fffffffu);
			g->mt[i] = g->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
		}
		g->idx = 0;
	}
	uint32_t y = g->mt[g->idx++];
	y ^= y >> 11;
	y ^= (y << 7) & 0x9d2c5680u;
	y ^= (y << 15) & 0xefc60000u;
	y ^= y >> 18;
	return y;
}

static float canonical_float(hvo_mt19937* g)
{
	const float sum = (float)mt_next(g); /* __sum += float(urng() - min) * 1.0f */
	float ret = sum / 4294967296.0f;     /* __tmp = 2^32 as float */
	if (ret >= 1.0f)
		ret = nextafterf(1.0f, 0.0f);
	return ret;
}

/* Utility::RandomNumber(min, max) */
static float random_number(hvo_mt19937* g, float mn, float mx)
{
	return mn + (mx - mn) * canonical_float(g);
}

void hvo_init_lights(uint32_t seed, uint32_t count, hvcf_light_data* out)
{
	hvo_mt19937 g;
	mt_seed(&g, seed);
	const float pi2 = 6.28318530717958647692528676655900576f; /* glm::two_pi<float>() */
	memset(out, 0, sizeof(hvcf_light_data) * (size_t)count);
	for (uint32_t i = 0; i < count; ++i)
	{
		const float yPos = random_number(&g, -2.f, 10.0f);
		const float radius = random_number(&g, 0.0f, 10.0f);
		const float rad = random_number(&g, 0.0f, pi2);
		out[i].position[0] = cosf(rad) * radius;
		out[i].position[1] = yPos;
		out[i].position[2] = sinf(rad) * radius;
		out[i].position[3] = 1.f;
		out[i].color[0] = random_number(&g, 0.0f, 1.0f);
		out[i].color[1] = random_number(&g, 0.0f, 1.0f);
		out[i].color[2] = random_number(&g, 0.0f, 1.0f);
		out[i].color[3] = 1.f;
		out[i].radius = random_number(&g, 0.5f, 2.0f);
	}
}

/* ------------------------------------------------------------------------------------------
 * glm restatements (External/glm/glm, v0.9.9.9), column-major m[col][row].
 * ------------------------------------------------------------------------------------------ */
typedef struct { float x, y, z; } v3;

static float dot3(v3 a, v3 b) /* detail/func_geometric.inl:48-55 */
{
	const float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
	return tx + ty + tz;
}

static v3 cross3(v3 x, v3 y) /* detail/func_geometric.inl:73-84 */
{
	v3 r;
	r.x = x.y * y.z - y.y * x.z;
	r.y = x.z * y.x - y.z * x.x;
	r.z = x.x * y.y - y.x * x.y;
	return r;
}

static v3 normalize3(v3 v) /* v * inversesqrt(dot(v,v)); inversesqrt = 1 / sqrt */
{
	const float s = 1.0f / sqrtf(dot3(v, v));
	v3 r = { v.x * s, v.y * s, v.z * s };
	return r;
}

static void mat4_inverse(const float m[4][4], float out[4][4]) /* detail/func_matrix.inl:347-405 */
{
	const float Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
	const float Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
	const float Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
	const float Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
	const float Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
	const float Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
	const float Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
	const float Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
	const float Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
	const float Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
	const float Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
	const float Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
	const float Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
	const float Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
	const float Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
	const float Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
	const float Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
	const float Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];

	const float Fac0[4] = { Coef00, Coef00, Coef02, Coef03 };
	const float Fac1[4] = { Coef04, Coef04, Coef06, Coef07 };
	const float Fac2[4] = { Coef08, Coef08, Coef10, Coef11 };
	const float Fac3[4] = { Coef12, Coef12, Coef14, Coef15 };
	const float Fac4[4] = { Coef16, Coef16, Coef18, Coef19 };
	const float Fac5[4] = { Coef20, Coef20, Coef22, Coef23 };
	const float Vec0[4] = { m[1][0], m[0][0], m[0][0], m[0][0] };
	const float Vec1[4] = { m[1][1], m[0][1], m[0][1], m[0][1] };
	const float Vec2[4] = { m[1][2], m[0][2], m[0][2], m[0][2] };
	const float Vec3[4] = { m[1][3], m[0][3], m[0][3], m[0][3] };
	static const float SignA[4] = { +1.f, -1.f, +1.f, -1.f };
	static const float SignB[4] = { -1.f, +1.f, -1.f, +1.f };

	float Inverse[4][4];
	for (int i = 0; i < 4; ++i)
	{
		const float Inv0 = (Vec1[i] * Fac0[i] - Vec2[i] * Fac1[i]) + Vec3[i] * Fac2[i];
		const float Inv1 = (Vec0[i] * Fac0[i] - Vec2[i] * Fac3[i]) + Vec3[i] * Fac4[i];
		const float Inv2 = (Vec0[i] * Fac1[i] - Vec1[i] * Fac3[i]) + Vec3[i] * Fac5[i];
		const float Inv3 = (Vec0[i] * Fac2[i] - Vec1[i] * Fac4[i]) + Vec2[i] * Fac5[i];
		Inverse[0][i] = Inv0 * SignA[i];
		Inverse[1][i] = Inv1 * SignB[i];
		Inverse[2][i] = Inv2 * SignA[i];
		Inverse[3][i] = Inv3 * SignB[i];
	}
	const float Dot0x = m[0][0] * Inverse[0][0];
	const float Dot0y = m[0][1] * Inverse[1][0];
	const float Dot0z = m[0][2] * Inverse[2][0];
	const float Dot0w = m[0][3] * Inverse[3][0];
	const float Dot1 = (Dot0x + Dot0y) + (Dot0z + Dot0w);
	const float OneOverDeterminant = 1.0f / Dot1;
	for (int c = 0; c < 4; ++c)
		for (int r = 0; r < 4; ++r)
			out[c][r] = Inverse[c][r] * OneOverDeterminant;
}

/* Header/Configs.h:33-44 */
#define HVO_CAMERA_NEAR 0.1f
#define HVO_CAMERA_FAR 100.0f
#define HVO_CAMERA_ZOOM 0.785398f

/* Camera::SetPositionAndTarget + UpdateInternal: projection (Y-flipped), its inverse and the view matrix */
static void camera_matrices(const float pos[3], const float target[3], float width, float height,
	float proj[4][4], float inv[4][4], float view[4][4])
{
	/* Camera::SetPositionAndTarget, Camera.cpp:81-93.  The unqualified asin/atan2/cos/sin calls
	 * in Camera.cpp bind to the C library's DOUBLE functions under g++ (verified with decltype);
	 * results narrow to float on assignment. */
	const v3 position = { pos[0], pos[1], pos[2] };
	v3 d = { target[0] - pos[0], target[1] - pos[1], target[2] - pos[2] };
	const v3 direction = normalize3(d);
	float pitch = (float)asin((double)direction.y);
	const float yaw = (float)atan2((double)direction.z, (double)direction.x);
	const float pitchLimit = 1.55334f; /* ConstrainPitch, Camera.cpp:95-100; glm::clamp = min(max(x,lo),hi) */
	pitch = fminf(fmaxf(pitch, -pitchLimit), pitchLimit);

	/* Camera::UpdateInternal, Camera.cpp:113-139 */
	v3 front;
	front.x = (float)(cos((double)yaw) * cos((double)pitch));
	front.y = (float)sin((double)pitch);
	front.z = (float)(sin((double)yaw) * cos((double)pitch));
	front = normalize3(front);
	const v3 worldUp = { 0.0f, 1.0f, 0.0f };
	const v3 right = normalize3(cross3(front, worldUp));
	const v3 up = normalize3(cross3(right, front));

	/* glm::perspective -> perspectiveRH_ZO (GLM_FORCE_DEPTH_ZERO_TO_ONE), ext/matrix_clip_space.inl:233-246 */
	const float aspect = width / height;
	const float tanHalfFovy = tanf(HVO_CAMERA_ZOOM / 2.0f);
	memset(proj, 0, 64);
	proj[0][0] = 1.0f / (aspect * tanHalfFovy);
	proj[1][1] = 1.0f / (tanHalfFovy);
	proj[2][2] = HVO_CAMERA_FAR / (HVO_CAMERA_NEAR - HVO_CAMERA_FAR);
	proj[2][3] = -1.0f;
	proj[3][2] = -(HVO_CAMERA_FAR * HVO_CAMERA_NEAR) / (HVO_CAMERA_FAR - HVO_CAMERA_NEAR);
	proj[1][1] *= -1; /* Camera.cpp:132 */
	mat4_inverse(proj, inv);

	/* glm::lookAt -> lookAtRH(position, position + front, up), ext/matrix_transform.inl:153-173 */
	const v3 center = { position.x + front.x, position.y + front.y, position.z + front.z };
	v3 cf = { center.x - position.x, center.y - position.y, center.z - position.z };
	const v3 f = normalize3(cf);
	const v3 s = normalize3(cross3(f, up));
	const v3 u = cross3(s, f);
	memset(view, 0, 64);
	view[0][0] = view[1][1] = view[2][2] = view[3][3] = 1.0f;
	view[0][0] = s.x; view[1][0] = s.y; view[2][0] = s.z;
	view[0][1] = u.x; view[1][1] = u.y; view[2][1] = u.z;
	view[0][2] = -f.x; view[1][2] = -f.y; view[2][2] = -f.z;
	view[3][0] = -dot3(s, position);
	view[3][1] = -dot3(u, position);
	view[3][2] = dot3(f, position);
}

void hvo_camera_ubo(const float pos[3], const float target[3], float width, float height,
	uint32_t X, uint32_t Y, uint32_t Z, hvcf_cluster_forward_ubo* out)
{
	float proj[4][4], inv[4][4], view[4][4];
	camera_matrices(pos, target, width, height, proj, inv, view);
	/* Camera::GetClusterForwardUBO, Camera.cpp:178-197 */
	memset(out, 0, sizeof(*out));
	memcpy(out->camera_inverse_projection, inv, 64);
	memcpy(out->camera_view, view, 64);
	out->screen_size[0] = width;
	out->screen_size[1] = height;
	const float zFloat = (float)Z;
	const float log2FarDivNear = log2f(HVO_CAMERA_FAR / HVO_CAMERA_NEAR);
	const float log2Near = log2f(HVO_CAMERA_NEAR);
	out->slice_scaling = zFloat / log2FarDivNear;
	out->slice_bias = -(zFloat * log2Near / log2FarDivNear);
	out->camera_near = HVO_CAMERA_NEAR;
	out->camera_far = HVO_CAMERA_FAR;
	out->slice_count_x = X;
	out->slice_count_y = Y;
	out->slice_count_z = Z;
}

/* ------------------------------------------------------------------------------------------
 * mat4 * vec4 in glm's association order (detail/type_mat4x4.inl:561-572):
 *   (m[0]*v0 + m[1]*v1) + (m[2]*v2 + m[3]*v3), per component, each product and sum rounded.
 * `m` is 16 floats, column-major.
 * ------------------------------------------------------------------------------------------ */
static void mat4_mul_vec4(const float* m, const float v[4], float out[4])
{
	for (int r = 0; r < 4; ++r)
	{
		const float mul0 = m[0 + r] * v[0];
		const float mul1 = m[4 + r] * v[1];
		const float add0 = mul0 + mul1;
		const float mul2 = m[8 + r] * v[2];
		const float mul3 = m[12 + r] * v[3];
		const float add1 = mul2 + mul3;
		out[r] = add0 + add1;
	}
}

/* AABBGenerator.comp:18-32 */
static void screen_to_view(const hvcf_cluster_forward_ubo* ubo, const float screen[4], float view[4])
{
	float clip[4];
	clip[0] = screen[0] * 2.0f - 1.0f;
	clip[1] = screen[1] * 2.0f - 1.0f;
	clip[2] = screen[2];
	clip[3] = screen[3];
	float v[4];
	mat4_mul_vec4(ubo->camera_inverse_projection, clip, v);
	const float w = v[3];
	for (int i = 0; i < 4; ++i)
		view[i] = v[i] / w;
}

/* AABBGenerator.comp:34-37: incident * zOffset / incident.z, evaluated left to right */
static void ray_intersect_z_plane(const float incident[3], float zOffset, float out[3])
{
	for (int i = 0; i < 3; ++i)
		out[i] = (incident[i] * zOffset) / incident[2];
}

static float min4(float a, float b, float c, float d) { return fminf(fminf(a, b), fminf(c, d)); }
static float max4(float a, float b, float c, float d) { return fmaxf(fmaxf(a, b), fmaxf(c, d)); }

/* AABBGenerator.comp:39-90, one invocation per (x, y, z) */
void hvo_generate_aabbs(const hvcf_cluster_forward_ubo* ubo, hvcf_aabb* out)
{
	const uint32_t X = ubo->slice_count_x, Y = ubo->slice_count_y, Z = ubo->slice_count_z;
	const float normX = 1.0f / (float)X;
	const float normY = 1.0f / (float)Y;
	for (uint32_t z = 0; z < Z; ++z)
		for (uint32_t y = 0; y < Y; ++y)
			for (uint32_t x = 0; x < X; ++x)
			{
				const uint32_t tileIndex = x + y * X + z * X * Y;
				const float maxSS[4] = { ((float)x + 1.0f) * normX, ((float)y + 1.0f) * normY, 0.0f, 1.0f };
				const float minSS[4] = { (float)x * normX, (float)y * normY, 0.0f, 1.0f };
				float maxVS[4], minVS[4];
				screen_to_view(ubo, maxSS, maxVS);
				screen_to_view(ubo, minSS, minVS);
				const float tileNear = -ubo->camera_near *
					powf(ubo->camera_far / ubo->camera_near, (float)z / (float)Z);
				const float tileFar = -ubo->camera_near *
					powf(ubo->camera_far / ubo->camera_near, (float)(z + 1) / (float)Z);
				float minNear[3], minFar[3], maxNear[3], maxFar[3];
				ray_intersect_z_plane(minVS, tileNear, minNear);
				ray_intersect_z_plane(minVS, tileFar, minFar);
				ray_intersect_z_plane(maxVS, tileNear, maxNear);
				ray_intersect_z_plane(maxVS, tileFar, maxFar);
				for (int i = 0; i < 3; ++i)
				{
					out[tileIndex].min_point[i] = min4(minNear[i], minFar[i], maxNear[i], maxFar[i]);
					out[tileIndex].max_point[i] = max4(minNear[i], minFar[i], maxNear[i], maxFar[i]);
				}
				out[tileIndex].min_point[3] = 0.0f;
				out[tileIndex].max_point[3] = 0.0f;
			}
}

/* LightCulling.comp:51-56: view-space position and r*r; depends on the light only */
void hvo_transform_lights(const hvcf_cluster_forward_ubo* ubo, const hvcf_light_data* lights,
	uint32_t count, float* out)
{
	for (uint32_t i = 0; i < count; ++i)
	{
		float pv[4];
		mat4_mul_vec4(ubo->camera_view, lights[i].position, pv);
		out[4 * (size_t)i + 0] = pv[0];
		out[4 * (size_t)i + 1] = pv[1];
		out[4 * (size_t)i + 2] = pv[2];
		out[4 * (size_t)i + 3] = lights[i].radius * lights[i].radius;
	}
}

/* LightCulling.comp:26-47 */
static inline float sq_dist_point_aabb(const float p[3], const hvcf_aabb* c)
{
	float sqDist = 0.0f;
	for (int i = 0; i < 3; ++i)
	{
		const float v = p[i];
		if (v < c->min_point[i])
		{
			const float diff = c->min_point[i] - v;
			sqDist += (diff * diff);
		}
		else if (v > c->max_point[i])
		{
			const float diff = c->max_point[i] - v;
			sqDist += (diff * diff);
		}
	}
	return sqDist;
}

/* the same with the product-sum contracted into fmaf, as a driver MAY do */
static inline float sq_dist_point_aabb_fma(const float p[3], const hvcf_aabb* c)
{
	float sqDist = 0.0f;
	for (int i = 0; i < 3; ++i)
	{
		const float v = p[i];
		if (v < c->min_point[i])
		{
			const float diff = c->min_point[i] - v;
			sqDist = fmaf(diff, diff, sqDist);
		}
		else if (v > c->max_point[i])
		{
			const float diff = c->max_point[i] - v;
			sqDist = fmaf(diff, diff, sqDist);
		}
	}
	return sqDist;
}

/* LightCulling.comp:59-77 for one cluster, on pre-transformed lights {x,y,z,r*r}: the transform
 * is a pure function of the light, so hoisting it out of the pair loop changes no bit. */
static uint32_t cull_one_xyzr2(const hvcf_aabb* aabb, const float* xyzr2, uint32_t num_lights,
	uint32_t cap, uint32_t* out_indices, uint64_t* executed, int use_fma)
{
	uint32_t lightIter = 0;
	uint32_t i = 0;
	for (; i < num_lights && lightIter < cap; ++i)
	{
		const float* l = xyzr2 + 4 * (size_t)i;
		const float sq = use_fma ? sq_dist_point_aabb_fma(l, aabb) : sq_dist_point_aabb(l, aabb);
		if (sq < l[3])
			out_indices[lightIter++] = i;
	}
	if (executed)
		*executed = i;
	return lightIter;
}

uint32_t hvo_cull_one(const hvcf_cluster_forward_ubo* ubo, const hvcf_aabb* aabb, const hvcf_light_data* lights,
	uint32_t num_lights, uint32_t cap, uint32_t* out_indices, uint64_t* out_executed)
{
	/* literal form: transform inside the pair loop, as TestSphereAABB does (LightCulling.comp:49-57) */
	uint32_t lightIter = 0;
	uint32_t i = 0;
	for (; i < num_lights && lightIter < cap; ++i)
	{
		const float lightRadius = lights[i].radius;
		float posView[4];
		mat4_mul_vec4(ubo->camera_view, lights[i].position, posView);
		const float squaredDistance = sq_dist_point_aabb(posView, aabb);
		if (squaredDistance < (lightRadius * lightRadius))
			out_indices[lightIter++] = i;
	}
	if (out_executed)
		*out_executed = i;
	return lightIter;
}

void hvo_cull(const hvcf_cluster_forward_ubo* ubo, const hvcf_aabb* aabbs, const hvcf_light_data* lights,
	uint32_t num_lights, uint32_t cap, uint32_t cluster_begin, uint32_t cluster_end,
	hvcf_light_cell* cells, uint32_t* indices, uint64_t* out_total, uint64_t* out_executed,
	int threads, int use_fma)
{
	const uint32_t n = cluster_end > cluster_begin ? cluster_end - cluster_begin : 0;
	float* xyzr2 = (float*)malloc(sizeof(float) * 4 * (size_t)(num_lights ? num_lights : 1));
	uint32_t* staging = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)cap * (size_t)(n ? n : 1));
	uint64_t* exec = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(n ? n : 1));
	hvo_transform_lights(ubo, lights, num_lights, xyzr2);
	const int nt = threads > 1 ? threads : 1;
	/* clusters per grab: 16 for whole frames, fewer for a small sample so that every thread still gets several */
	const int chunk = (int64_t)n >= 64 * (int64_t)nt ? 16 : ((int)n / (4 * nt) > 1 ? (int)n / (4 * nt) : 1);
	(void)nt;
	(void)chunk;
#pragma omp parallel for schedule(dynamic, chunk) num_threads(nt)
	for (int64_t c = 0; c < (int64_t)n; ++c)
	{
		cells[c].count = cull_one_xyzr2(&aabbs[cluster_begin + c], xyzr2, num_lights, cap,
			staging + (size_t)c * cap, &exec[c], use_fma);
	}
	uint64_t total = 0, executed = 0;
	for (uint32_t c = 0; c < n; ++c)
	{
		cells[c].offset = (uint32_t)total;
		memcpy(indices + total, staging + (size_t)c * cap, sizeof(uint32_t) * cells[c].count);
		total += cells[c].count;
		executed += exec[c];
	}
	if (out_total)
		*out_total = total;
	if (out_executed)
		*out_executed = executed;
	free(exec);
	free(staging);
	free(xyzr2);
}

void hvo_boundary_report(const hvcf_cluster_forward_ubo* ubo, const hvcf_aabb* aabbs,
	const hvcf_light_data* lights, uint32_t num_lights, uint32_t cluster_begin, uint32_t cluster_end,
	uint64_t* out_within_1ulp, uint64_t* out_fma_flips, int threads)
{
	float* xyzr2 = (float*)malloc(sizeof(float) * 4 * (size_t)(num_lights ? num_lights : 1));
	hvo_transform_lights(ubo, lights, num_lights, xyzr2);
	uint64_t near1 = 0, flips = 0;
	const int nt = threads > 1 ? threads : 1;
	(void)nt;
#pragma omp parallel for schedule(dynamic, 16) num_threads(nt) reduction(+ : near1, flips)
	for (int64_t c = (int64_t)cluster_begin; c < (int64_t)cluster_end; ++c)
	{
		for (uint32_t i = 0; i < num_lights; ++i)
		{
			const float* l = xyzr2 + 4 * (size_t)i;
			const float sq = sq_dist_point_aabb(l, &aabbs[c]);
			const float sqf = sq_dist_point_aabb_fma(l, &aabbs[c]);
			const float r2 = l[3];
			if (sq == r2 || sq == nextafterf(r2, INFINITY) || sq == nextafterf(r2, -INFINITY))
				++near1;
			if ((sq < r2) != (sqf < r2))
				++flips;
		}
	}
	*out_within_1ulp = near1;
	*out_fma_flips = flips;
	free(xyzr2);
}

/* Scene.frag:55-78 */
uint32_t hvo_cluster_index(const hvcf_cluster_forward_ubo* ubo, float frag_x, float frag_y, float frag_depth)
{
	const float near = ubo->camera_near, far = ubo->camera_far;
	const float linDepth = near * far / (far + frag_depth * (near - far));
	const float zf = fmaxf(log2f(linDepth) * ubo->slice_scaling + ubo->slice_bias, 0.0f);
	const uint32_t zIndex = (uint32_t)zf;
	const float tileX = ubo->screen_size[0] / (float)ubo->slice_count_x;
	const float tileY = ubo->screen_size[1] / (float)ubo->slice_count_y;
	const uint32_t cx = (uint32_t)(frag_x / tileX);
	const uint32_t cy = (uint32_t)(frag_y / tileY);
	return cx + cy * ubo->slice_count_x + zIndex * ubo->slice_count_x * ubo->slice_count_y;
}

/* ------------------------------------------------------------------------------------------
 * Consumer: light loop of the PBR cluster-forward fragment shader
 * ------------------------------------------------------------------------------------------ */
#define HVO_PI 3.1415926535897932384626433832795f /* Shaders/PBR/PBRHeader.glsl:10-12 (a 32-bit float literal in GLSL) */

static float distribution_ggx(float NoH, float roughness) /* PBRHeader.glsl:16-27 */
{
	const float alpha = roughness * roughness;
	const float alpha2 = alpha * alpha;
	const float NoH2 = NoH * NoH;
	const float nominator = alpha2;
	float denominator = (NoH2 * (alpha2 - 1.0f) + 1.0f);
	denominator = HVO_PI * denominator * denominator;
	return nominator / denominator;
}

static float alpha_direct_lighting(float roughness) /* PBRHeader.glsl:30-35 */
{
	const float r = (roughness + 1.0f);
	return (r * r) / 8.0f;
}

static float geometry_schlick_ggx(float NoL, float NoV, float alpha) /* PBRHeader.glsl:41-46 */
{
	const float GL = NoL / (NoL * (1.0f - alpha) + alpha);
	const float GV = NoV / (NoV * (1.0f - alpha) + alpha);
	return GL * GV;
}

static v3 fresnel_schlick(float cosTheta, v3 F0) /* PBRHeader.glsl:50-53 */
{
	const float c = fminf(fmaxf(1.0f - cosTheta, 0.0f), 1.0f);
	const float p = powf(c, 5.0f);
	v3 r = { F0.x + (1.0f - F0.x) * p, F0.y + (1.0f - F0.y) * p, F0.z + (1.0f - F0.z) * p };
	return r;
}

/* Shaders/ClusteredForward/Radiance.glsl:2-57 */
static v3 radiance(v3 worldPos, const hvcf_push_const_pbr* pc, v3 albedo, v3 N, v3 V, v3 F0, float metallic,
	float roughness, float alphaRoughness, float NoV, const hvcf_light_data* light)
{
	const v3 toLight = { light->position[0] - worldPos.x, light->position[1] - worldPos.y, light->position[2] - worldPos.z };
	const v3 L = normalize3(toLight);
	const v3 VL = { V.x + L.x, V.y + L.y, V.z + L.z };
	const v3 H = normalize3(VL);
	const float NoH = fmaxf(dot3(N, H), 0.0f);
	const float NoL = fmaxf(dot3(N, L), 0.0f);
	const float HoV = fmaxf(dot3(H, V), 0.0f);
	const float distance = sqrtf(dot3(toLight, toLight));
	const float attenuation = fmaxf(1.0f - (distance / light->radius), 0.0f) / powf(distance, pc->light_falloff);
	const v3 rad = { light->color[0] * attenuation * pc->light_intensity, light->color[1] * attenuation * pc->light_intensity,
		light->color[2] * attenuation * pc->light_intensity };
	const float D = distribution_ggx(NoH, roughness);
	const float G = geometry_schlick_ggx(NoL, NoV, alphaRoughness);
	const v3 F = fresnel_schlick(HoV, F0);
	const float DG = D * G;
	const float denominator = 4.0f * NoV * NoL + 0.0001f;
	const v3 specular = { DG * F.x / denominator, DG * F.y / denominator, DG * F.z / denominator };
	const float om = 1.0f - metallic;
	const v3 kD = { (1.0f - F.x) * om, (1.0f - F.y) * om, (1.0f - F.z) * om };
	const v3 diffuse = { kD.x * albedo.x / HVO_PI, kD.y * albedo.y / HVO_PI, kD.z * albedo.z / HVO_PI };
	v3 Lo = { (diffuse.x + specular.x) * rad.x * NoL, (diffuse.y + specular.y) * rad.y * NoL, (diffuse.z + specular.z) * rad.z * NoL };
	return Lo;
}

void hvo_shade(const hvcf_cluster_forward_ubo* ubo, const hvcf_light_cell* cells, const uint32_t* indices,
	const hvcf_light_data* lights, const hvcf_push_const_pbr* pc, const float camera_position[3],
	uint32_t width, uint32_t height, uint64_t pixel_begin, uint64_t pixel_end, uint64_t pixel_stride,
	const float* frag_depth, const float* world_pos, const float* normal_rough, const float* albedo_metal,
	float* out_color, uint32_t* out_cluster, int threads)
{
	const int nt = threads > 1 ? threads : 1;
	(void)nt;
	(void)height;
	if (pixel_stride == 0)
		pixel_stride = 1;
	const uint32_t X = ubo->slice_count_x, Y = ubo->slice_count_y, Z = ubo->slice_count_z;
	const int64_t n = (int64_t)((pixel_end - pixel_begin + pixel_stride - 1) / pixel_stride);
#pragma omp parallel for schedule(dynamic, 256) num_threads(nt)
	for (int64_t k = 0; k < n; ++k)
	{
		const uint64_t p = pixel_begin + (uint64_t)k * pixel_stride;
		const float fx = (float)(p % width) + 0.5f, fy = (float)(p / width) + 0.5f;
		/* Scene.frag:64-78 */
		uint32_t clusterIdx = hvo_cluster_index(ubo, fx, fy, frag_depth[p]);
		if (clusterIdx >= X * Y * Z) /* guard, see header */
			clusterIdx = (clusterIdx % (X * Y)) + (Z - 1) * X * Y;
		const uint32_t lightCount = cells[clusterIdx].count;
		const uint32_t lightIndexOffset = cells[clusterIdx].offset;
		const v3 worldPos = { world_pos[4 * p], world_pos[4 * p + 1], world_pos[4 * p + 2] };
		const v3 N = { normal_rough[4 * p], normal_rough[4 * p + 1], normal_rough[4 * p + 2] };
		const float roughness = normal_rough[4 * p + 3];
		const v3 albedo = { albedo_metal[4 * p], albedo_metal[4 * p + 1], albedo_metal[4 * p + 2] };
		const float metallic = albedo_metal[4 * p + 3];
		const float alphaRoughness = alpha_direct_lighting(roughness); /* Scene.frag:103 */
		const v3 toCam = { camera_position[0] - worldPos.x, camera_position[1] - worldPos.y, camera_position[2] - worldPos.z };
		const v3 V = normalize3(toCam); /* :107 */
		const float NoV = fmaxf(dot3(N, V), 0.0f); /* :108 */
		/* :112-113  F0 = mix(vec3(baseReflectivity), albedo, metallic) = x*(1-a) + y*a */
		const float b = pc->base_reflectivity;
		const v3 F0 = { b * (1.0f - metallic) + albedo.x * metallic, b * (1.0f - metallic) + albedo.y * metallic,
			b * (1.0f - metallic) + albedo.z * metallic };
		v3 Lo = { albedo.x * pc->albedo_multipler, albedo.y * pc->albedo_multipler, albedo.z * pc->albedo_multipler }; /* :117 */
		for (uint32_t i = 0; i < lightCount; ++i) /* :119-134 */
		{
			const uint32_t lightIndex = indices[i + lightIndexOffset];
			const v3 r = radiance(worldPos, pc, albedo, N, V, F0, metallic, roughness, alphaRoughness, NoV, &lights[lightIndex]);
			Lo.x += r.x;
			Lo.y += r.y;
			Lo.z += r.z;
		}
		if (out_color)
		{
			out_color[4 * p] = Lo.x;
			out_color[4 * p + 1] = Lo.y;
			out_color[4 * p + 2] = Lo.z;
			out_color[4 * p + 3] = 1.0f;
		}
		if (out_cluster)
			out_cluster[p] = clusterIdx;
	}
}

/* ------------------------------------------------------------------------------------------
 * Frustum culling
 * ------------------------------------------------------------------------------------------ */
/* glm mat4 * mat4, detail/type_mat4x4.inl:630-648: Result[c] = ((A0*b0 + A1*b1) + A2*b2) + A3*b3 */
static void mat4_mul_mat4(const float a[4][4], const float b[4][4], float out[4][4])
{
	for (int c = 0; c < 4; ++c)
		for (int r = 0; r < 4; ++r)
		{
			const float t0 = a[0][r] * b[c][0], t1 = a[1][r] * b[c][1], t2 = a[2][r] * b[c][2], t3 = a[3][r] * b[c][3];
			out[c][r] = ((t0 + t1) + t2) + t3;
		}
}

/* Camera::GetFrustumUBO, Source/Camera.cpp:199-234 */
void hvo_frustum_ubo(const float pos[3], const float target[3], float width, float height, hvcf_frustum_ubo* out)
{
	float proj[4][4], inv[4][4], view[4][4], projView[4][4], t[4][4], invProjView[4][4];
	camera_matrices(pos, target, width, height, proj, inv, view);
	mat4_mul_mat4(proj, view, projView);
	for (int c = 0; c < 4; ++c) /* glm::transpose */
		for (int r = 0; r < 4; ++r)
			t[c][r] = projView[r][c];
	memset(out, 0, sizeof(*out));
	for (int k = 0; k < 4; ++k)
	{
		out->planes[0][k] = t[3][k] + t[0][k]; /* left */
		out->planes[1][k] = t[3][k] - t[0][k]; /* right */
		out->planes[2][k] = t[3][k] + t[1][k]; /* bottom */
		out->planes[3][k] = t[3][k] - t[1][k]; /* top */
		out->planes[4][k] = t[3][k] + t[2][k]; /* near */
		out->planes[5][k] = t[3][k] - t[2][k]; /* far */
	}
	mat4_inverse(projView, invProjView);
	static const float corners[8][4] = {
		{ -1, -1, -1, 1 }, { 1, -1, -1, 1 }, { 1, 1, -1, 1 }, { -1, 1, -1, 1 },
		{ -1, -1, 1, 1 }, { 1, -1, 1, 1 }, { 1, 1, 1, 1 }, { -1, 1, 1, 1 } };
	for (int i = 0; i < 8; ++i)
	{
		float q[4];
		mat4_mul_vec4(&invProjView[0][0], corners[i], q);
		for (int k = 0; k < 4; ++k)
			out->corners[i][k] = q[k] / q[3];
	}
}

/* dot(vec4, vec4) as glm evaluates it off MSVC: (x + y) + (z + w), detail/func_geometric.inl:57-68 */
static inline float plane_dot(const float p[4], float x, float y, float z)
{
	const float tx = p[0] * x, ty = p[1] * y, tz = p[2] * z, tw = p[3] * 1.0f;
	return (tx + ty) + (tz + tw);
}

/* FrustumCulling.comp:14-44 */
static int is_box_in_frustum(const hvcf_frustum_ubo* f, const hvcf_aabb* box)
{
	const float* mn = box->min_point;
	const float* mx = box->max_point;
	for (int i = 0; i < 6; i++)
	{
		int r = 0;
		r += (plane_dot(f->planes[i], mn[0], mn[1], mn[2]) < 0.0f) ? 1 : 0;
		r += (plane_dot(f->planes[i], mx[0], mn[1], mn[2]) < 0.0f) ? 1 : 0;
		r += (plane_dot(f->planes[i], mn[0], mx[1], mn[2]) < 0.0f) ? 1 : 0;
		r += (plane_dot(f->planes[i], mx[0], mx[1], mn[2]) < 0.0f) ? 1 : 0;
		r += (plane_dot(f->planes[i], mn[0], mn[1], mx[2]) < 0.0f) ? 1 : 0;
		r += (plane_dot(f->planes[i], mx[0], mn[1], mx[2]) < 0.0f) ? 1 : 0;
		r += (plane_dot(f->planes[i], mn[0], mx[1], mx[2]) < 0.0f) ? 1 : 0;
		r += (plane_dot(f->planes[i], mx[0], mx[1], mx[2]) < 0.0f) ? 1 : 0;
		if (r == 8)
			return 0;
	}
	for (int axis = 0; axis < 3; ++axis)
	{
		int r = 0;
		for (int i = 0; i < 8; i++)
			r += (f->corners[i][axis] > mx[axis]) ? 1 : 0;
		if (r == 8)
			return 0;
		r = 0;
		for (int i = 0; i < 8; i++)
			r += (f->corners[i][axis] < mn[axis]) ? 1 : 0;
		if (r == 8)
			return 0;
	}
	return 1;
}

/* FrustumCulling.comp:46-58; only instanceCount is written */
void hvo_frustum_cull(const hvcf_frustum_ubo* frustum, const hvcf_aabb* boxes, uint32_t count,
	hvcf_draw_indirect_command* commands, int threads)
{
	const int nt = threads > 1 ? threads : 1;
	(void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
	for (int64_t i = 0; i < (int64_t)count; ++i)
		commands[i].instance_count = is_box_in_frustum(frustum, &boxes[i]) ? 1u : 0u;
}

int hvo_max_threads(void)
{
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}
