// hvcf_shade.cu -- the consumer of the light grid and the sibling culling pass, as CUDA kernels for sm_100a.
//
// cluster_shade_kernel: one thread per pixel of a device G-buffer.
//   * cluster lookup (Scene.frag:55-78) is EXACT: linear depth with explicit round-to-nearest ops (this file is
//     compiled with FMA contraction on), the z slice from a host-built threshold table (the GLSL log2 has no
//     reproducible rounding; the table is built with the same libm call the oracle makes, so
//     zIndex = #{k : depth >= threshold[k]} is bit-identical by construction), tile x/y by IEEE division.
//   * light loop (Scene.frag:119-134, Radiance.glsl, PBRHeader.glsl): float tolerance.  A warp covers 32
//     consecutive pixels; lanes are grouped by cluster, and each group's light list is staged 32 lights at a
//     time into shared memory by ALL lanes of the warp (index load + 48-byte light gather in parallel instead
//     of a dependent chain per light), then evaluated from shared memory in list order.
//     Lights that cannot contribute are skipped exactly as the shader's arithmetic would zero them:
//     attenuation = max(1 - d/r, 0) is 0 beyond the radius, and the term is multiplied by NoL = max(N.L, 0).
//
// frustum_cull_kernel: one thread per box, planes in the constant bank, bit-exact (explicit _rn ops in glm's
// association order), one 4-byte store per 16-byte indirect command.
#include "hvcf_shade.cuh"

namespace hvcf
{

namespace
{
constexpr uint32_t kShadeWarps = 8;
constexpr float kPi = 3.1415926535897932384626433832795f; // PBRHeader.glsl:10-12

// glm dot(vec3, vec3): (x*x' + y*y') + z*z', every product and sum rounded (detail/func_geometric.inl:48-55).  The
// directions L, H, V and the cosines built from them are kept bit-identical to the shader compiled as C++: the GGX
// denominator NoH^2 * (alpha^2 - 1) + 1 cancels down to ~alpha^2 at a highlight, so one ulp of NoH is percents of D.
__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz)
{
	return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by)), __fmul_rn(az, bz));
}
// glm inversesqrt = 1 / sqrt(x), both correctly rounded
__device__ __forceinline__ float inv_sqrt_exact(float x, float* root)
{
	*root = __fsqrt_rn(x);
	return __frcp_rn(*root);
}

// Scene.frag:55-58 with every operation rounded on its own
__device__ __forceinline__ float linear_depth(float z, float n, float f)
{
	return __fdiv_rn(__fmul_rn(n, f), __fadd_rn(f, __fmul_rn(z, __fsub_rn(n, f))));
}
} // namespace

__global__ void __launch_bounds__(kShadeWarps * 32) cluster_shade_kernel(const ShadeParams p)
{
	__shared__ float4 s_pos[kShadeWarps][32]; // light position.xyz, 1 / radius
	__shared__ float4 s_col[kShadeWarps][32]; // light color.rgb, radius * radius
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t total = static_cast<uint64_t>(p.width) * p.height;
	const uint64_t pix = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
	const bool in_range = pix < total;
	const uint64_t q = in_range ? pix : total - 1;

	// ---- Scene.frag:64-78 -------------------------------------------------------------------------------
	const uint32_t px = static_cast<uint32_t>(q % p.width), py = static_cast<uint32_t>(q / p.width);
	const float fx = static_cast<float>(px) + 0.5f, fy = static_cast<float>(py) + 0.5f;
	const float lin = linear_depth(__ldg(p.frag_depth + q), p.near_plane, p.far_plane);
	uint32_t zi = min(static_cast<uint32_t>(fmaxf(__log2f(lin) * p.slice_scaling + p.slice_bias, 0.0f)), p.Z);
	while (zi < p.Z && lin >= __ldg(p.slice_thresholds + zi + 1)) // the guess is within one slice of the table
		++zi;
	while (zi > 0 && !(lin >= __ldg(p.slice_thresholds + zi)))
		--zi;
	const uint32_t cx = __float2uint_rz(__fdiv_rn(fx, p.tile_w)), cy = __float2uint_rz(__fdiv_rn(fy, p.tile_h));
	uint32_t cluster = cx + cy * p.X + zi * p.X * p.Y;
	if (cluster >= p.X * p.Y * p.Z) // only reachable with frag_depth == 1 (see include/hvcf.h)
		cluster = (cluster % (p.X * p.Y)) + (p.Z - 1) * p.X * p.Y;
	const uint2 cell = __ldg(reinterpret_cast<const uint2*>(p.cells) + cluster); // {offset, count}

	// ---- per-pixel terms (Scene.frag:103-117) -----------------------------------------------------------
	const float4 wp = __ldg(p.world_pos + q), nr = __ldg(p.normal_rough + q), am = __ldg(p.albedo_metal + q);
	const float roughness = nr.w, metallic = am.w;
	const float ar = (roughness + 1.0f) * (roughness + 1.0f) / 8.0f; // AlphaDirectLighting
	float vx = p.cam[0] - wp.x, vy = p.cam[1] - wp.y, vz = p.cam[2] - wp.z;
	float vlen;
	const float vinv = inv_sqrt_exact(dot3(vx, vy, vz, vx, vy, vz), &vlen);
	vx = __fmul_rn(vx, vinv), vy = __fmul_rn(vy, vinv), vz = __fmul_rn(vz, vinv);
	const float NoV = fmaxf(dot3(nr.x, nr.y, nr.z, vx, vy, vz), 0.0f);
	const float b = p.pc.base_reflectivity;
	const float f0x = b * (1.0f - metallic) + am.x * metallic, f0y = b * (1.0f - metallic) + am.y * metallic,
				f0z = b * (1.0f - metallic) + am.z * metallic;
	float lox = am.x * p.pc.albedo_multipler, loy = am.y * p.pc.albedo_multipler, loz = am.z * p.pc.albedo_multipler;
	// loop invariants of Radiance()
	const float alpha = __fmul_rn(roughness, roughness), alpha2 = __fmul_rn(alpha, alpha), alpha2m1 = __fsub_rn(alpha2, 1.0f);
	const float GV = NoV / (NoV * (1.0f - ar) + ar);
	const float om = 1.0f - metallic;
	const float falloff = p.pc.light_falloff, intensity = p.pc.light_intensity;
	const bool unit_falloff = falloff == 1.0f;

	// ---- lanes grouped by cluster; every group's list staged by the whole warp ---------------------------
	uint32_t remaining = __ballot_sync(0xffffffffu, in_range && cell.y != 0);
	while (remaining)
	{
		const int leader = __ffs(remaining) - 1;
		const uint32_t lc = __shfl_sync(0xffffffffu, cluster, leader);
		const uint32_t off = __shfl_sync(0xffffffffu, cell.x, leader), cnt = __shfl_sync(0xffffffffu, cell.y, leader);
		const bool mine = in_range && cluster == lc;
		remaining &= ~__ballot_sync(0xffffffffu, mine);
		for (uint32_t base = 0; base < cnt; base += 32)
		{
			const uint32_t n = min(32u, cnt - base);
			if (lane < n)
			{
				const uint32_t li = __ldg(p.indices + off + base + lane);
				const float4* lp = reinterpret_cast<const float4*>(p.lights + li);
				const float4 pos = __ldg(lp), col = __ldg(lp + 1);
				const float radius = __ldg(reinterpret_cast<const float*>(lp + 2));
				s_pos[warp][lane] = make_float4(pos.x, pos.y, pos.z, 1.0f / radius);
				s_col[warp][lane] = make_float4(col.x, col.y, col.z, radius * radius);
			}
			__syncwarp();
			if (mine)
			{
				for (uint32_t k = 0; k < n; ++k)
				{
					const float4 lp = s_pos[warp][k], lcol = s_col[warp][k];
					const float lx = __fsub_rn(lp.x, wp.x), ly = __fsub_rn(lp.y, wp.y), lz = __fsub_rn(lp.z, wp.z);
					const float d2 = dot3(lx, ly, lz, lx, ly, lz);
					if (!(d2 < lcol.w))
						continue; // attenuation = max(1 - d/r, 0) = 0
					float dist;
					const float dinv = inv_sqrt_exact(d2, &dist);
					const float Lx = __fmul_rn(lx, dinv), Ly = __fmul_rn(ly, dinv), Lz = __fmul_rn(lz, dinv);
					const float NoL = fmaxf(dot3(nr.x, nr.y, nr.z, Lx, Ly, Lz), 0.0f);
					if (!(NoL > 0.0f))
						continue; // the term is multiplied by NoL
					float hx = __fadd_rn(vx, Lx), hy = __fadd_rn(vy, Ly), hz = __fadd_rn(vz, Lz), hlen;
					const float hinv = inv_sqrt_exact(dot3(hx, hy, hz, hx, hy, hz), &hlen);
					hx = __fmul_rn(hx, hinv), hy = __fmul_rn(hy, hinv), hz = __fmul_rn(hz, hinv);
					const float NoH = fmaxf(dot3(nr.x, nr.y, nr.z, hx, hy, hz), 0.0f);
					const float HoV = fmaxf(dot3(hx, hy, hz, vx, vy, vz), 0.0f);
					const float powd = unit_falloff ? dist : __powf(dist, falloff);
					const float att = __fdividef(fmaxf(1.0f - dist * lp.w, 0.0f), powd) * intensity;
					const float den = __fadd_rn(__fmul_rn(__fmul_rn(NoH, NoH), alpha2m1), 1.0f);
					const float D = __fdividef(alpha2, kPi * den * den);
					const float G = __fdividef(NoL, NoL * (1.0f - ar) + ar) * GV;
					const float c = fminf(fmaxf(1.0f - HoV, 0.0f), 1.0f), c2 = c * c, p5 = c2 * c2 * c;
					const float Fx = f0x + (1.0f - f0x) * p5, Fy = f0y + (1.0f - f0y) * p5, Fz = f0z + (1.0f - f0z) * p5;
					const float sden = __fdividef(D * G, 4.0f * NoV * NoL + 0.0001f);
					const float w = att * NoL;
					lox += ((1.0f - Fx) * om * am.x * (1.0f / kPi) + Fx * sden) * (lcol.x * w);
					loy += ((1.0f - Fy) * om * am.y * (1.0f / kPi) + Fy * sden) * (lcol.y * w);
					loz += ((1.0f - Fz) * om * am.z * (1.0f / kPi) + Fz * sden) * (lcol.z * w);
				}
			}
			__syncwarp();
		}
	}
	if (in_range)
	{
		p.out_color[pix] = make_float4(lox, loy, loz, 1.0f);
		if (p.out_cluster)
			p.out_cluster[pix] = cluster;
	}
}

cudaError_t launch_cluster_shade(const ShadeParams& p, cudaStream_t stream)
{
	const uint64_t total = static_cast<uint64_t>(p.width) * p.height;
	if (total == 0)
		return cudaSuccess;
	const uint32_t threads = kShadeWarps * 32;
	const uint64_t blocks = (total + threads - 1) / threads;
	if (blocks > 0x7fffffffull)
		return cudaErrorInvalidValue;
	cluster_shade_kernel<<<static_cast<uint32_t>(blocks), threads, 0, stream>>>(p);
	return cudaGetLastError();
}

// ======================================================================================================
// FrustumCulling.comp:14-57.  dot(plane, vec4(x, y, z, 1)) in glm's association order (x + y) + (z + w), every
// product and sum rounded on its own; the 8 corner sums share their 6 products and 6 partial sums.
// ======================================================================================================
__global__ void __launch_bounds__(256) frustum_cull_kernel(const FrustumParams f, const hvcf_aabb* __restrict__ boxes, uint32_t count,
	hvcf_draw_indirect_command* __restrict__ commands)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count)
		return;
	const float4* b = reinterpret_cast<const float4*>(boxes + i);
	const float4 mn = __ldg(b), mx = __ldg(b + 1);
	bool inside = true;
#pragma unroll
	for (int k = 0; k < 6; ++k)
	{
		const float px = f.planes[k][0], py = f.planes[k][1], pz = f.planes[k][2], pw = f.planes[k][3];
		const float x0 = __fmul_rn(px, mn.x), x1 = __fmul_rn(px, mx.x);
		const float y0 = __fmul_rn(py, mn.y), y1 = __fmul_rn(py, mx.y);
		const float z0 = __fadd_rn(__fmul_rn(pz, mn.z), pw), z1 = __fadd_rn(__fmul_rn(pz, mx.z), pw); // pw * 1.0 is exact
		const float xy00 = __fadd_rn(x0, y0), xy10 = __fadd_rn(x1, y0), xy01 = __fadd_rn(x0, y1), xy11 = __fadd_rn(x1, y1);
		const bool all_out = (__fadd_rn(xy00, z0) < 0.0f) & (__fadd_rn(xy10, z0) < 0.0f) & (__fadd_rn(xy01, z0) < 0.0f) &
							 (__fadd_rn(xy11, z0) < 0.0f) & (__fadd_rn(xy00, z1) < 0.0f) & (__fadd_rn(xy10, z1) < 0.0f) &
							 (__fadd_rn(xy01, z1) < 0.0f) & (__fadd_rn(xy11, z1) < 0.0f);
		inside = inside && !all_out;
	}
	// frustum outside/inside box (:36-41): all eight corners beyond one face
	inside = inside && !(f.corner_min[0] > mx.x) && !(f.corner_max[0] < mn.x) && !(f.corner_min[1] > mx.y) &&
			 !(f.corner_max[1] < mn.y) && !(f.corner_min[2] > mx.z) && !(f.corner_max[2] < mn.z);
	commands[i].instance_count = inside ? 1u : 0u;
}

cudaError_t launch_frustum_cull(const FrustumParams& f, const hvcf_aabb* boxes, uint32_t count, hvcf_draw_indirect_command* commands,
	cudaStream_t stream)
{
	if (count == 0)
		return cudaSuccess;
	frustum_cull_kernel<<<(count + 255) / 256, 256, 0, stream>>>(f, boxes, count, commands);
	return cudaGetLastError();
}

} // namespace hvcf
