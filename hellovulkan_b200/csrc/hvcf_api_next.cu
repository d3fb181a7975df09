// hvcf_api_next.cu -- C ABI of the rows next to the path (kernels in hvcf_shade.cu): the consumer of the light grid
// (PipelinePBRClusterForward::FillCommandBuffer, Source/Pipelines/PipelinePBRClusterForward.cpp:88-143, for Scene.frag:55-78,103-134),
// compute frustum culling (PipelineFrustumCulling::Execute, Source/Pipelines/PipelineFrustumCulling.cpp:32-69) and the measured FP32
// issue rate that bench.py's roofline divides by.
#include "hvcf_ctx.cuh"

#include <cmath>
#include <cstring>

#include "hvcf_shade.cuh"

using namespace hvcf;

extern "C" {

// ---------------------------------------------------------------------------------------------------------
// consumer: Scene.frag light loop over a device G-buffer
// ---------------------------------------------------------------------------------------------------------
namespace
{
// Scene.frag:65 on the host, with the libm the oracle uses: uint(max(log2(linDepth) * scaling + bias, 0.0))
inline uint32_t host_z_index(float lin, float scaling, float bias)
{
	const float v = fmaxf(log2f(lin) * scaling + bias, 0.0f);
	return v >= 4294967040.0f ? 0xffffffffu : static_cast<uint32_t>(v);
}

// thr[k] = the smallest positive float whose zIndex is >= k (bisection over the bit patterns of positive floats:
// log2f is monotone, and so is the rounded multiply-add), +inf when no finite depth reaches slice k.
void build_slice_thresholds(float scaling, float bias, uint32_t Z, float* thr)
{
	thr[0] = 0.0f;
	for (uint32_t k = 1; k <= Z; ++k)
	{
		uint32_t lo = 0x00800000u, hi = 0x7f7fffffu; // smallest normal .. FLT_MAX
		float f;
		memcpy(&f, &hi, 4);
		if (host_z_index(f, scaling, bias) < k)
		{
			thr[k] = INFINITY;
			continue;
		}
		memcpy(&f, &lo, 4);
		if (host_z_index(f, scaling, bias) >= k)
		{
			thr[k] = f;
			continue;
		}
		while (hi - lo > 1) // invariant: zIndex(lo) < k <= zIndex(hi)
		{
			const uint32_t mid = lo + (hi - lo) / 2;
			memcpy(&f, &mid, 4);
			if (host_z_index(f, scaling, bias) >= k)
				hi = mid;
			else
				lo = mid;
		}
		memcpy(&thr[k], &hi, 4);
	}
}
} // namespace

int hvcf_shade(hvcf_ctx* ctx, uint32_t frame, const hvcf_gbuffer* g, const hvcf_push_const_pbr* pc, const float camera_position[3],
	float* out_color, uint32_t* out_cluster_index, void* stream)
{
	if (!ctx || !g || !pc || !camera_position || !out_color)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_shade: null argument");
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_shade: frame %u >= frame_count %u", frame, ctx->frame_count);
	if (!ctx->ubo_set[frame])
		return fail(HVCF_ERR_NOT_READY, "hvcf_shade: hvcf_set_ubo(frame %u) has not been called", frame);
	if (!g->width || !g->height || !g->frag_depth || !g->world_pos || !g->normal_rough || !g->albedo_metal)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_shade: incomplete G-buffer");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	const size_t tab = ctx->Z + 1;
	if (!ctx->d_slice_thr)
	{
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_slice_thr), sizeof(float) * tab * ctx->frame_count));
		HVCF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&ctx->h_slice_thr), sizeof(float) * tab * ctx->frame_count));
	}
	const hvcf_cluster_forward_ubo& u = ctx->h_ubo[frame];
	if (!ctx->slice_valid[frame] || ctx->slice_key[frame][0] != u.slice_scaling || ctx->slice_key[frame][1] != u.slice_bias)
	{
		// the slot's previous table may still be in flight on `stream` only if the caller reuses a frame slot before its
		// work finished, which the 2-deep frame ring forbids (AppBase.cpp:96-100)
		build_slice_thresholds(u.slice_scaling, u.slice_bias, ctx->Z, ctx->h_slice_thr + tab * frame);
		HVCF_CUDA(cudaMemcpyAsync(ctx->d_slice_thr + tab * frame, ctx->h_slice_thr + tab * frame, sizeof(float) * tab,
			cudaMemcpyHostToDevice, s));
		ctx->slice_key[frame][0] = u.slice_scaling;
		ctx->slice_key[frame][1] = u.slice_bias;
		ctx->slice_valid[frame] = true;
	}
	ShadeParams p{};
	p.cells = ctx->d_cells;
	p.indices = ctx->d_indices;
	p.lights = ctx->bound_lights ? ctx->bound_lights : ctx->d_lights;
	p.slice_thresholds = ctx->d_slice_thr + tab * frame;
	p.pc = *pc;
	p.cam[0] = camera_position[0], p.cam[1] = camera_position[1], p.cam[2] = camera_position[2];
	p.near_plane = u.camera_near, p.far_plane = u.camera_far;
	p.slice_scaling = u.slice_scaling, p.slice_bias = u.slice_bias;
	p.tile_w = u.screen_size[0] / static_cast<float>(u.slice_count_x); // Scene.frag:66-68
	p.tile_h = u.screen_size[1] / static_cast<float>(u.slice_count_y);
	p.X = ctx->X, p.Y = ctx->Y, p.Z = ctx->Z;
	p.width = g->width, p.height = g->height;
	p.frag_depth = g->frag_depth;
	p.world_pos = reinterpret_cast<const float4*>(g->world_pos);
	p.normal_rough = reinterpret_cast<const float4*>(g->normal_rough);
	p.albedo_metal = reinterpret_cast<const float4*>(g->albedo_metal);
	p.out_color = reinterpret_cast<float4*>(out_color);
	p.out_cluster = out_cluster_index;
	HVCF_CUDA(launch_cluster_shade(p, s));
	return HVCF_OK;
}

// ---------------------------------------------------------------------------------------------------------
// FrustumCulling.comp
// ---------------------------------------------------------------------------------------------------------
int hvcf_frustum_cull(int device, const hvcf_frustum_ubo* frustum, const hvcf_aabb* boxes, uint32_t count,
	hvcf_draw_indirect_command* commands, void* stream)
{
	if (!frustum || (count && (!boxes || !commands)))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_frustum_cull: null argument");
	if (device >= 0)
		HVCF_CUDA(cudaSetDevice(device));
	FrustumParams f{};
	memcpy(f.planes, frustum->planes, sizeof(f.planes));
	for (int a = 0; a < 3; ++a)
	{
		float mn = INFINITY, mx = -INFINITY;
		bool nan = false;
		for (int i = 0; i < 8; ++i)
		{
			const float c = frustum->corners[i][a];
			nan = nan || c != c;
			mn = fminf(mn, c);
			mx = fmaxf(mx, c);
		}
		// a NaN corner makes every `corner > max` / `corner < min` sum fall short of 8 in the shader (:36-41)
		f.corner_min[a] = nan ? -INFINITY : mn;
		f.corner_max[a] = nan ? INFINITY : mx;
	}
	HVCF_CUDA(launch_frustum_cull(f, boxes, count, commands, static_cast<cudaStream_t>(stream)));
	return HVCF_OK;
}

int hvcf_measure_fp32_issue_rate(int device, double* out_lane_ops_per_second)
{
	if (!out_lane_ops_per_second)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_measure_fp32_issue_rate: null argument");
	if (device >= 0)
		HVCF_CUDA(cudaSetDevice(device));
	cudaDeviceProp prop;
	int dev = 0;
	HVCF_CUDA(cudaGetDevice(&dev));
	HVCF_CUDA(cudaGetDeviceProperties(&prop, dev));
	float* sink = nullptr;
	HVCF_CUDA(cudaMalloc(&sink, 16));
	cudaEvent_t e0, e1;
	HVCF_CUDA(cudaEventCreate(&e0));
	HVCF_CUDA(cudaEventCreate(&e1));
	const uint32_t blocks = static_cast<uint32_t>(prop.multiProcessorCount) * 8, iters = 1u << 15;
	double best = 0.0;
	for (int rep = 0; rep < 5; ++rep)
	{
		HVCF_CUDA(cudaEventRecord(e0, 0));
		HVCF_CUDA(launch_fp32_issue_rate(sink, blocks, iters, 0));
		HVCF_CUDA(cudaEventRecord(e1, 0));
		HVCF_CUDA(cudaEventSynchronize(e1));
		float ms = 0.0f;
		HVCF_CUDA(cudaEventElapsedTime(&ms, e0, e1));
		const double ops = static_cast<double>(blocks) * 256.0 * iters * 16.0;
		if (rep > 0 && ms > 0.0f && ops / (ms * 1e-3) > best)
			best = ops / (ms * 1e-3);
	}
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	cudaFree(sink);
	*out_lane_ops_per_second = best;
	return HVCF_OK;
}

} // extern "C"
