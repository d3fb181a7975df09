// hvcf_api_io.cu -- C ABI of everything that reads results back or moves raw buffers (include/hvcf.h): VulkanBuffer::DownloadBufferData
// (Source/Vulkan/VulkanBuffer.cpp:169-178) as hvcf_download_*, the device pointers of the SSBOs, work counters / timings / debug
// views of a frame, and the raw device and pinned host buffers behind include/hv/Context.h (VulkanBuffer.cpp:6-50, :148-178).
#include "hvcf_ctx.cuh"

#include <cstring>

using namespace hvcf;

extern "C" {

int hvcf_shard_total_ptr(hvcf_ctx* ctx, uint32_t** out_device_ptr)
{
	if (!ctx || !out_device_ptr)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_shard_total_ptr: null argument");
	*out_device_ptr = ctx->d_shard_total;
	return HVCF_OK;
}

int hvcf_get_device_buffers(hvcf_ctx* ctx, hvcf_buffers* out)
{
	if (!ctx || !out)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_get_device_buffers: null argument");
	memset(out, 0, sizeof(*out));
	out->aabb = ctx->d_aabb;
	out->light_cells = ctx->d_cells;
	out->light_indices = ctx->d_indices;
	out->global_index_count = ctx->d_global_index;
	out->lights = ctx->bound_lights ? const_cast<hvcf_light_data*>(ctx->bound_lights) : ctx->d_lights;
	out->lights_view = reinterpret_cast<float*>(ctx->d_lights_view);
	out->light_indices_capacity = static_cast<uint64_t>(ctx->cluster_total) * ctx->cap;
	out->cluster_count = ctx->cluster_total;
	out->light_count = ctx->num_lights;
	out->cluster_begin = ctx->cluster_begin;
	out->cluster_end = ctx->cluster_end;
	return HVCF_OK;
}

// Sharded contexts: a download is the caller's next synchronisation point after a frame, so a wait that timed out inside
// the frame (hvcf_peer.cu; the push kernels then skipped their stores) is reported here instead of handing out a stale grid.
// Called after the stream has been synchronised.  One 4-byte read, only when peers are attached.
int hvcf_peer_error_after_sync(hvcf_ctx* ctx, const char* what)
{
	if (!ctx->d_mail || ctx->peer.world < 2)
		return HVCF_OK;
	unsigned int err = 0;
	HVCF_CUDA(cudaMemcpy(&err, &ctx->d_mail->error, sizeof(err), cudaMemcpyDeviceToHost));
	if (err)
		return fail(HVCF_ERR_NOT_READY, "%s: the sharded frame failed on rank %u (peer error 0x%08x: kind %u, rank %u, frame %u mod 2^20); hvcf_peer_status has the details",
			what, ctx->peer.rank, err, err >> 28, (err >> 20) & 0xffu, err & 0xfffffu);
	return HVCF_OK;
}

static int download(hvcf_ctx* ctx, void* dst, const void* src, size_t bytes, size_t limit, void* stream, const char* what)
{
	if (!ctx || (bytes && !dst))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: null argument", what);
	if (bytes > limit)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: %zu bytes > buffer size %zu", what, bytes, limit);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	if (bytes)
		HVCF_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s));
	HVCF_CUDA(cudaStreamSynchronize(s));
	return hvcf_peer_error_after_sync(ctx, what);
}

int hvcf_download_aabbs(hvcf_ctx* ctx, void* dst, size_t bytes, void* stream)
{
	return download(ctx, dst, ctx ? ctx->d_aabb : nullptr, bytes, ctx ? sizeof(hvcf_aabb) * ctx->cluster_total : 0, stream,
		"hvcf_download_aabbs");
}
int hvcf_download_light_cells(hvcf_ctx* ctx, void* dst, size_t bytes, void* stream)
{
	return download(ctx, dst, ctx ? ctx->d_cells : nullptr, bytes, ctx ? sizeof(hvcf_light_cell) * ctx->cluster_total : 0, stream,
		"hvcf_download_light_cells");
}
int hvcf_download_light_indices(hvcf_ctx* ctx, void* dst, size_t bytes, void* stream)
{
	return download(ctx, dst, ctx ? ctx->d_indices : nullptr, bytes,
		ctx ? sizeof(uint32_t) * static_cast<size_t>(ctx->cluster_total) * ctx->cap : 0, stream, "hvcf_download_light_indices");
}
int hvcf_download_total(hvcf_ctx* ctx, uint32_t frame, uint32_t* out_total, void* stream)
{
	if (ctx && frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_download_total: frame %u >= frame_count %u", frame, ctx->frame_count);
	return download(ctx, out_total, ctx ? &ctx->d_global_index[frame] : nullptr, sizeof(uint32_t), sizeof(uint32_t), stream,
		"hvcf_download_total");
}

// is p pinned host memory the device can store to at the same address?  Asked on every call (well under a microsecond): a
// remembered answer would go stale when the caller frees a pinned buffer and pageable memory takes its address.
static bool mapped_host_ptr(const void* p)
{
	if (!p || (reinterpret_cast<uintptr_t>(p) & 15u))
		return false;
	cudaPointerAttributes a{};
	const bool ok = cudaPointerGetAttributes(&a, p) == cudaSuccess && a.type == cudaMemoryTypeHost && a.devicePointer == p;
	cudaGetLastError(); // unregistered memory is not an error here
	return ok;
}

int hvcf_download_grid(hvcf_ctx* ctx, uint32_t frame, void* cells_dst, size_t cells_bytes, void* indices_dst,
	size_t indices_capacity_bytes, uint32_t* out_total, void* stream)
{
	if (!ctx || !out_total || (cells_bytes && !cells_dst))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_download_grid: null argument");
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_download_grid: frame %u >= frame_count %u", frame, ctx->frame_count);
	if (cells_bytes > sizeof(hvcf_light_cell) * static_cast<size_t>(ctx->cluster_total))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_download_grid: %zu bytes of cells > buffer size", cells_bytes);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	if (!ctx->h_total)
		HVCF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&ctx->h_total), sizeof(uint32_t) * 16));
	// both destinations in pinned (mapped) host memory, 16 B aligned: one kernel stores count, cells and indices[0, count)
	// across PCIe -- no host round trip between "how many" and "copy them" (HVCF_DOWNLOAD_KERNEL=0: the two-copy path)
	if (ctx->download_kernel && cells_bytes % 4 == 0 && mapped_host_ptr(cells_dst) && (indices_capacity_bytes == 0 || mapped_host_ptr(indices_dst)))
	{
		const size_t cap_words = indices_capacity_bytes / sizeof(uint32_t);
		HVCF_CUDA(launch_grid_download(ctx->d_cells, cells_bytes, ctx->d_indices, &ctx->d_global_index[frame],
			static_cast<uint32_t>(cap_words < 0xffffffffull ? cap_words : 0xffffffffull), ctx->download_max_words, cells_dst, static_cast<uint32_t*>(indices_dst),
			&ctx->h_total[frame], ctx->download_blocks, s));
		HVCF_CUDA(cudaStreamSynchronize(s));
		if (const int prc = hvcf_peer_error_after_sync(ctx, "hvcf_download_grid"))
			return prc;
		const uint32_t total = ctx->h_total[frame];
		*out_total = total;
		if (sizeof(uint32_t) * static_cast<size_t>(total) > indices_capacity_bytes)
			return fail(HVCF_ERR_CAPACITY, "hvcf_download_grid: %u indices do not fit in %zu bytes", total, indices_capacity_bytes);
		const uint32_t by_kernel = ctx->download_max_words & ~3u;
		if (total > by_kernel) // a long list: the copy engine takes what the kernel left
		{
			HVCF_CUDA(cudaMemcpyAsync(static_cast<uint32_t*>(indices_dst) + by_kernel, ctx->d_indices + by_kernel,
				sizeof(uint32_t) * static_cast<size_t>(total - by_kernel), cudaMemcpyDeviceToHost, s));
			HVCF_CUDA(cudaStreamSynchronize(s));
		}
		return HVCF_OK;
	}
	HVCF_CUDA(cudaMemcpyAsync(&ctx->h_total[frame], &ctx->d_global_index[frame], sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	if (cells_bytes)
		HVCF_CUDA(cudaMemcpyAsync(cells_dst, ctx->d_cells, cells_bytes, cudaMemcpyDeviceToHost, s));
	HVCF_CUDA(cudaStreamSynchronize(s));
	if (const int prc = hvcf_peer_error_after_sync(ctx, "hvcf_download_grid"))
		return prc;
	const uint32_t total = ctx->h_total[frame];
	*out_total = total;
	const size_t need = sizeof(uint32_t) * static_cast<size_t>(total);
	if (need > indices_capacity_bytes || (need && !indices_dst))
		return fail(HVCF_ERR_CAPACITY, "hvcf_download_grid: %u indices do not fit in %zu bytes", total, indices_capacity_bytes);
	if (need)
	{
		HVCF_CUDA(cudaMemcpyAsync(indices_dst, ctx->d_indices, need, cudaMemcpyDeviceToHost, s));
		HVCF_CUDA(cudaStreamSynchronize(s));
	}
	return HVCF_OK;
}

int hvcf_get_stats(hvcf_ctx* ctx, uint32_t frame, hvcf_stats* out, void* stream)
{
	if (!ctx || !out)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_get_stats: null argument");
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_get_stats: frame %u >= frame_count %u", frame, ctx->frame_count);
	DeviceStats ds;
	const int rc = download(ctx, &ds, &ctx->d_stats[frame], sizeof(ds), sizeof(ds), stream, "hvcf_get_stats");
	if (rc != HVCF_OK)
		return rc;
	memset(out, 0, sizeof(*out));
	out->nominal_tests = static_cast<uint64_t>(ctx->cluster_count) * ctx->num_lights;
	out->executed_tests = ds.executed_tests;
	out->prefilter_tests = ds.prefilter_tests;
	out->bin_tests = ds.bin_tests;
	out->write_pass_tests = ds.write_tests;
	{
		const size_t C = ctx->cluster_total;
		const uint64_t ref_bytes = sizeof(hvcf_light_data) * static_cast<uint64_t>(ctx->max_lights) + (sizeof(hvcf_aabb) + sizeof(hvcf_light_cell)) * C +
			sizeof(uint32_t) * C * ctx->cap + (sizeof(uint32_t) + sizeof(hvcf_cluster_forward_ubo)) * ctx->frame_count;
		out->device_bytes = ctx->device_bytes;
		out->scratch_bytes = ctx->device_bytes > ref_bytes ? ctx->device_bytes - ref_bytes : 0;
	}
	if (ctx->hier)
	{
		out->plan_groups = ctx->hp.ncb;
		out->plan_blocks_per_program = ctx->hp.K;
		out->plan_segments = ctx->h_num_segs;
		out->plan_rounds = ctx->h_num_rounds >= 2 ? ctx->h_num_rounds : 1;
		out->plan_flags = ctx->last_plan_flags;
	}
	out->total_indices = ds.total_indices;
	out->overflow_clusters = ds.overflow_clusters;
	out->kernel_launches = ctx->last_launches[frame];
	return HVCF_OK;
}

int hvcf_debug_trace(hvcf_ctx* ctx, unsigned long long* out_words, size_t max_words, size_t* out_ctas_per_pass)
{
	if (!ctx || !out_words || !out_ctas_per_pass)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_debug_trace: null argument");
	if (!ctx->d_trace)
		return fail(HVCF_ERR_NOT_READY, "hvcf_debug_trace: create the context with HVCF_HIER_TRACE=1 (hierarchical mode)");
	const size_t ctas = static_cast<size_t>(ctx->hp.ncb) * 2 * ctx->h_num_segs;
	const size_t words = ctas * 8 < max_words ? ctas * 8 : max_words;
	HVCF_CUDA(cudaSetDevice(ctx->device));
	HVCF_CUDA(cudaDeviceSynchronize());
	HVCF_CUDA(cudaMemcpy(out_words, ctx->d_trace, words * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
	*out_ctas_per_pass = ctas;
	return HVCF_OK;
}

int hvcf_debug_kernel_times(hvcf_ctx* ctx, uint32_t frame, float out_ms[5])
{
	if (!ctx || !out_ms || frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_debug_kernel_times: bad argument");
	if (!ctx->ktimes || !ctx->kev[frame][5])
		return fail(HVCF_ERR_NOT_READY, "hvcf_debug_kernel_times: needs HVCF_KERNEL_TIMES=1 at hvcf_create and a hierarchical hvcf_cull on this frame");
	HVCF_CUDA(cudaEventSynchronize(ctx->kev[frame][5]));
	for (int i = 0; i < 5; ++i)
		HVCF_CUDA(cudaEventElapsedTime(&out_ms[i], ctx->kev[frame][i], ctx->kev[frame][i + 1]));
	if (ctx->h_num_rounds >= 2 && ctx->kev_round[frame][0][2]) // rounds: bin = the rounds' alive + binning launches, count = their count passes
	{
		out_ms[1] = out_ms[2] = 0.0f;
		for (uint32_t r = 0; r < ctx->h_num_rounds; ++r)
		{
			float a = 0.0f, b = 0.0f;
			HVCF_CUDA(cudaEventElapsedTime(&a, ctx->kev_round[frame][r][0], ctx->kev_round[frame][r][1]));
			HVCF_CUDA(cudaEventElapsedTime(&b, ctx->kev_round[frame][r][1], ctx->kev_round[frame][r][2]));
			out_ms[1] += a, out_ms[2] += b;
		}
		HVCF_CUDA(cudaEventElapsedTime(&out_ms[3], ctx->kev_round[frame][ctx->h_num_rounds - 1][2], ctx->kev[frame][4]));
		if (ctx->routed_frame && ctx->kev_phase2[frame][1]) // the second route of a two-phase frame sits inside a round's binning interval
		{
			float ph = 0.0f;
			if (cudaEventElapsedTime(&ph, ctx->kev_phase2[frame][0], ctx->kev_phase2[frame][1]) == cudaSuccess && ph > 0.0f && ph < out_ms[1])
				out_ms[0] += ph, out_ms[1] -= ph;
			(void)cudaGetLastError();
		}
	}
	if (ctx->kev_exchange[frame]) // sharded frames: the wait for the other ranks' totals sits between finalize and the write pass
		HVCF_CUDA(cudaEventElapsedTime(&out_ms[4], ctx->kev[frame][6], ctx->kev[frame][5]));
	return HVCF_OK;
}

int hvcf_debug_rounds(hvcf_ctx* ctx, uint32_t plan[68], uint32_t what, void* dst, size_t bytes)
{
	if (!ctx || !plan)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_debug_rounds: null argument");
	if (!ctx->hier)
		return fail(HVCF_ERR_NOT_READY, "hvcf_debug_rounds: hierarchical mode only");
	memset(plan, 0, sizeof(uint32_t) * 68);
	plan[0] = ctx->hp.ncb, plan[1] = ctx->hp.K, plan[2] = ctx->tiles_cap, plan[3] = ctx->h_num_rounds;
	for (uint32_t r = 0; r < ctx->h_num_rounds && r < kHMaxChunks; ++r)
	{
		const hvcf_ctx::Round& rd = ctx->h_rounds[r];
		uint32_t* o = plan + 4 + 8 * r;
		o[0] = rd.seg_begin, o[1] = rd.seg_end, o[2] = rd.tile_begin, o[3] = rd.tile_end, o[4] = rd.tiles_per_seg, o[5] = rd.pipe, o[6] = rd.macro_shift;
	}
	if (!dst)
		return HVCF_OK;
	HVCF_CUDA(cudaSetDevice(ctx->device));
	HVCF_CUDA(cudaDeviceSynchronize());
	const void* src = nullptr;
	size_t have = 0;
	switch (what)
	{
	case 0: src = ctx->d_group_alive, have = sizeof(uint32_t) * ctx->hp.ncb * kHMaxChunks; break;
	case 1: src = ctx->d_bin_count, have = sizeof(uint32_t) * static_cast<size_t>(ctx->hp.ncb) * ctx->tiles_cap; break;
	case 2: src = ctx->d_alive_union, have = sizeof(float) * 8 * ctx->hp.ncb; break;
	default: return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_debug_rounds: what = %u", what);
	}
	if (bytes > have)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_debug_rounds: %zu bytes > %zu", bytes, have);
	HVCF_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
	return HVCF_OK;
}

int hvcf_set_timing(hvcf_ctx* ctx, int enabled)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_timing: null context");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	if (enabled)
		for (uint32_t f = 0; f < ctx->frame_count; ++f)
			for (auto& e : ctx->ev[f])
				if (!e)
					HVCF_CUDA(cudaEventCreate(&e));
	ctx->timing = enabled != 0;
	return HVCF_OK;
}

int hvcf_get_timing(hvcf_ctx* ctx, uint32_t frame, float* out_cull_kernel_ms, float* out_total_ms)
{
	if (!ctx || !out_cull_kernel_ms || !out_total_ms)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_get_timing: null argument");
	if (frame >= ctx->frame_count || !ctx->timing)
		return fail(HVCF_ERR_NOT_READY, "hvcf_get_timing: timing disabled or bad frame");
	HVCF_CUDA(cudaEventSynchronize(ctx->ev[frame][3]));
	*out_cull_kernel_ms = 0.0f;
	if (ctx->ev_valid[frame])
		HVCF_CUDA(cudaEventElapsedTime(out_cull_kernel_ms, ctx->ev[frame][1], ctx->ev[frame][2]));
	HVCF_CUDA(cudaEventElapsedTime(out_total_ms, ctx->ev[frame][0], ctx->ev[frame][3]));
	return HVCF_OK;
}

int hvcf_device_alloc(int device, size_t bytes, void** out_device_ptr)
{
	if (!out_device_ptr)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_device_alloc: null argument");
	if (device >= 0)
		HVCF_CUDA(cudaSetDevice(device));
	HVCF_CUDA(cudaMalloc(out_device_ptr, bytes ? bytes : 16));
	return HVCF_OK;
}

int hvcf_device_free(void* device_ptr)
{
	if (device_ptr)
		HVCF_CUDA(cudaFree(device_ptr));
	return HVCF_OK;
}

int hvcf_device_upload(void* device_dst, size_t offset, const void* host_src, size_t bytes, void* stream)
{
	if (!device_dst || (bytes && !host_src))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_device_upload: null argument");
	if (bytes)
		HVCF_CUDA(cudaMemcpyAsync(static_cast<char*>(device_dst) + offset, host_src, bytes, cudaMemcpyHostToDevice,
			static_cast<cudaStream_t>(stream)));
	return HVCF_OK;
}

int hvcf_device_download(void* host_dst, const void* device_src, size_t offset, size_t bytes, void* stream)
{
	if (!device_src || (bytes && !host_dst))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_device_download: null argument");
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	if (bytes)
		HVCF_CUDA(cudaMemcpyAsync(host_dst, static_cast<const char*>(device_src) + offset, bytes, cudaMemcpyDeviceToHost, s));
	HVCF_CUDA(cudaStreamSynchronize(s));
	return HVCF_OK;
}

int hvcf_stream_synchronize(void* stream)
{
	HVCF_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
	return HVCF_OK;
}

int hvcf_host_alloc(size_t bytes, void** out_ptr)
{
	if (!out_ptr)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_host_alloc: null argument");
	HVCF_CUDA(cudaMallocHost(out_ptr, bytes ? bytes : 16));
	return HVCF_OK;
}

int hvcf_host_free(void* ptr)
{
	if (ptr)
		HVCF_CUDA(cudaFreeHost(ptr));
	return HVCF_OK;
}

} // extern "C"
