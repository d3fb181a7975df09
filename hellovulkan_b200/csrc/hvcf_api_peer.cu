// hvcf_api_peer.cu -- C ABI of the exchange step of the sharded grid over peer memory (include/hvcf.h: hvcf_peer_*,
// hvcf_cull_sharded; kernels in hvcf_peer.cu).  The reference is single-GPU: what this replaces, at scale, is the one dispatch of
// PipelineLightCulling::Execute (Source/Pipelines/PipelineLightCulling.cpp:58-119); SURVEY.md 8(e).
#include "hvcf_ctx.cuh"

#include <cstring>

using namespace hvcf;

void peer_unmap(hvcf_ctx* c)
{
	if (c->peer_ipc)
		for (uint32_t r = 0; r < c->peer.world; ++r)
			if (r != c->peer.rank)
			{
				cudaIpcCloseMemHandle(c->peer.cells[r]);
				cudaIpcCloseMemHandle(c->peer.indices[r]);
				cudaIpcCloseMemHandle(c->peer.mail[r]);
				if (c->shard.inbox_peer[r])
					cudaIpcCloseMemHandle(c->shard.inbox_peer[r]);
			}
	for (auto& q : c->shard.inbox_peer)
		q = nullptr;
	c->peer = PeerParams{};
	c->peer_ipc = false;
}

extern "C" {

// ---- exchange step of the sharded grid over peer memory ---------------------------------------------------
static int peer_alloc(hvcf_ctx* ctx)
{
	if (ctx->d_mail && ctx->d_index_base && ctx->d_peer_totals)
		return HVCF_OK; // nothing to do -- and no synchronisation: getters may be called between stream-ordered peer calls
	HVCF_CUDA(cudaSetDevice(ctx->device));
	if (!ctx->d_mail)
	{
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_mail), sizeof(PeerMailbox)));
		HVCF_CUDA(cudaMemset(ctx->d_mail, 0, sizeof(PeerMailbox)));
	}
	if (!ctx->d_index_base)
	{
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_index_base), sizeof(uint32_t)));
		HVCF_CUDA(cudaMemset(ctx->d_index_base, 0, sizeof(uint32_t)));
	}
	if (!ctx->d_peer_totals)
	{
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_peer_totals), sizeof(uint32_t) * kMaxPeers));
		HVCF_CUDA(cudaMemset(ctx->d_peer_totals, 0, sizeof(uint32_t) * kMaxPeers));
	}
	HVCF_CUDA(cudaDeviceSynchronize());
	return HVCF_OK;
}

int hvcf_peer_export(hvcf_ctx* ctx, hvcf_peer_handle* out)
{
	static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(hvcf_peer_handle) == 256, "hvcf_peer_handle holds four 64-byte IPC handles");
	static_assert(HVCF_MAX_PEERS == kMaxPeers, "HVCF_MAX_PEERS");
	if (!ctx || !out)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_export: null argument");
	const int rc = peer_alloc(ctx);
	if (rc != HVCF_OK)
		return rc;
	HVCF_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->cells), ctx->d_cells));
	HVCF_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->indices), ctx->d_indices));
	HVCF_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->mailbox), ctx->d_mail));
	memset(out->inbox, 0, sizeof(out->inbox));
	if (ctx->shard.on)
		HVCF_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->inbox), ctx->shard.d_inbox));
	return HVCF_OK;
}

static int peer_check_attach(hvcf_ctx* ctx, uint32_t rank, uint32_t world, const void* p, const char* who)
{
	if (!ctx || !p)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: null argument", who);
	if (world == 0 || world > kMaxPeers || rank >= world)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: rank %u of world %u (at most %u peers)", who, rank, world, kMaxPeers);
	if (ctx->peer.world)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: already attached; call hvcf_peer_detach first", who);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	HVCF_CUDA(preload_basic_kernels());
	HVCF_CUDA(preload_hier_kernels());
	HVCF_CUDA(preload_bin_kernels());
	HVCF_CUDA(preload_small_kernels());
	HVCF_CUDA(preload_peer_kernels());
	HVCF_CUDA(preload_shard_kernels());
	if (ctx->shard.on && (ctx->shard.world != world || ctx->shard.rank != rank))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: rank %u of %u, but the context was created as rank %u of %u", who, rank, world, ctx->shard.rank,
			ctx->shard.world);
	return peer_alloc(ctx);
}

int hvcf_peer_attach(hvcf_ctx* ctx, uint32_t rank, uint32_t world, const hvcf_peer_handle* handles)
{
	const int rc = peer_check_attach(ctx, rank, world, handles, "hvcf_peer_attach");
	if (rc != HVCF_OK)
		return rc;
	PeerParams pp{};
	pp.world = world;
	pp.rank = rank;
	for (uint32_t r = 0; r < world; ++r)
	{
		if (r == rank)
		{
			pp.cells[r] = ctx->d_cells;
			pp.indices[r] = ctx->d_indices;
			pp.mail[r] = ctx->d_mail;
			continue;
		}
		void* ptr[4] = {nullptr, nullptr, nullptr, nullptr};
		const unsigned char* h[4] = {handles[r].cells, handles[r].indices, handles[r].mailbox, handles[r].inbox};
		const int nh = ctx->shard.on ? 4 : 3;
		for (int k = 0; k < nh; ++k)
		{
			cudaIpcMemHandle_t handle;
			memcpy(&handle, h[k], sizeof(handle));
			const cudaError_t e = cudaIpcOpenMemHandle(&ptr[k], handle, cudaIpcMemLazyEnablePeerAccess);
			if (e != cudaSuccess)
			{
				for (int j = 0; j < k; ++j)
					cudaIpcCloseMemHandle(ptr[j]);
				for (uint32_t q = 0; q < r; ++q) // the ranks mapped so far
					if (q != rank)
					{
						cudaIpcCloseMemHandle(pp.cells[q]);
						cudaIpcCloseMemHandle(pp.indices[q]);
						cudaIpcCloseMemHandle(pp.mail[q]);
						if (ctx->shard.inbox_peer[q])
							cudaIpcCloseMemHandle(ctx->shard.inbox_peer[q]);
						ctx->shard.inbox_peer[q] = nullptr;
					}
				return fail(HVCF_ERR_CUDA, "hvcf_peer_attach: cudaIpcOpenMemHandle(rank %u, buffer %d): %s", r, k, cudaGetErrorString(e));
			}
		}
		pp.cells[r] = static_cast<hvcf_light_cell*>(ptr[0]);
		pp.indices[r] = static_cast<uint32_t*>(ptr[1]);
		pp.mail[r] = static_cast<PeerMailbox*>(ptr[2]);
		ctx->shard.inbox_peer[r] = static_cast<unsigned char*>(ptr[3]);
	}
	if (ctx->shard.on)
		ctx->shard.inbox_peer[rank] = ctx->shard.d_inbox;
	ctx->peer = pp;
	ctx->peer_ipc = true;
	ctx->peer_epoch = 0;
	ctx->shard.epoch = 0;
	return HVCF_OK;
}

int hvcf_peer_attach_local(hvcf_ctx* ctx, uint32_t rank, uint32_t world, hvcf_ctx* const* contexts)
{
	const int rc = peer_check_attach(ctx, rank, world, contexts, "hvcf_peer_attach_local");
	if (rc != HVCF_OK)
		return rc;
	PeerParams pp{};
	pp.world = world;
	pp.rank = rank;
	for (uint32_t r = 0; r < world; ++r)
	{
		hvcf_ctx* o = r == rank ? ctx : contexts[r];
		if (!o)
			return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_attach_local: contexts[%u] is null", r);
		if (o->cluster_total != ctx->cluster_total || o->cap != ctx->cap)
			return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_attach_local: contexts[%u] has a different grid or cap", r);
		const int rc2 = peer_alloc(o);
		if (rc2 != HVCF_OK)
			return rc2;
		if (o->device != ctx->device)
		{
			int can = 0;
			HVCF_CUDA(cudaDeviceCanAccessPeer(&can, ctx->device, o->device));
			if (!can)
				return fail(HVCF_ERR_UNSUPPORTED, "hvcf_peer_attach_local: device %d cannot access device %d", ctx->device, o->device);
			HVCF_CUDA(cudaSetDevice(ctx->device));
			const cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0);
			if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
				return fail(HVCF_ERR_CUDA, "hvcf_peer_attach_local: cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
			(void)cudaGetLastError();
		}
		pp.cells[r] = o->d_cells;
		pp.indices[r] = o->d_indices;
		pp.mail[r] = o->d_mail;
		if (ctx->shard.on)
		{
			if (!o->shard.on || o->shard.parity_bytes != ctx->shard.parity_bytes || o->shard.world != world)
				return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_attach_local: contexts[%u] was not created with the same hvcf_shard_desc", r);
			ctx->shard.inbox_peer[r] = o->shard.d_inbox;
		}
	}
	ctx->peer = pp;
	ctx->peer_ipc = false;
	ctx->peer_epoch = 0;
	ctx->shard.epoch = 0;
	return HVCF_OK;
}

int hvcf_peer_detach(hvcf_ctx* ctx)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_detach: null context");
	cudaSetDevice(ctx->device);
	cudaDeviceSynchronize();
	peer_unmap(ctx);
	if (ctx->d_mail)
		cudaMemset(ctx->d_mail, 0, sizeof(PeerMailbox));
	return HVCF_OK;
}

int hvcf_peer_totals_step(hvcf_ctx* ctx, void* stream, uint32_t what)
{
	if (!ctx || !ctx->peer.world)
		return fail(HVCF_ERR_NOT_READY, "hvcf_peer_exchange_totals: no peers attached");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	if (what & 1u) // a new exchange starts with the post
	{
		ctx->peer_epoch = ctx->peer_epoch >= 0x7fffffffu ? 1u : ctx->peer_epoch + 1u;
		ctx->peer.epoch = ctx->peer_epoch;
	}
	HVCF_CUDA(launch_peer_totals(ctx->peer, ctx->d_shard_total, ctx->d_index_base, ctx->d_peer_totals, static_cast<cudaStream_t>(stream), what));
	if (what == 3u && ctx->ktimes && ctx->range_frame < ctx->frame_count) // (posted early / waited late: the write pass does not wait)
	{
		HVCF_CUDA(hvcf_kmark(ctx, ctx->range_frame, 6, static_cast<cudaStream_t>(stream)));
		ctx->kev_exchange[ctx->range_frame] = true;
	}
	return HVCF_OK;
}

int hvcf_peer_exchange_totals(hvcf_ctx* ctx, void* stream) { return hvcf_peer_totals_step(ctx, stream, 3u); }

int hvcf_peer_push_grid(hvcf_ctx* ctx, void* stream)
{
	if (!ctx || !ctx->peer.world || !ctx->peer.epoch)
		return fail(HVCF_ERR_NOT_READY, "hvcf_peer_push_grid: no peers attached or no totals exchanged");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	// enough CTAs to keep the NVLink ports busy: 4 per SM by default (HVCF_PEER_BLOCKS overrides, for sweeps)
	static const uint32_t blocks = static_cast<uint32_t>(env_int("HVCF_PEER_BLOCKS", 592) > 0 ? env_int("HVCF_PEER_BLOCKS", 592) : 592);
	HVCF_CUDA(launch_peer_push(ctx->peer, ctx->d_shard_total, ctx->d_index_base, ctx->cluster_begin, ctx->cluster_count, blocks,
		static_cast<cudaStream_t>(stream)));
	return HVCF_OK;
}

int hvcf_peer_wait_grid(hvcf_ctx* ctx, void* stream)
{
	if (!ctx || !ctx->peer.world || !ctx->peer.epoch)
		return fail(HVCF_ERR_NOT_READY, "hvcf_peer_wait_grid: no peers attached or no totals exchanged");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	HVCF_CUDA(launch_peer_wait(ctx->peer, ctx->d_peer_totals, &ctx->d_global_index[ctx->range_frame < ctx->frame_count ? ctx->range_frame : 0],
		static_cast<cudaStream_t>(stream)));
	return HVCF_OK;
}

int hvcf_cull_sharded(hvcf_ctx* ctx, uint32_t frame, void* stream)
{
	if (!ctx || !ctx->peer.world)
		return fail(HVCF_ERR_NOT_READY, "hvcf_cull_sharded: no peers attached (hvcf_peer_attach)");
	int rc = hvcf_cull_begin(ctx, frame, stream);
	if (rc == HVCF_OK)
		rc = hvcf_peer_exchange_totals(ctx, stream);
	if (rc == HVCF_OK)
		rc = hvcf_cull_finish(ctx, frame, ctx->d_index_base, stream);
	if (rc == HVCF_OK)
		rc = hvcf_peer_push_grid(ctx, stream);
	if (rc == HVCF_OK)
		rc = hvcf_peer_wait_grid(ctx, stream);
	return rc;
}

int hvcf_peer_totals_ptr(hvcf_ctx* ctx, uint32_t** out_device_ptr)
{
	if (!ctx || !out_device_ptr)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_totals_ptr: null argument");
	const int rc = peer_alloc(ctx);
	if (rc != HVCF_OK)
		return rc;
	*out_device_ptr = ctx->d_peer_totals;
	return HVCF_OK;
}

int hvcf_peer_index_base_ptr(hvcf_ctx* ctx, uint32_t** out_device_ptr)
{
	if (!ctx || !out_device_ptr)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_index_base_ptr: null argument");
	const int rc = peer_alloc(ctx);
	if (rc != HVCF_OK)
		return rc;
	*out_device_ptr = ctx->d_index_base;
	return HVCF_OK;
}

int hvcf_peer_status(hvcf_ctx* ctx, void* stream, int* out_error)
{
	if (!ctx || !out_error)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_peer_status: null argument");
	*out_error = 0;
	if (!ctx->d_mail)
		return HVCF_OK;
	HVCF_CUDA(cudaSetDevice(ctx->device));
	HVCF_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
	unsigned int err = 0;
	HVCF_CUDA(cudaMemcpy(&err, &ctx->d_mail->error, sizeof(err), cudaMemcpyDeviceToHost));
	*out_error = static_cast<int>(err);
	if (err)
	{
		static const char* const kinds[] = {"?", "index total", "grid segment", "routed lights", "error report"};
		const unsigned int kind = err >> 28;
		return fail(HVCF_ERR_NOT_READY, "hvcf_peer_status: rank %u %s the %s of rank %u in sharded frame %u (mod 2^20)", ctx->peer.rank,
			kind == 4 ? "received" : "waited in vain for", kinds[kind < 5 ? kind : 0], (err >> 20) & 0xffu, err & 0xfffffu);
	}
	return HVCF_OK;
}

} // extern "C"
