// hvcf_shard_api.cu -- C ABI of the sharded frame with the lights sharded too (include/hvcf.h, "hvcf_cull_routed";
// kernels in hvcf_shard.cu).  The reference is single-GPU: what this replaces, at scale, is the one dispatch of
// PipelineLightCulling::Execute (Source/Pipelines/PipelineLightCulling.cpp:58-119).
#include "hvcf_ctx.cuh"

#include <cstring>

using namespace hvcf;

namespace
{
uint32_t default_block_tiles(uint32_t world, uint32_t total_lights)
{
	// big frames: 16 tiles (the largest cluster of CTAs).  The late rounds of a routed frame keep one survivor slot per group and
	// route block, and the records of a 16,384-light block fill whole inbox tiles where those of a 2,048-light block leave most
	// of every tile empty (2 GPUs, 16M lights: 1.52 -> 1.28 ms per frame, profiles/r03b)
	if (total_lights >= (1u << 20))
		return kMaxRouteBlockTiles;
	// a route block is compacted by one CTA per destination, so a block of B tiles arrives as about B x (share of the lights that
	// touch the slab) dense inbox tiles: B = world keeps that near two tiles for the reference's light distribution
	uint32_t b = 1;
	while (b < world && b < 8u)
		b <<= 1;
	return b;
}
} // namespace

extern "C" {

void hvcf_shard_free(hvcf_ctx* c)
{
	cudaFree(c->shard.d_inbox);
	cudaFree(c->shard.d_ranges);
	cudaFree(c->shard.d_slab_box);
	cudaFree(c->shard.d_tile_cnt);
	cudaFree(c->shard.d_tile_ids);
	cudaFree(c->shard.d_seg);
	cudaFree(c->shard.d_cnt_last);
	c->shard = ShardState{};
}

int hvcf_create_sharded(const hvcf_config* config, const hvcf_shard_desc* desc, hvcf_ctx** out_ctx)
{
	if (!config || !desc || !out_ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: null argument");
	if (desc->struct_size != sizeof(hvcf_shard_desc))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: struct_size %u != %zu", desc->struct_size, sizeof(hvcf_shard_desc));
	*out_ctx = nullptr;
	const uint32_t world = desc->world, rank = desc->rank;
	if (world == 0 || world > HVCF_MAX_PEERS || rank >= world)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: rank %u of world %u (at most %u ranks)", rank, world, HVCF_MAX_PEERS);
	const uint32_t B = desc->block_tiles ? desc->block_tiles : default_block_tiles(world, desc->total_lights);
	if (B > kMaxRouteBlockTiles || (B & (B - 1u)))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: block_tiles %u is not a power of two <= %u", B, kMaxRouteBlockTiles);
	const uint32_t Z = config->slice_count_z ? config->slice_count_z : HVCF_DEFAULT_SLICE_COUNT_Z;
	const uint32_t Yc = config->slice_count_y ? config->slice_count_y : HVCF_DEFAULT_SLICE_COUNT_Y;
	// the cuts in cluster rows (row = y + z * Y): row_cuts if given, else the z cuts
	uint32_t rows[HVCF_MAX_PEERS + 1];
	const bool by_rows = desc->row_cuts[world] != 0;
	for (uint32_t r = 0; r <= world; ++r)
		rows[r] = by_rows ? desc->row_cuts[r] : desc->z_cuts[r] * Yc;
	if (rows[0] != 0 || rows[world] != Yc * Z)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: the cuts must run from 0 to %s = %u", by_rows ? "slice_count_y x slice_count_z" : "slice_count_z",
			by_rows ? Yc * Z : Z);
	for (uint32_t r = 0; r < world; ++r)
		if (rows[r] > rows[r + 1])
			return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: the cuts must ascend");
	if (rows[rank] == rows[rank + 1])
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: rank %u owns no cluster", rank);
	if (config->mode == HVCF_MODE_BRUTE)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_create_sharded: hierarchical mode only");

	// ownership plan: whole route blocks, the same number on every rank (the last ranks may own fewer lights, or none)
	const uint64_t n = desc->total_lights;
	const uint64_t tiles = (n + kTileLights - 1) / kTileLights;
	const uint64_t blocks_total = tiles ? (tiles + B - 1) / B : 1;
	const uint64_t bpr = (blocks_total + world - 1) / world;
	const uint64_t blocks_phys = bpr * world, tiles_phys = blocks_phys * B;
	if (tiles_phys * kTileLights > 0xffff0000ull)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create_sharded: %u lights are too many", desc->total_lights);

	hvcf_config cfg = *config;
	cfg.z_begin = cfg.z_end = 0;
	cfg.row_begin = rows[rank];
	cfg.row_end = rows[rank + 1];
	cfg.mode = HVCF_MODE_HIERARCHICAL;
	const uint32_t need = static_cast<uint32_t>(tiles_phys * kTileLights);
	if (cfg.max_lights < need)
		cfg.max_lights = need; // the survivor slots are indexed by inbox tile
	hvcf_ctx* c = nullptr;
	int rc = hvcf_create(&cfg, &c);
	if (rc != HVCF_OK)
		return rc;
	if (!c->bin_ns)
	{
		hvcf_destroy(c);
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_create_sharded: routed frames bin with the two-level kernel (HVCF_BIN2=0 / HVCF_HIER_V1 switch it off)");
	}
	ShardState& sh = c->shard;
	sh.rank = rank, sh.world = world, sh.total_lights = desc->total_lights;
	sh.block_tiles = B;
	sh.two_phase = env_int("HVCF_ROUTE_TWO_PHASE", 1) != 0;
	sh.blocks_per_rank = static_cast<uint32_t>(bpr);
	sh.blocks_phys = static_cast<uint32_t>(blocks_phys);
	sh.tiles_phys = static_cast<uint32_t>(tiles_phys);
	memcpy(sh.row_cuts, rows, sizeof(uint32_t) * (world + 1));
	sh.off_mask = static_cast<size_t>(tiles_phys) * kTileBytes;
	sh.off_cnt = sh.off_mask + static_cast<size_t>(tiles_phys) * 32 * sizeof(uint32_t);
	sh.parity_bytes = (sh.off_cnt + static_cast<size_t>(blocks_phys) * sizeof(uint32_t) + 255u) & ~static_cast<size_t>(255u);
	uint32_t ranges[2 * HVCF_MAX_PEERS];
	for (uint32_t r = 0; r < world; ++r)
	{
		ranges[2 * r] = rows[r] * c->X;
		ranges[2 * r + 1] = rows[r + 1] * c->X;
	}
	cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&sh.d_inbox), 2 * sh.parity_bytes);
	if (e == cudaSuccess)
		e = cudaMemset(sh.d_inbox, 0, 2 * sh.parity_bytes);
	if (e == cudaSuccess)
		e = cudaMalloc(reinterpret_cast<void**>(&sh.d_ranges), sizeof(ranges));
	if (e == cudaSuccess)
		e = cudaMemcpy(sh.d_ranges, ranges, sizeof(uint32_t) * 2 * world, cudaMemcpyHostToDevice);
	if (e == cudaSuccess)
		e = cudaMalloc(reinterpret_cast<void**>(&sh.d_slab_box), sizeof(float) * 8 * world);
	if (e == cudaSuccess)
		e = cudaMalloc(reinterpret_cast<void**>(&sh.d_tile_cnt), sizeof(uint32_t) * tiles_phys);
	if (e == cudaSuccess)
		e = cudaMemset(sh.d_tile_cnt, 0, sizeof(uint32_t) * tiles_phys);
	if (e == cudaSuccess)
		e = cudaMalloc(reinterpret_cast<void**>(&sh.d_seg), sizeof(uint32_t) * static_cast<size_t>(c->cluster_count) * c->cap + 64);
	if (e == cudaSuccess)
		e = cudaMalloc(reinterpret_cast<void**>(&sh.d_tile_ids), sizeof(uint16_t) * tiles_phys * kTileLights);
	if (e == cudaSuccess)
		e = cudaMalloc(reinterpret_cast<void**>(&sh.d_cnt_last), sizeof(uint32_t) * blocks_phys);
	if (e == cudaSuccess)
		e = cudaMemset(sh.d_cnt_last, 0, sizeof(uint32_t) * blocks_phys);
	if (e != cudaSuccess)
	{
		hvcf_destroy(c);
		return fail(e == cudaErrorMemoryAllocation ? HVCF_ERR_OUT_OF_MEMORY : HVCF_ERR_CUDA, "hvcf_create_sharded: inbox of %zu bytes: %s",
			2 * sh.parity_bytes, cudaGetErrorString(e));
	}
	c->device_bytes += 2 * sh.parity_bytes + sizeof(ranges) + sizeof(float) * 8 * world + sizeof(uint32_t) * tiles_phys +
		sizeof(uint32_t) * static_cast<size_t>(c->cluster_count) * c->cap + 64 + sizeof(uint16_t) * tiles_phys * kTileLights + sizeof(uint32_t) * blocks_phys;
	sh.on = true;
	c->num_lights = desc->total_lights; // the frame is over all the lights of the scene; this rank holds its own range of them
	c->aabb_dirty = true;
	*out_ctx = c;
	return HVCF_OK;
}

int hvcf_shard_light_range(hvcf_ctx* ctx, uint32_t* out_begin, uint32_t* out_end)
{
	if (!ctx || !out_begin || !out_end)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_shard_light_range: null argument");
	if (!ctx->shard.on)
		return fail(HVCF_ERR_NOT_READY, "hvcf_shard_light_range: not a context of hvcf_create_sharded");
	const ShardState& sh = ctx->shard;
	const uint64_t per = static_cast<uint64_t>(sh.blocks_per_rank) * sh.block_tiles * kTileLights;
	const uint64_t b = per * sh.rank, e = per * (sh.rank + 1);
	*out_begin = static_cast<uint32_t>(b < sh.total_lights ? b : sh.total_lights);
	*out_end = static_cast<uint32_t>(e < sh.total_lights ? e : sh.total_lights);
	return HVCF_OK;
}

int hvcf_set_lights_range(hvcf_ctx* ctx, const void* lights, uint32_t begin, uint32_t count, int src_is_device, void* stream)
{
	if (!ctx || (count && !lights))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_lights_range: null argument");
	if (static_cast<uint64_t>(begin) + count > ctx->max_lights)
		return fail(HVCF_ERR_CAPACITY, "hvcf_set_lights_range: [%u, %u + %u) beyond max_lights %u", begin, begin, count, ctx->max_lights);
	if (ctx->bound_lights)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_lights_range: an external light buffer is bound (hvcf_bind_lights); write to it directly");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	if (count)
		HVCF_CUDA(cudaMemcpyAsync(ctx->d_lights + begin, lights, sizeof(hvcf_light_data) * static_cast<size_t>(count),
			src_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
	if (!ctx->shard.on && begin + count > ctx->num_lights)
		ctx->num_lights = begin + count;
	return HVCF_OK;
}

int hvcf_cull_routed(hvcf_ctx* ctx, uint32_t frame, void* stream)
{
	if (!ctx || !ctx->shard.on)
		return fail(HVCF_ERR_NOT_READY, "hvcf_cull_routed: not a context of hvcf_create_sharded");
	ShardState& sh = ctx->shard;
	if (ctx->anim_len)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_cull_routed: device-side light animation is not routed (frame-partitioned replicas animate)");
	if (!ctx->peer.world)
	{
		if (sh.world != 1)
			return fail(HVCF_ERR_NOT_READY, "hvcf_cull_routed: no peers attached (hvcf_peer_attach)");
		hvcf_ctx* self[1] = {ctx};
		const int rc = hvcf_peer_attach_local(ctx, 0, 1, self);
		if (rc != HVCF_OK)
			return rc;
	}
	NvtxZone zone("Light_Culling");
	int rc = hvcf_cull_open(ctx, frame, stream, "hvcf_cull_routed", true, true);
	if (rc != HVCF_OK)
		return rc;
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	sh.epoch = sh.epoch >= 0x7fffffffu ? 1u : sh.epoch + 1u;
	const uint32_t parity = sh.epoch & 1u;
	DeviceStats* stats = &ctx->d_stats[frame];

	// ---- route: own lights -> view space -> every slab that they touch.  A frame that is counted in rounds is routed in TWO phases:
	// the lights of the first two rounds against the slabs' union boxes, and -- once those rounds are counted on every rank and every
	// rank has published the boxes of its clusters that are still below their cap (slab_alive_kernel) -- the rest against those.
	// A slab that is full receives nothing more; on the reference scene that is every slab in front of slice 64 for 97 % of the lights. ----
	uint32_t lb = 0, le = 0;
	hvcf_shard_light_range(ctx, &lb, &le);
	const uint32_t own0 = sh.rank * sh.blocks_per_rank, own1 = own0 + sh.blocks_per_rank; // own route blocks (global numbering)
	uint32_t phase_round = 0, head_blocks = sh.blocks_phys; // first round of the second phase; route blocks of the first phase
	if (ctx->h_num_rounds >= 3 && sh.two_phase && ctx->h_rounds[2].tile_begin % sh.block_tiles == 0)
	{
		phase_round = 2;
		head_blocks = ctx->h_rounds[2].tile_begin / sh.block_tiles;
	}
	auto route = [&](uint32_t b0, uint32_t b1, uint32_t phase) -> int { // own blocks [b0, b1) of the frame's phase
		if (b0 >= b1)
			return HVCF_OK; // nothing of this phase is ours; the wait kernel that follows raises the flag all the same
		RouteParams rp{};
		rp.ubo = &ctx->d_ubo[frame];
		// indexed by GLOBAL light index; a bound buffer (hvcf_bind_lights) starts at the first owned light, lb
		rp.lights = ctx->bound_lights ? reinterpret_cast<const hvcf_light_data*>(reinterpret_cast<uintptr_t>(ctx->bound_lights) -
			sizeof(hvcf_light_data) * static_cast<uintptr_t>(lb)) : ctx->d_lights;
		const uint64_t end_light = static_cast<uint64_t>(b1) * sh.block_tiles * kTileLights;
		rp.light_end = end_light < le ? static_cast<uint32_t>(end_light) : le;
		rp.block_tiles = sh.block_tiles;
		rp.block0 = b0;
		rp.blocks = b1 - b0;
		rp.world = sh.world, rp.rank = sh.rank, rp.epoch = sh.epoch, rp.parity = parity;
		rp.phase = phase;
		rp.boxes = phase == 0 ? 1u : kSlabBoxes;
		rp.slab_box = phase == 0 ? sh.d_slab_box : &ctx->d_mail->alive_box[parity][0][0][0];
		for (uint32_t r = 0; r < sh.world; ++r)
		{
			unsigned char* box = sh.inbox_peer[r] + static_cast<size_t>(parity) * sh.parity_bytes;
			rp.rec[r] = reinterpret_cast<float4*>(box);
			rp.mask[r] = reinterpret_cast<uint32_t*>(box + sh.off_mask);
			rp.cnt[r] = reinterpret_cast<uint32_t*>(box + sh.off_cnt);
			rp.mail[r] = ctx->peer.mail[r];
		}
		if (phase == 0)
		{
			rp.zero = ctx->d_hscratch; // ResetGlobalIndex (PipelineLightCulling.cpp:38-43) and the early-exit counters
			rp.zero_words = ctx->cluster_count * kHMaxChunks;
		}
		rp.stats = stats;
		HVCF_CUDA(launch_shard_route(rp, s));
		return HVCF_OK;
	};
	const uint32_t head_end = own1 < head_blocks ? own1 : (own0 > head_blocks ? own0 : head_blocks); // own blocks [own0, head_end) belong to the first phase
	if (own0 >= head_end) // nothing to route first: the frame's counters are cleared (and its serial bumped) by a clear-only launch
		HVCF_CUDA(launch_light_transform(&ctx->d_ubo[frame], nullptr, 0, 0, ctx->d_lights_view, ctx->d_hscratch, ctx->cluster_count * kHMaxChunks, stats, s, nullptr));
	rc = route(own0, head_end, 0u);
	if (rc != HVCF_OK)
		return rc;
	HVCF_CUDA(hvcf_kmark(ctx, frame, 7, s)); // HVCF_KERNEL_TIMES: the route kernel alone (hvcf_debug_route_ms)
	ctx->range_cleared = true;
	PeerParams pp = ctx->peer;
	pp.epoch = sh.epoch;
	HVCF_CUDA(launch_shard_wait_routed(pp, parity, s, 0u));
	HVCF_CUDA(hvcf_kmark(ctx, frame, 1, s)); // HVCF_KERNEL_TIMES: "transform" = route + wait for the other ranks' shares

	// ---- level 0 on the inbox: every tile against the cluster groups of this slab ----
	unsigned char* box = sh.d_inbox + static_cast<size_t>(parity) * sh.parity_bytes;
	BinParams b{};
	b.lights_view = reinterpret_cast<const float4*>(box);
	b.cb_union = ctx->d_cb_union;
	b.ncb = ctx->hp.ncb;
	b.tiles_cap = ctx->tiles_cap;
	b.bin_count = ctx->d_bin_count;
	b.bin_rec = ctx->d_bin_rec;
	b.bin_idx = ctx->d_bin_idx;
	b.stats = stats;
	b.tile_begin = 0;
	b.sb_union = ctx->d_sb_union;
	b.sg = ctx->bin_sg;
	b.ns = ctx->bin_ns;
	b.in_mask = reinterpret_cast<const uint32_t*>(box + sh.off_mask);
	b.tile_ids = sh.d_tile_ids;
	b.in_cnt = reinterpret_cast<const uint32_t*>(box + sh.off_cnt);
	b.block_tiles = sh.block_tiles;
	b.tile_cnt = sh.d_tile_cnt;
	ctx->open_launches += 2;
	if (ctx->h_num_rounds >= 2)
	{
		// the inbox is in light order, so it takes rounds like any other frame: each round's binning skips the groups of this
		// slab that the earlier rounds have filled (hier_alive_kernel, hvcf_bin.cu)
		const int rrc = hvcf_hier_rounds(ctx, frame, s, &ctx->open_launches, [&](uint32_t r) {
			const hvcf_ctx::Round& rd = ctx->h_rounds[r];
			if (phase_round && r == phase_round)
			{
				// between the phases (hier_alive_kernel of this round has just been launched): publish this slab's live boxes, wait
				// for everybody's, route the rest of the own lights against them, wait for everybody's second share
				if (ctx->ktimes)
				{
					for (auto& e : ctx->kev_phase2[frame])
						if (!e)
							HVCF_CUDA(cudaEventCreate(&e));
					HVCF_CUDA(cudaEventRecord(ctx->kev_phase2[frame][0], s));
				}
				ctx->last_plan_flags |= HVCF_PLAN_TWO_PHASE_ROUTE;
				HVCF_CUDA(launch_slab_alive(pp, parity, ctx->bin_alive, ctx->d_alive_union, ctx->hp.ncb, s));
				HVCF_CUDA(launch_shard_wait_routed(pp, parity, s, 1u));
				const int prc = route(head_end, own1, 1u);
				if (prc != HVCF_OK)
					return prc;
				HVCF_CUDA(launch_shard_wait_routed(pp, parity, s, 2u));
				if (ctx->ktimes)
					HVCF_CUDA(cudaEventRecord(ctx->kev_phase2[frame][1], s));
				ctx->open_launches += 4;
			}
			BinParams br = b;
			br.tile_begin = rd.tile_begin;
			br.macro = ctx->bin_macro;
			br.group_alive = ctx->bin_alive;
			br.cb_union = ctx->bin_alive ? ctx->d_alive_union : ctx->d_cb_union;
			const cudaError_t e = launch_hier_bin(br, rd.tile_end - rd.tile_begin, s);
			if (e != cudaSuccess)
				return fail(HVCF_ERR_CUDA, "hvcf_cull_routed: binning of round %u: %s", r, cudaGetErrorString(e));
			return static_cast<int>(HVCF_OK);
		});
		if (rrc != HVCF_OK)
			return rrc;
	}
	else
	{
		HVCF_CUDA(launch_hier_bin(b, sh.tiles_phys, s));
		++ctx->open_launches;
	}
	HVCF_CUDA(hvcf_kmark(ctx, frame, 2, s));
	// the route blocks' record counts of this parity are consumed: cleared here, so that the senders of the next frame of this
	// parity only store the non-zero ones (nobody writes them before this rank has posted the totals of the frame in between)
	HVCF_CUDA(cudaMemcpyAsync(sh.d_cnt_last, box + sh.off_cnt, sizeof(uint32_t) * sh.blocks_phys, cudaMemcpyDeviceToDevice, s)); // (hvcf_shard_inbox_records)
	HVCF_CUDA(cudaMemsetAsync(box + sh.off_cnt, 0, sizeof(uint32_t) * sh.blocks_phys, s));
	ctx->range_transformed = ctx->range_binned = sh.tiles_phys * kTileLights;

	// ---- count pass + scan; this rank's total is POSTED at once, but nobody waits yet: the write pass runs on local positions
	// (own segment buffer, shard-local offsets) while the other ranks finish counting; the push then adds the index base and
	// places cells + segment on every rank, this one included ----
	rc = hvcf_hier_close(ctx, frame, s);
	if (rc == HVCF_OK)
		rc = hvcf_peer_totals_step(ctx, stream, 1u);
	if (rc == HVCF_OK)
	{
		sh.seg_local = true;
		rc = hvcf_cull_finish(ctx, frame, nullptr, stream);
		sh.seg_local = false;
	}
	if (rc == HVCF_OK)
		rc = hvcf_peer_totals_step(ctx, stream, 2u);
	if (rc == HVCF_OK)
	{
		static const uint32_t blocks = static_cast<uint32_t>(env_int("HVCF_PEER_BLOCKS", 592) > 0 ? env_int("HVCF_PEER_BLOCKS", 592) : 592);
		HVCF_CUDA(launch_peer_push_relocate(ctx->peer, ctx->d_shard_total, ctx->d_index_base, sh.d_seg, ctx->cluster_begin, ctx->cluster_count, blocks, s));
	}
	if (rc == HVCF_OK)
		rc = hvcf_peer_wait_grid(ctx, stream);
	return rc;
}

int hvcf_download_shard(hvcf_ctx* ctx, void* cells_dst, size_t cells_bytes, void* indices_dst, size_t indices_capacity_bytes, uint32_t* out_index_base,
	uint32_t* out_total, void* stream)
{
	if (!ctx || !out_index_base || !out_total || (cells_bytes && !cells_dst))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_download_shard: null argument");
	if (cells_bytes > sizeof(hvcf_light_cell) * static_cast<size_t>(ctx->cluster_count))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_download_shard: %zu bytes of cells > the shard's %u clusters", cells_bytes, ctx->cluster_count);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	if (!ctx->h_total)
		HVCF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&ctx->h_total), sizeof(uint32_t) * 16));
	uint32_t* h = ctx->h_total; // [8] = index base, [9] = shard total (slots 0..7 belong to hvcf_download_grid's frames)
	h[8] = 0u;
	if (ctx->peer.world && ctx->d_index_base)
		HVCF_CUDA(cudaMemcpyAsync(&h[8], ctx->d_index_base, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	HVCF_CUDA(cudaMemcpyAsync(&h[9], ctx->d_shard_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	if (cells_bytes)
		HVCF_CUDA(cudaMemcpyAsync(cells_dst, ctx->d_cells + ctx->cluster_begin, cells_bytes, cudaMemcpyDeviceToHost, s));
	HVCF_CUDA(cudaStreamSynchronize(s));
	if (const int prc = hvcf_peer_error_after_sync(ctx, "hvcf_download_shard"))
		return prc;
	*out_index_base = h[8];
	*out_total = h[9];
	const size_t need = sizeof(uint32_t) * static_cast<size_t>(h[9]);
	if (need > indices_capacity_bytes || (need && !indices_dst))
		return fail(HVCF_ERR_CAPACITY, "hvcf_download_shard: %u indices do not fit in %zu bytes", h[9], indices_capacity_bytes);
	if (need)
	{
		HVCF_CUDA(cudaMemcpyAsync(indices_dst, ctx->d_indices + h[8], need, cudaMemcpyDeviceToHost, s));
		HVCF_CUDA(cudaStreamSynchronize(s));
	}
	return HVCF_OK;
}

int hvcf_debug_route_ms(hvcf_ctx* ctx, uint32_t frame, float* out_route_ms)
{
	if (!ctx || !out_route_ms || frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_debug_route_ms: bad argument");
	if (!ctx->ktimes || !ctx->kev[frame][7] || !ctx->kev[frame][0])
		return fail(HVCF_ERR_NOT_READY, "hvcf_debug_route_ms: needs HVCF_KERNEL_TIMES=1 at hvcf_create and a routed frame on this slot");
	HVCF_CUDA(cudaEventSynchronize(ctx->kev[frame][7]));
	HVCF_CUDA(cudaEventElapsedTime(out_route_ms, ctx->kev[frame][0], ctx->kev[frame][7]));
	return HVCF_OK;
}

int hvcf_shard_inbox_records(hvcf_ctx* ctx, void* stream, uint64_t* out_records)
{
	if (!ctx || !out_records)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_shard_inbox_records: null argument");
	if (!ctx->shard.on)
		return fail(HVCF_ERR_NOT_READY, "hvcf_shard_inbox_records: not a context of hvcf_create_sharded");
	const ShardState& sh = ctx->shard;
	HVCF_CUDA(cudaSetDevice(ctx->device));
	HVCF_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
	std::vector<uint32_t> cnt(sh.blocks_phys);
	HVCF_CUDA(cudaMemcpy(cnt.data(), sh.d_cnt_last, sizeof(uint32_t) * sh.blocks_phys, cudaMemcpyDeviceToHost));
	uint64_t sum = 0;
	for (uint32_t v : cnt)
		sum += v;
	*out_records = sum;
	return HVCF_OK;
}

} // extern "C"
