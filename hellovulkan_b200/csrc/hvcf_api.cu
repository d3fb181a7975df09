// hvcf_api.cu -- the extern "C" boundary of libhvcf.so (include/hvcf.h).  Owns the device buffers in
// the reference's layouts, stages UBOs through pinned memory, and strings the kernels of
// hvcf_kernels.cu together on the caller's stream.  No CPU fallback: every entry point fails with
// HVCF_ERR_CUDA when no sm_100 device is usable.
#include "hvcf_ctx.cuh"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <new>

#include "hvcf_shade.cuh"

using namespace hvcf;

namespace hvcf
{
static thread_local char g_last_error[512] = "";

int fail(int code, const char* fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
	va_end(ap);
	return code;
}
} // namespace hvcf

static void free_ctx(hvcf_ctx* c)
{
	if (!c)
		return;
	peer_unmap(c);
	hvcf_shard_free(c);
	cudaFree(c->d_mail);
	cudaFree(c->d_index_base);
	cudaFree(c->d_peer_totals);
	cudaFree(c->d_lights);
	cudaFree(c->d_aabb);
	cudaFree(c->d_cells);
	cudaFree(c->d_indices);
	cudaFree(c->d_global_index);
	cudaFree(c->d_ubo);
	cudaFree(c->d_lights_view);
	cudaFree(c->d_zplanes);
	cudaFree(c->d_seg_counts);
	cudaFree(c->d_seg_lists);
	cudaFree(c->d_counts);
	cudaFree(c->d_block_sums);
	cudaFree(c->d_block_offsets);
	cudaFree(c->d_shard_total);
	cudaFree(c->d_stats);
	cudaFree(c->d_hseg_counts);
	cudaFree(c->d_hseg_prefix);
	cudaFree(c->d_hunit_any);
	cudaFree(c->d_hscratch);
	cudaFree(c->d_hoffsets);
	cudaFree(c->d_hlane_cluster);
	cudaFree(c->d_cb_union);
	cudaFree(c->d_sb_union);
	cudaFree(c->d_bin_count);
	cudaFree(c->d_bin_rec);
	cudaFree(c->d_small_state);
	cudaFree(c->d_small_over);
	cudaFree(c->d_small_ctl);
	cudaFree(c->d_group_alive);
	cudaFree(c->d_alive_union);
	cudaFree(c->d_bin_idx);
	cudaFree(c->d_scan_state);
	cudaFree(c->d_trace);
	for (int i = 0; i < 8; ++i)
	{
		if (c->ubo_events[i])
			cudaEventDestroy(c->ubo_events[i]);
		for (int j = 0; j < 8; ++j)
			if (c->kev[i][j])
				cudaEventDestroy(c->kev[i][j]);
	}
	cudaFree(c->d_slice_thr);
	cudaFree(c->d_anim_table);
	cudaFree(c->d_base_y);
	cudaFree(c->d_upd_idx);
	cudaFree(c->d_upd_xyz);
	for (int i = 0; i < 2; ++i)
	{
		cudaFreeHost(c->h_upd_idx[i]);
		cudaFreeHost(c->h_upd_xyz[i]);
		if (c->upd_events[i])
			cudaEventDestroy(c->upd_events[i]);
	}
	cudaFreeHost(c->h_slice_thr);
	cudaFreeHost(c->h_ubo);
	cudaFreeHost(c->h_zplanes);
	cudaFreeHost(c->h_one_light);
	cudaFreeHost(c->h_total);
	for (auto& e : c->one_light_events)
		if (e)
			cudaEventDestroy(e);
	for (auto& fe : c->ev)
		for (auto& e : fe)
			if (e)
				cudaEventDestroy(e);
	for (auto& st : c->write_streams)
		if (st)
			cudaStreamDestroy(st);
	for (auto& e : c->write_ev)
		if (e)
			cudaEventDestroy(e);
	for (auto& fe : c->kev_phase2)
		for (auto& e : fe)
			if (e)
				cudaEventDestroy(e);
	for (auto& fe : c->kev_round)
		for (auto& re : fe)
			for (auto& e : re)
				if (e)
					cudaEventDestroy(e);
	delete[] c->ubo_set;
	delete c;
}

static HierParams hier_params(hvcf_ctx* ctx, uint32_t frame)
{
	HierParams p{};
	p.lights_view = ctx->d_lights_view;
	p.aabbs = ctx->d_aabb;
	p.lane_cluster = ctx->d_hlane_cluster;
	p.cluster_begin = ctx->cluster_begin;
	p.cluster_count = ctx->cluster_count;
	p.ncb = ctx->hp.ncb;
	p.tiles_total = ctx->h_tiles_total;
	p.tiles_per_seg = ctx->h_tiles_per_seg;
	p.num_segs = ctx->h_num_segs;
	p.parts = ctx->h_parts;
	// long segments walk many slots per warp program: pipelined variant (measured: wins at 32 tiles per segment, loses at 4)
	p.pipe = ctx->h_pipe;
	p.seg_begin = 0;
	p.seg_end = ctx->h_num_segs;
	p.cap = ctx->cap;
	p.seg_counts = ctx->d_hseg_counts;
	p.seg_prefix = ctx->d_hseg_prefix;
	p.cum = ctx->d_hscratch;
	p.offsets = ctx->d_hoffsets;
	p.counts = ctx->d_counts;
	p.index_base = nullptr;
	p.shard_total = ctx->d_shard_total;
	p.cells = ctx->d_cells;
	p.indices = ctx->d_indices;
	p.frame_total = &ctx->d_global_index[frame];
	p.stats = &ctx->d_stats[frame];
	p.trace = ctx->d_trace;
	p.bin_count = ctx->d_bin_count;
	p.bin_rec = ctx->d_bin_rec;
	p.bin_idx = ctx->d_bin_idx;
	p.tiles_cap = ctx->tiles_cap;
	p.unit_any = ctx->d_hunit_any;
	if (ctx->shard.seg_local) // the write pass runs before the totals exchange: local positions, relocated by the push
		p.indices = ctx->shard.d_seg;
	if (ctx->routed_frame) // the tiles are this rank's inbox of the frame's parity
	{
		const ShardState& sh = ctx->shard;
		unsigned char* box = sh.d_inbox + static_cast<size_t>(sh.epoch & 1u) * sh.parity_bytes;
		p.lights_view = reinterpret_cast<const float4*>(box);
		p.tile_ids = sh.d_tile_ids;
		p.tile_cnt = sh.d_tile_cnt;
		p.tile_mask = ~(sh.block_tiles - 1u);
	}
	return p;
}

extern "C" {

const char* hvcf_last_error_string(void) { return g_last_error; }
const char* hvcf_version(void) { return "hvcf 0.1 (sm_100a)"; }

int hvcf_create(const hvcf_config* config, hvcf_ctx** out_ctx)
{
	if (!config || !out_ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create: null argument");
	if (config->struct_size != sizeof(hvcf_config))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create: struct_size %u != %zu", config->struct_size, sizeof(hvcf_config));
	*out_ctx = nullptr;
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
		return fail(HVCF_ERR_CUDA, "hvcf_create: no CUDA device (libhvcf has no CPU fallback)");
	int dev = config->device;
	if (dev < 0)
		HVCF_CUDA(cudaGetDevice(&dev));
	if (dev >= ndev)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create: device %d of %d", dev, ndev);
	cudaDeviceProp prop;
	HVCF_CUDA(cudaGetDeviceProperties(&prop, dev));
	if (prop.major != 10)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_create: device %d is sm_%d%d; libhvcf is built for sm_100a only", dev,
			prop.major, prop.minor);
	HVCF_CUDA(cudaSetDevice(dev));

	hvcf_ctx* c = new (std::nothrow) hvcf_ctx();
	if (!c)
		return fail(HVCF_ERR_OUT_OF_MEMORY, "hvcf_create: host allocation failed");
	c->cfg = *config;
	c->device = dev;
	c->X = config->slice_count_x ? config->slice_count_x : HVCF_DEFAULT_SLICE_COUNT_X;
	c->Y = config->slice_count_y ? config->slice_count_y : HVCF_DEFAULT_SLICE_COUNT_Y;
	c->Z = config->slice_count_z ? config->slice_count_z : HVCF_DEFAULT_SLICE_COUNT_Z;
	c->cap = config->max_lights_per_cluster ? config->max_lights_per_cluster : HVCF_DEFAULT_MAX_LIGHTS_PER_CLUSTER;
	c->frame_count = config->frame_count ? config->frame_count : HVCF_DEFAULT_FRAME_COUNT;
	c->max_lights = config->max_lights ? config->max_lights : 1024;
	const uint64_t total64 = static_cast<uint64_t>(c->X) * c->Y * c->Z;
	if (total64 == 0 || total64 > (1ull << 26) || c->frame_count > 8 || c->cap > (1u << 16) ||
		total64 * c->cap >= (1ull << 32) || c->max_lights > 0xffff0000u)
	{
		delete c;
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create: grid %ux%ux%u cap %u frames %u out of range", config->slice_count_x,
			config->slice_count_y, config->slice_count_z, config->max_lights_per_cluster, config->frame_count);
	}
	c->cluster_total = static_cast<uint32_t>(total64);
	uint32_t zb = config->z_begin, ze = config->z_end;
	if (zb == 0 && ze == 0)
		ze = c->Z;
	if (zb >= ze || ze > c->Z)
	{
		delete c;
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create: z shard [%u,%u) outside [0,%u)", zb, ze, c->Z);
	}
	uint32_t rb = zb * c->Y, re = ze * c->Y; // the shard in cluster rows
	if (config->row_end)
	{
		rb = config->row_begin, re = config->row_end;
		if (rb >= re || re > c->Y * c->Z)
		{
			delete c;
			return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_create: row shard [%u,%u) outside [0,%u)", rb, re, c->Y * c->Z);
		}
		zb = rb / c->Y;                  // the slices that cover the rows: the plan's box
		ze = (re + c->Y - 1) / c->Y;
	}
	uint32_t mode = config->mode;
	if (mode > HVCF_MODE_HIERARCHICAL)
	{
		delete c;
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_create: unknown mode %u", config->mode);
	}
	if (mode == HVCF_MODE_AUTO)
	{
		const char* m = getenv("HVCF_MODE");
		mode = (m && !strcmp(m, "brute")) ? HVCF_MODE_BRUTE : HVCF_MODE_HIERARCHICAL;
	}
	c->hier = mode == HVCF_MODE_HIERARCHICAL;
	c->cluster_begin = rb * c->X;
	c->cluster_end = re * c->X;
	c->cluster_count = c->cluster_end - c->cluster_begin;
	c->padded_capacity = round_up_u32(c->max_lights ? c->max_lights : 1, kTileLights);
	c->download_kernel = env_int("HVCF_DOWNLOAD_KERNEL", 1) != 0;
	{
		const int db = env_int("HVCF_DOWNLOAD_BLOCKS", 64);
		c->download_blocks = static_cast<uint32_t>(db < 1 ? 1 : (db > 4096 ? 4096 : db));
		const int mb = env_int("HVCF_DOWNLOAD_KERNEL_MAX_MB", 4);
		c->download_max_words = static_cast<uint32_t>(mb < 0 ? 0 : (mb > 2048 ? 2048 : mb)) << 18;
	}
	// light segments: enough CTAs to fill 148 SMs x 4 resident CTAs about twice
	const uint32_t per_cta = kCullClustersPerCta;
	const uint32_t groups = (c->cluster_count + per_cta - 1) / per_cta;
	uint32_t segs = (2 * 148 * 4 + groups - 1) / groups;
	segs = segs < 1 ? 1 : (segs > 32 ? 32 : segs);
	const int env_segs = env_int("HVCF_SEGMENTS", 0);
	if (env_segs > 0)
		segs = static_cast<uint32_t>(env_segs > 64 ? 64 : env_segs);
	c->max_segs = segs;
	c->count_blocks = (c->cluster_count + 255) / 256;
	if (c->hier)
	{
		uint32_t shape[3] = {0, 0, 0};
		if (const char* sh = getenv("HVCF_HIER_SHAPE"))
			sscanf(sh, "%u,%u,%u", &shape[0], &shape[1], &shape[2]);
		hier_plan_grid(c->X, c->Y, ze - zb, env_int("HVCF_HIER_K", 0), shape, c->max_lights, &c->hp);
		c->hp.brick = env_int("HVCF_HIER_BRICK", 1) != 0 ? 1u : 0u;
		// blocks of the covering box that hold none of the shard's clusters are dropped here, before anything is sized by the plan
		hier_build_lane_map(&c->hp, c->X, c->Y, ze - zb, c->cluster_begin - zb * c->X * c->Y, c->cluster_end - zb * c->X * c->Y, &c->lane_map);
		const uint32_t max_tiles = c->padded_capacity / kTileLights;
		// most light segments a frame may be cut into (hier_plan_lights picks the actual number).  Measured on B200
		// (scripts/sweep_hier.py): grid A x 1M lights 0.181 ms at 256 -> 0.131 ms at 512 (half the executed tests: shorter
		// segments publish their counts sooner); grid B x 1M keeps 256 by the plan's own target, forcing 512 there costs
		// 0.240 -> 0.339 ms (twice the warp-program prologues)
		uint32_t seg_limit = static_cast<uint32_t>(env_int("HVCF_HIER_MAX_SEGS", 512));
		seg_limit = seg_limit < 1 ? 1 : (seg_limit > 2048 ? 2048 : seg_limit);
		c->h_plan_max_segs = max_tiles < seg_limit ? max_tiles : seg_limit;
		// frames of few tiles cut every tile into light segments of a part of its survivor slot (hier2_kernel): measured on B200,
		// grid A x 65,536 lights
		{
			const int parts = env_int("HVCF_HIER_PARTS", 0); // 0 = by the number of tiles
			c->h_parts_cfg = static_cast<uint32_t>(parts < 0 ? 0 : (parts > static_cast<int>(kHMaxParts) ? static_cast<int>(kHMaxParts) : parts));
		}
		c->h_max_segs = static_cast<uint64_t>(max_tiles) * kHMaxParts < seg_limit ? max_tiles * kHMaxParts : seg_limit;
		c->h_max_chunks = hier_num_chunks(c->h_max_segs);
		c->h_tps_override = env_int("HVCF_HIER_TPS", 0);
		c->h_split_head = static_cast<uint32_t>(env_int("HVCF_HIER_SPLIT", 0));
		c->ktimes = env_int("HVCF_KERNEL_TIMES", 0) != 0;
		c->tiles_cap = max_tiles;
		// two-level level-0 binning (HVCF_BIN2=0/1 forces it off/on; otherwise chosen per frame, see hier_range)
		c->bin2_mode = env_int("HVCF_BIN2", -1);
		c->xbin_max_lights = static_cast<uint32_t>(env_int("HVCF_XBIN_MAX", 262144));
		c->rounds_mode = env_int("HVCF_ROUNDS", -1);
		c->rounds_min_tiles = static_cast<uint32_t>(env_int("HVCF_ROUNDS_MIN_TILES", 2048));
		c->rounds_head = static_cast<uint32_t>(env_int("HVCF_ROUNDS_HEAD", 64));
		c->rounds_growth = static_cast<uint32_t>(env_int("HVCF_ROUNDS_GROWTH", 8));
		c->rounds_macro = env_int("HVCF_ROUNDS_MACRO", -1);
		c->write_fork = env_int("HVCF_WRITE_FORK", 1) != 0;
		c->small_mode = env_int("HVCF_SMALL", 1) != 0;
		c->rounds_late_segs = static_cast<uint32_t>(env_int("HVCF_ROUNDS_LATE_SEGS", 32));
		c->rounds_early_segs = static_cast<uint32_t>(env_int("HVCF_ROUNDS_EARLY_SEGS", 16));
		if (c->rounds_early_segs == 0)
			c->rounds_early_segs = 1;
		if (c->rounds_late_segs == 0)
			c->rounds_late_segs = 1;
		c->pipe_mode = env_int("HVCF_BATCH_PIPE", -1);
		c->pipe_min_tps = static_cast<uint32_t>(env_int("HVCF_PIPE_MIN_TPS", 16));
		if (c->bin2_mode != 0)
		{
			c->bin_sg = (c->hp.ncb + 31) / 32 > 8 ? (c->hp.ncb + 31) / 32 : 8;
			c->bin_ns = (c->hp.ncb + c->bin_sg - 1) / c->bin_sg;
		}
	}

#define HVCF_ALLOC(ptr, bytes)                                                              \
	do                                                                                      \
	{                                                                                       \
		cudaError_t e_ = cudaMalloc(reinterpret_cast<void**>(&(ptr)), (bytes) ? (bytes) : 16); \
		c->device_bytes += (bytes) ? (bytes) : 16;                                           \
		if (e_ != cudaSuccess)                                                              \
		{                                                                                   \
			free_ctx(c);                                                                    \
			return fail(HVCF_ERR_OUT_OF_MEMORY, "hvcf_create: cudaMalloc(%zu) for %s: %s", static_cast<size_t>(bytes), #ptr, \
				cudaGetErrorString(e_));                                                    \
		}                                                                                   \
	} while (0)
	const size_t C = c->cluster_total, Cs = c->cluster_count;
	HVCF_ALLOC(c->d_lights, sizeof(hvcf_light_data) * static_cast<size_t>(c->max_lights));
	HVCF_ALLOC(c->d_aabb, sizeof(hvcf_aabb) * C);
	HVCF_ALLOC(c->d_cells, sizeof(hvcf_light_cell) * C);
	HVCF_ALLOC(c->d_indices, sizeof(uint32_t) * C * c->cap);
	HVCF_ALLOC(c->d_global_index, sizeof(uint32_t) * c->frame_count);
	HVCF_ALLOC(c->d_ubo, sizeof(hvcf_cluster_forward_ubo) * c->frame_count);
	HVCF_ALLOC(c->d_lights_view, static_cast<size_t>(c->padded_capacity) * 16);
	HVCF_ALLOC(c->d_zplanes, sizeof(float) * (c->Z + 1) * c->frame_count);
	if (c->hier)
	{
		HVCF_ALLOC(c->d_hseg_counts, sizeof(uint32_t) * Cs * c->h_max_segs);
		HVCF_ALLOC(c->d_hseg_prefix, sizeof(uint32_t) * Cs * c->h_max_segs);
		HVCF_ALLOC(c->d_hscratch, sizeof(uint32_t) * Cs * kHMaxChunks);
		HVCF_ALLOC(c->d_hoffsets, sizeof(uint32_t) * Cs);
		HVCF_ALLOC(c->d_hunit_any, sizeof(uint32_t) * static_cast<size_t>(c->h_max_segs) * c->hp.ncb * 8);
		HVCF_ALLOC(c->d_hlane_cluster, sizeof(uint32_t) * 32 * static_cast<size_t>(c->hp.ncb) * 8 * c->hp.K);
		{
			const size_t slots = static_cast<size_t>(c->hp.ncb) * c->tiles_cap;
			{
				// the survivor slots are DENSE -- groups x tiles x kBinCap records, whatever the filter lets through (64x36x96 clusters
				// x 16M lights: 3.5 M slots, 32 GB) -- so say what is missing before cudaMalloc fails with a bare "out of memory"
				const size_t need = slots * kBinCap * (sizeof(float4) + sizeof(uint16_t)) + slots * sizeof(uint32_t);
				size_t free_b = 0, total_b = 0;
				if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && need + (static_cast<size_t>(1) << 30) > free_b)
				{
					free_ctx(c);
					return fail(HVCF_ERR_OUT_OF_MEMORY, "hvcf_create: the level-0 survivor slots of %u cluster groups x %u light tiles need %.1f GB, %.1f GB are free: "
						"lower max_lights, or shard the grid (hvcf_create_sharded: a rank's slots shrink with its slab)", c->hp.ncb, c->tiles_cap, need / 1e9, free_b / 1e9);
				}
			}
			HVCF_ALLOC(c->d_cb_union, sizeof(float) * 8 * c->hp.ncb);
			HVCF_ALLOC(c->d_sb_union, sizeof(float) * 8 * (c->bin_ns ? c->bin_ns : 1));
			HVCF_ALLOC(c->d_bin_count, sizeof(uint32_t) * slots);
			HVCF_ALLOC(c->d_group_alive, sizeof(uint32_t) * c->hp.ncb * kHMaxChunks);
			HVCF_ALLOC(c->d_alive_union, sizeof(float) * 8 * c->hp.ncb);
			HVCF_ALLOC(c->d_bin_rec, sizeof(float4) * slots * kBinCap);
			HVCF_ALLOC(c->d_bin_idx, sizeof(uint16_t) * slots * kBinCap);
		}
		HVCF_ALLOC(c->d_scan_state, sizeof(unsigned long long) * ((Cs + 31) / 32) * c->frame_count);
		HVCF_ALLOC(c->d_small_state, sizeof(unsigned long long) * ((Cs + 31) / 32));
		HVCF_ALLOC(c->d_small_over, sizeof(uint32_t) * ((Cs + 31) / 32));
		HVCF_ALLOC(c->d_small_ctl, sizeof(uint32_t) * 4);
		if (env_int("HVCF_HIER_TRACE", 0))
			HVCF_ALLOC(c->d_trace, sizeof(unsigned long long) * 16 * static_cast<size_t>(c->hp.ncb) * c->h_max_segs);
	}
	else
	{
		HVCF_ALLOC(c->d_seg_counts, sizeof(uint32_t) * Cs * c->max_segs);
		HVCF_ALLOC(c->d_seg_lists, sizeof(uint32_t) * Cs * c->max_segs * c->cap);
	}
	HVCF_ALLOC(c->d_counts, sizeof(uint32_t) * Cs);
	HVCF_ALLOC(c->d_block_sums, sizeof(uint32_t) * c->count_blocks);
	HVCF_ALLOC(c->d_block_offsets, sizeof(uint32_t) * c->count_blocks);
	HVCF_ALLOC(c->d_shard_total, sizeof(uint32_t));
	HVCF_ALLOC(c->d_stats, sizeof(DeviceStats) * c->frame_count);
#undef HVCF_ALLOC
	cudaError_t e = cudaMallocHost(reinterpret_cast<void**>(&c->h_ubo), sizeof(hvcf_cluster_forward_ubo) * c->frame_count);
	if (e == cudaSuccess)
		e = cudaMallocHost(reinterpret_cast<void**>(&c->h_zplanes), sizeof(float) * (c->Z + 1) * c->frame_count);
	if (e == cudaSuccess)
		e = cudaMallocHost(reinterpret_cast<void**>(&c->h_one_light), sizeof(hvcf_light_data) * 16);
	for (int i = 0; i < 16 && e == cudaSuccess; ++i)
		e = cudaEventCreateWithFlags(&c->one_light_events[i], cudaEventDisableTiming);
	if (e == cudaSuccess)
		e = cudaMemset(c->d_cells, 0, sizeof(hvcf_light_cell) * C);
	if (e == cudaSuccess)
		e = cudaMemset(c->d_aabb, 0, sizeof(hvcf_aabb) * C);
	if (e == cudaSuccess)
		e = cudaMemset(c->d_global_index, 0, sizeof(uint32_t) * c->frame_count);
	if (e == cudaSuccess)
		e = cudaMemset(c->d_stats, 0, sizeof(DeviceStats) * c->frame_count);
	if (e == cudaSuccess)
		e = cudaMemset(c->d_shard_total, 0, sizeof(uint32_t));
	if (e == cudaSuccess && c->hier)
	{
		e = cudaMemcpy(c->d_hlane_cluster, c->lane_map.data(), c->lane_map.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
		std::vector<uint32_t>().swap(c->lane_map);
	}
	if (e == cudaSuccess && c->hier)
		e = cudaMemset(c->d_scan_state, 0, sizeof(unsigned long long) * ((Cs + 31) / 32) * c->frame_count);
	if (e == cudaSuccess && c->hier)
		e = cudaMemset(c->d_small_state, 0, sizeof(unsigned long long) * ((Cs + 31) / 32));
	if (e == cudaSuccess && c->hier)
	{
		const uint32_t ctl[4] = {1u, 0u, 0u, 0u}; // serial (never 0: the zeroed state words carry tag 0), ticket counter
		e = cudaMemcpy(c->d_small_ctl, ctl, sizeof(ctl), cudaMemcpyHostToDevice);
	}
	if (e != cudaSuccess)
	{
		free_ctx(c);
		return fail(HVCF_ERR_CUDA, "hvcf_create: pinned staging / init: %s", cudaGetErrorString(e));
	}
	c->ubo_set = new bool[c->frame_count]();
	*out_ctx = c;
	return HVCF_OK;
}

int hvcf_destroy(hvcf_ctx* ctx)
{
	if (!ctx)
		return HVCF_OK;
	cudaSetDevice(ctx->device);
	cudaDeviceSynchronize();
	free_ctx(ctx);
	return HVCF_OK;
}

int hvcf_set_lights(hvcf_ctx* ctx, const void* lights, uint32_t n, int src_is_device, void* stream)
{
	if (!ctx || (n && !lights))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_lights: null argument");
	if (n > ctx->max_lights)
		return fail(HVCF_ERR_CAPACITY, "hvcf_set_lights: %u lights > max_lights %u", n, ctx->max_lights);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	if (n)
		HVCF_CUDA(cudaMemcpyAsync(ctx->d_lights, lights, sizeof(hvcf_light_data) * static_cast<size_t>(n),
			src_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
	ctx->num_lights = n;
	ctx->bound_lights = nullptr;
	if (ctx->anim_len) // new lights, new rest positions
	{
		HVCF_CUDA(launch_capture_base_y(ctx->d_lights, n, ctx->d_base_y, static_cast<cudaStream_t>(stream)));
		ctx->base_y_count = n;
	}
	return HVCF_OK;
}

int hvcf_bind_lights(hvcf_ctx* ctx, const void* device_lights, uint32_t n)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_bind_lights: null context");
	if (n > ctx->max_lights)
		return fail(HVCF_ERR_CAPACITY, "hvcf_bind_lights: %u lights > max_lights %u", n, ctx->max_lights);
	if (ctx->shard.on)
	{
		// a context of hvcf_create_sharded binds exactly the lights it OWNS: device_lights[0] is light `begin` of
		// hvcf_shard_light_range.  Only hvcf_cull_routed reads such a buffer; the scene's light count stays what it was.
		uint32_t lb = 0, le = 0;
		hvcf_shard_light_range(ctx, &lb, &le);
		if (device_lights && n != le - lb)
			return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_bind_lights: a sharded context binds the %u lights it owns, not %u", le - lb, n);
		if (device_lights && ctx->anim_len)
			return fail(HVCF_ERR_UNSUPPORTED, "hvcf_bind_lights: device-side light animation is not routed");
		ctx->bound_lights = static_cast<const hvcf_light_data*>(device_lights);
		return HVCF_OK;
	}
	const bool changed = ctx->bound_lights != static_cast<const hvcf_light_data*>(device_lights) || n > ctx->base_y_count;
	ctx->bound_lights = static_cast<const hvcf_light_data*>(device_lights);
	ctx->num_lights = device_lights ? n : 0;
	// device-side animation moves the lights about their REST height: a new buffer, or more lights than were captured, needs a new
	// capture (on the legacy default stream, which orders itself against every blocking stream).  The moved
	// positions are written back into the bound buffer, as UpdateLightPosition writes them into the SSBO.
	if (ctx->anim_len && device_lights && n && changed)
	{
		HVCF_CUDA(cudaSetDevice(ctx->device));
		HVCF_CUDA(launch_capture_base_y(ctx->bound_lights, n, ctx->d_base_y, nullptr));
		ctx->base_y_count = n;
	}
	return HVCF_OK;
}

int hvcf_update_light(hvcf_ctx* ctx, uint32_t index, const float xyz[3], void* stream)
{
	if (!ctx || !xyz)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_update_light: null argument");
	if (index >= ctx->num_lights)
		return HVCF_OK; // ResourcesLight::UpdateLightPosition ignores an out-of-range index (ResourcesLight.cpp:22-25)
	if (ctx->shard.on && ctx->bound_lights)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_update_light: the owned lights live in a bound buffer (hvcf_bind_lights); write to it directly");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	const uint32_t slot = ctx->one_light_cursor++ & 15u;
	HVCF_CUDA(cudaEventSynchronize(ctx->one_light_events[slot])); // slot's previous copy has been consumed
	float* h = ctx->h_one_light[slot].position;
	h[0] = xyz[0], h[1] = xyz[1], h[2] = xyz[2], h[3] = 1.0f;
	hvcf_light_data* base = ctx->bound_lights ? const_cast<hvcf_light_data*>(ctx->bound_lights) : ctx->d_lights;
	HVCF_CUDA(cudaMemcpyAsync(base[index].position, h, 16, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
	HVCF_CUDA(cudaEventRecord(ctx->one_light_events[slot], static_cast<cudaStream_t>(stream)));
	return HVCF_OK;
}

int hvcf_update_lights(hvcf_ctx* ctx, const uint32_t* indices, const float* xyz, uint32_t count, void* stream)
{
	if (!ctx || (count && (!indices || !xyz)))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_update_lights: null argument");
	if (count == 0)
		return HVCF_OK;
	if (ctx->shard.on && ctx->bound_lights)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_update_lights: the owned lights live in a bound buffer (hvcf_bind_lights); write to it directly");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	if (count > ctx->upd_capacity)
	{
		// grow: everything in flight on the old staging must be done first
		HVCF_CUDA(cudaDeviceSynchronize());
		uint32_t cap = ctx->upd_capacity ? ctx->upd_capacity : 1024;
		while (cap < count)
			cap *= 2;
		cudaFree(ctx->d_upd_idx), cudaFree(ctx->d_upd_xyz);
		ctx->d_upd_idx = nullptr, ctx->d_upd_xyz = nullptr;
		for (int i = 0; i < 2; ++i)
		{
			cudaFreeHost(ctx->h_upd_idx[i]), cudaFreeHost(ctx->h_upd_xyz[i]);
			ctx->h_upd_idx[i] = nullptr, ctx->h_upd_xyz[i] = nullptr;
		}
		ctx->upd_capacity = 0;
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_upd_idx), sizeof(uint32_t) * 2 * cap));
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_upd_xyz), sizeof(float) * 3 * 2 * cap));
		for (int i = 0; i < 2; ++i)
		{
			HVCF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&ctx->h_upd_idx[i]), sizeof(uint32_t) * cap));
			HVCF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&ctx->h_upd_xyz[i]), sizeof(float) * 3 * cap));
			if (!ctx->upd_events[i])
				HVCF_CUDA(cudaEventCreateWithFlags(&ctx->upd_events[i], cudaEventDisableTiming));
		}
		ctx->upd_capacity = cap;
	}
	// double-buffered pinned staging: the caller's arrays are free again when this returns, and the copy of batch k
	// overlaps whatever the GPU still does with batch k-1
	const uint32_t slot = ctx->upd_cursor++ & 1u;
	HVCF_CUDA(cudaEventSynchronize(ctx->upd_events[slot]));
	memcpy(ctx->h_upd_idx[slot], indices, sizeof(uint32_t) * count);
	memcpy(ctx->h_upd_xyz[slot], xyz, sizeof(float) * 3 * count);
	uint32_t* di = ctx->d_upd_idx + static_cast<size_t>(slot) * ctx->upd_capacity;
	float* dx = ctx->d_upd_xyz + static_cast<size_t>(slot) * 3 * ctx->upd_capacity;
	HVCF_CUDA(cudaMemcpyAsync(di, ctx->h_upd_idx[slot], sizeof(uint32_t) * count, cudaMemcpyHostToDevice, s));
	HVCF_CUDA(cudaMemcpyAsync(dx, ctx->h_upd_xyz[slot], sizeof(float) * 3 * count, cudaMemcpyHostToDevice, s));
	HVCF_CUDA(cudaEventRecord(ctx->upd_events[slot], s));
	hvcf_light_data* lights = const_cast<hvcf_light_data*>(ctx->bound_lights ? ctx->bound_lights : ctx->d_lights);
	HVCF_CUDA(launch_scatter_positions(lights, ctx->num_lights, di, dx, count, ctx->anim_len ? ctx->d_base_y : nullptr, s));
	return HVCF_OK;
}

int hvcf_set_light_animation(hvcf_ctx* ctx, const float* table, uint32_t table_len, float amplitude, void* stream)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_light_animation: null context");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	if (!table || table_len == 0)
	{
		ctx->anim_len = 0; // off; the light buffer keeps the positions of the last animated frame
		return HVCF_OK;
	}
	if (ctx->shard.on && ctx->bound_lights)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_set_light_animation: not with a bound buffer on a sharded context");
	if (table_len & (table_len - 1))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_light_animation: table_len %u is not a power of two", table_len);
	if (table_len != ctx->anim_len)
	{
		HVCF_CUDA(cudaStreamSynchronize(s));
		cudaFree(ctx->d_anim_table);
		ctx->d_anim_table = nullptr;
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_anim_table), sizeof(float) * table_len));
	}
	if (!ctx->d_base_y)
		HVCF_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->d_base_y), sizeof(float) * (ctx->max_lights ? ctx->max_lights : 1)));
	HVCF_CUDA(cudaMemcpyAsync(ctx->d_anim_table, table, sizeof(float) * table_len, cudaMemcpyHostToDevice, s));
	HVCF_CUDA(cudaStreamSynchronize(s)); // `table` may be pageable and go away
	const hvcf_light_data* lights = ctx->bound_lights ? ctx->bound_lights : ctx->d_lights;
	HVCF_CUDA(launch_capture_base_y(lights, ctx->num_lights, ctx->d_base_y, s));
	ctx->base_y_count = ctx->num_lights;
	ctx->anim_len = table_len;
	ctx->anim_amplitude = amplitude;
	return HVCF_OK;
}

int hvcf_set_animation_step(hvcf_ctx* ctx, uint32_t frame, uint32_t step)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_animation_step: null context");
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_animation_step: frame %u >= frame_count %u", frame, ctx->frame_count);
	ctx->anim_step[frame] = step;
	return HVCF_OK;
}

int hvcf_set_ubo(hvcf_ctx* ctx, uint32_t frame, const hvcf_cluster_forward_ubo* ubo, void* stream)
{
	if (!ctx || !ubo)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_ubo: null argument");
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_ubo: frame %u >= frame_count %u", frame, ctx->frame_count);
	if (ubo->slice_count_x != ctx->X || ubo->slice_count_y != ctx->Y || ubo->slice_count_z != ctx->Z)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_set_ubo: ubo grid %ux%ux%u != context grid %ux%ux%u", ubo->slice_count_x,
			ubo->slice_count_y, ubo->slice_count_z, ctx->X, ctx->Y, ctx->Z);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	// the slot's previous copy out of pinned staging must have executed before the slot is overwritten: the fence the
	// reference takes at frame start (vkWaitForFences, AppBase.cpp:96-100), scoped to the UBO slot
	if (!ctx->ubo_events[frame])
		HVCF_CUDA(cudaEventCreateWithFlags(&ctx->ubo_events[frame], cudaEventDisableTiming));
	else
		HVCF_CUDA(cudaEventSynchronize(ctx->ubo_events[frame]));
	ctx->h_ubo[frame] = *ubo;
	HVCF_CUDA(cudaMemcpyAsync(&ctx->d_ubo[frame], &ctx->h_ubo[frame], sizeof(hvcf_cluster_forward_ubo), cudaMemcpyHostToDevice,
		static_cast<cudaStream_t>(stream)));
	HVCF_CUDA(cudaEventRecord(ctx->ubo_events[frame], static_cast<cudaStream_t>(stream)));
	ctx->ubo_set[frame] = true;
	// AABBs depend on the inverse projection, near/far and the grid only
	float key[18];
	memcpy(key, ubo->camera_inverse_projection, 64);
	key[16] = ubo->camera_near;
	key[17] = ubo->camera_far;
	if (!ctx->have_aabb_key || memcmp(key, ctx->aabb_key, sizeof(key)) != 0)
		ctx->aabb_dirty = true;
	return HVCF_OK;
}

int hvcf_mark_aabb_dirty(hvcf_ctx* ctx)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_mark_aabb_dirty: null context");
	ctx->aabb_dirty = true;
	return HVCF_OK;
}

int hvcf_is_aabb_dirty(hvcf_ctx* ctx, int* out_dirty)
{
	if (!ctx || !out_dirty)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_is_aabb_dirty: null argument");
	*out_dirty = ctx->aabb_dirty ? 1 : 0;
	return HVCF_OK;
}

int hvcf_generate_aabbs(hvcf_ctx* ctx, uint32_t frame, void* stream)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_generate_aabbs: null context");
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_generate_aabbs: frame %u >= frame_count %u", frame, ctx->frame_count);
	if (!ctx->ubo_set[frame])
		return fail(HVCF_ERR_NOT_READY, "hvcf_generate_aabbs: hvcf_set_ubo(frame %u) has not been called", frame);
	NvtxZone zone("AABB_Generator");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	const hvcf_cluster_forward_ubo& u = ctx->h_ubo[frame];
	// slice-plane table: tileNear(z) = -near * pow(far/near, float(z)/Z)  (AABBGenerator.comp:72-75)
	float* hz = ctx->h_zplanes + static_cast<size_t>(frame) * (ctx->Z + 1);
	for (uint32_t z = 0; z <= ctx->Z; ++z)
		hz[z] = -u.camera_near * powf(u.camera_far / u.camera_near, static_cast<float>(z) / static_cast<float>(ctx->Z));
	float* dz = ctx->d_zplanes + static_cast<size_t>(frame) * (ctx->Z + 1);
	HVCF_CUDA(cudaMemcpyAsync(dz, hz, sizeof(float) * (ctx->Z + 1), cudaMemcpyHostToDevice, s));
	HVCF_CUDA(launch_aabb_generate(&ctx->d_ubo[frame], dz, ctx->d_aabb, ctx->cluster_total, s));
	if (ctx->hier) // level-0 union boxes of the cluster groups follow the AABBs
	{
		HVCF_CUDA(launch_hier_unions(ctx->d_hlane_cluster, ctx->d_aabb, ctx->cluster_begin, ctx->hp.K, ctx->hp.ncb, ctx->d_cb_union, s));
		if (ctx->bin_ns)
			HVCF_CUDA(launch_hier_super_unions(ctx->d_cb_union, ctx->hp.ncb, ctx->bin_sg, ctx->bin_ns, ctx->d_sb_union, s));
	}
	if (ctx->shard.on) // the level above the groups: union box of every rank's slab
		HVCF_CUDA(launch_slab_unions(ctx->d_aabb, ctx->shard.d_ranges, ctx->shard.world, ctx->shard.d_slab_box, s));
	memcpy(ctx->aabb_key, u.camera_inverse_projection, 64);
	ctx->aabb_key[16] = u.camera_near;
	ctx->aabb_key[17] = u.camera_far;
	ctx->have_aabb_key = true;
	ctx->aabb_dirty = false;
	return HVCF_OK;
}

// ---- hierarchical frame, in three steps so that the light exchange of the sharded grid can be pipelined with it:
//   open   AABBs if dirty, light-segment plan
//   range  transform + level-0 binning of the lights [begin, end) (whole tiles; the last range takes the padding)
//   close  count pass + finalize (per-cluster counts, offsets, shard total)
cudaError_t hvcf_kmark(hvcf_ctx* ctx, uint32_t frame, int i, cudaStream_t s)
{
	if (!ctx->ktimes)
		return cudaSuccess;
	if (!ctx->kev[frame][i])
	{
		cudaError_t e = cudaEventCreate(&ctx->kev[frame][i]);
		if (e != cudaSuccess)
			return e;
	}
	return cudaEventRecord(ctx->kev[frame][i], s);
}

int hvcf_cull_open(hvcf_ctx* ctx, uint32_t frame, void* stream, const char* who, bool routed, bool allow_rounds)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: null context", who);
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "%s: frame %u >= frame_count %u", who, frame, ctx->frame_count);
	if (!ctx->ubo_set[frame])
		return fail(HVCF_ERR_NOT_READY, "%s: hvcf_set_ubo(frame %u) has not been called", who, frame);
	if (!routed && ctx->shard.on && ctx->bound_lights)
		return fail(HVCF_ERR_UNSUPPORTED, "%s: the bound buffer of a sharded context holds the owned lights only (hvcf_cull_routed reads it)", who);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	ctx->ev_valid[frame] = false;
	ctx->open_launches = 0;
	ctx->routed_frame = routed;
	ctx->kev_exchange[frame] = false;
	if (ctx->timing)
		HVCF_CUDA(cudaEventRecord(ctx->ev[frame][0], s));
	if (ctx->aabb_dirty)
	{
		const int rc = hvcf_generate_aabbs(ctx, frame, stream);
		if (rc != HVCF_OK)
			return rc;
		++ctx->open_launches;
	}
	if (ctx->hier)
	{
		// at least one (all-padding) tile, so that pass 2 always writes the light cells
		const uint32_t n = ctx->num_lights;
		const uint32_t padded = round_up_u32(n ? n : 1, kTileLights);
		const uint32_t tiles_total = routed ? ctx->shard.tiles_phys : padded / kTileLights; // routed: the tiles of the inbox
		hier_plan_lights(ctx->hp, tiles_total, ctx->h_plan_max_segs, ctx->h_tps_override, &ctx->h_tiles_per_seg, &ctx->h_num_segs);
		ctx->h_pipe = ctx->pipe_mode >= 0 ? (ctx->pipe_mode ? 1u : 0u) : (ctx->h_tiles_per_seg >= ctx->pipe_min_tps ? 1u : 0u);
		// sub-tile segments: about 128 light segments per frame (measured: 64 tiles x 2 parts 0.0516 -> 0.0478 ms, x 4 0.0506;
		// 1 tile x 8 parts 0.0347 -> 0.0252 ms), at most kHMaxParts per tile; HVCF_HIER_PARTS fixes the number (1 = off)
		ctx->h_parts = 1;
		if (ctx->h_tiles_per_seg == 1 && !ctx->h_pipe && !routed)
		{
			uint32_t parts = ctx->h_parts_cfg ? ctx->h_parts_cfg : 128u / ctx->h_num_segs;
			parts = parts > kHMaxParts ? kHMaxParts : parts;
			while (parts > 1 && static_cast<uint64_t>(ctx->h_num_segs) * parts > ctx->h_max_segs)
				--parts;
			if (parts > 1)
			{
				ctx->h_parts = parts;
				ctx->h_num_segs *= parts;
			}
		}
		ctx->h_tiles_total = tiles_total;
		ctx->h_num_rounds = 0;
		ctx->last_plan_flags = (ctx->h_pipe ? HVCF_PLAN_PIPELINED_WALK : 0u) | (ctx->h_parts > 1 ? HVCF_PLAN_SUBTILE_SEGMENTS : 0u) |
			(routed ? HVCF_PLAN_TWO_LEVEL_BIN : 0u);
		if (allow_rounds && ctx->h_parts == 1 && ctx->rounds_mode != 0 && tiles_total >= 2 &&
			(ctx->rounds_mode > 0 || tiles_total >= ctx->rounds_min_tiles))
		{
			// rounds of growing length: head, head * growth, ...; the last one takes the rest.  Early rounds are cut into short
			// segments (that is where clusters fill up: most of the frame's hits, spread over many warp programs), late ones into
			// few long segments (sparse: only the groups that are still alive are walked at all).
			const uint32_t max_rounds = ctx->rounds_mode > 0 ? (static_cast<uint32_t>(ctx->rounds_mode) < kHMaxChunks ? static_cast<uint32_t>(ctx->rounds_mode) : kHMaxChunks) : kHMaxChunks;
			const uint32_t growth = ctx->rounds_growth < 2 ? 2 : ctx->rounds_growth;
			uint32_t t = 0, size = ctx->rounds_head ? ctx->rounds_head : 1, nr = 0, seg = 0;
			while (t < tiles_total && nr < max_rounds)
			{
				uint32_t end = static_cast<uint64_t>(t) + size < tiles_total ? t + size : tiles_total;
				if (nr + 1 == max_rounds || tiles_total - end < size) // the last round takes the rest (no tiny tail round)
					end = tiles_total;
				const uint32_t target = nr < 2 ? ctx->rounds_early_segs : ctx->rounds_late_segs;
				uint32_t tps = (end - t + target - 1) / target;
				// macro tiles (two-level binning only, pipelined walk only): late rounds are sparse, so M consecutive tiles share one
				// slot per group.  M by what keeps about four binning CTAs per SM, at most 64, and segments are whole macro tiles.
				uint32_t mshift = 0;
				const bool two_level = ctx->bin_ns && (ctx->bin2_mode > 0 || ctx->hp.ncb >= kBin2MinGroups ||
					static_cast<uint64_t>(ctx->hp.ncb) * ctx->num_lights >= (1ull << 28));
				if (nr >= 2 && two_level && ctx->pipe_mode != 0)
				{
					uint32_t want = ctx->rounds_macro >= 0 ? static_cast<uint32_t>(ctx->rounds_macro) : (end - t) / 592u;
					want = want > 64u ? 64u : want;
					if (routed) // a routed frame's macro tile is a route block: its records are dense from the block's first inbox tile
						want = ctx->rounds_macro == 0 || t % ctx->shard.block_tiles ? 1u : ctx->shard.block_tiles;
					while ((2u << mshift) <= want)
						++mshift;
					tps = ((tps + (1u << mshift) - 1u) >> mshift) << mshift;
				}
				hvcf_ctx::Round& r = ctx->h_rounds[nr];
				r.macro_shift = mshift;
				r.tile_begin = t, r.tile_end = end, r.tiles_per_seg = tps;
				r.seg_begin = seg;
				seg += (end - t + tps - 1) / tps;
				r.seg_end = seg;
				r.pipe = mshift ? 1u : (ctx->pipe_mode >= 0 ? (ctx->pipe_mode ? 1u : 0u) : (tps >= ctx->pipe_min_tps ? 1u : 0u));
				t = end;
				size = static_cast<uint64_t>(size) * growth > 0x7fffffffu ? 0x7fffffffu : size * growth;
				++nr;
			}
			if (nr >= 2 && seg <= ctx->h_max_segs)
			{
				ctx->h_num_rounds = nr;
				ctx->h_num_segs = seg;
				ctx->last_plan_flags &= ~HVCF_PLAN_PIPELINED_WALK;
				for (uint32_t r = 0; r < nr; ++r)
					ctx->last_plan_flags |= (ctx->h_rounds[r].pipe ? HVCF_PLAN_PIPELINED_WALK : 0u) | (ctx->h_rounds[r].macro_shift ? HVCF_PLAN_MACRO_TILES : 0u);
			}
		}
		if (ctx->h_num_rounds < 2 && hier_num_chunks(ctx->h_num_segs) > kHMaxChunks)
			return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull: %u light segments need more than %u early-exit chunks", ctx->h_num_segs, kHMaxChunks);
		HVCF_CUDA(hvcf_kmark(ctx, frame, 0, s));
	}
	ctx->range_frame = frame;
	ctx->range_transformed = ctx->range_binned = 0;
	ctx->range_cleared = false;
	ctx->range_open = true;
	return HVCF_OK;
}

static int hier_range(hvcf_ctx* ctx, uint32_t frame, uint32_t begin, uint32_t end, uint32_t flags, cudaStream_t s)
{
	const uint32_t n = ctx->num_lights;
	const bool last = end >= n;
	const uint32_t end_padded = last ? round_up_u32(n ? n : 1, kTileLights) : end;
	DeviceStats* stats = &ctx->d_stats[frame];
	// ResetGlobalIndex (PipelineLightCulling.cpp:38-43) and the early-exit scratch are cleared by the first transform launch
	uint32_t* zero = ctx->range_cleared ? nullptr : ctx->d_hscratch;
	const uint32_t zero_words = ctx->range_cleared ? 0u : ctx->cluster_count * kHMaxChunks;
	DeviceStats* clear_stats = ctx->range_cleared ? nullptr : stats;
	// few lights, whole frame in one call, one-level binning, no animation: transform + binning as one kernel
	if (flags == (HVCF_RANGE_TRANSFORM | HVCF_RANGE_BIN) && begin == 0 && last && !ctx->range_cleared && !ctx->anim_len &&
		ctx->xbin_max_lights && n <= ctx->xbin_max_lights &&
		!(ctx->bin_ns && (ctx->bin2_mode > 0 || ctx->hp.ncb >= kBin2MinGroups || static_cast<uint64_t>(ctx->hp.ncb) * n >= (1ull << 28))))
	{
		XBinParams q{};
		q.bin.lights_view = ctx->d_lights_view;
		q.bin.cb_union = ctx->d_cb_union;
		q.bin.ncb = ctx->hp.ncb;
		q.bin.tiles_cap = ctx->tiles_cap;
		q.bin.bin_count = ctx->d_bin_count;
		q.bin.bin_rec = ctx->d_bin_rec;
		q.bin.bin_idx = ctx->d_bin_idx;
		q.bin.stats = stats;
		q.ubo = &ctx->d_ubo[frame];
		q.lights = ctx->bound_lights ? ctx->bound_lights : ctx->d_lights;
		q.num_lights = n;
		q.zero = zero;
		q.zero_words = zero_words;
		HVCF_CUDA(launch_hier_xbin(q, end_padded / kTileLights, s));
		ctx->last_plan_flags |= HVCF_PLAN_FUSED_BIN;
		ctx->range_cleared = true;
		ctx->range_transformed = ctx->range_binned = end_padded;
		++ctx->open_launches;
		HVCF_CUDA(hvcf_kmark(ctx, frame, 1, s));
		HVCF_CUDA(hvcf_kmark(ctx, frame, 2, s));
		return HVCF_OK;
	}
	if (flags & HVCF_RANGE_TRANSFORM)
	{
		const hvcf_light_data* lights = ctx->bound_lights ? ctx->bound_lights : ctx->d_lights;
		LightAnim anim_storage{};
		const LightAnim* anim = nullptr;
		if (ctx->anim_len && n)
		{
			anim_storage.base_y = ctx->d_base_y + begin;
			anim_storage.table = ctx->d_anim_table;
			anim_storage.mask = ctx->anim_len - 1;
			anim_storage.step = ctx->anim_step[frame] + begin;
			anim_storage.amplitude = ctx->anim_amplitude;
			anim_storage.lights_rw = const_cast<hvcf_light_data*>(lights) + begin; // the moved position is written back, as UpdateLightPosition does
			anim = &anim_storage;
		}
		HVCF_CUDA(launch_light_transform(&ctx->d_ubo[frame], lights + begin, n - begin, end_padded - begin, ctx->d_lights_view + begin, zero,
			zero_words, clear_stats, s, anim));
		ctx->range_cleared = true;
		ctx->range_transformed = end_padded;
		++ctx->open_launches;
		HVCF_CUDA(hvcf_kmark(ctx, frame, 1, s));
	}
	if (flags & HVCF_RANGE_BIN)
	{
		if (!ctx->range_cleared) // the view records came from elsewhere: a clear-only launch
		{
			HVCF_CUDA(launch_light_transform(&ctx->d_ubo[frame], nullptr, 0, 0, ctx->d_lights_view, zero, zero_words, clear_stats, s, nullptr));
			ctx->range_cleared = true;
			++ctx->open_launches;
		}
		{
			BinParams b{};
			b.lights_view = ctx->d_lights_view;
			b.cb_union = ctx->bin_alive ? ctx->d_alive_union : ctx->d_cb_union; // rounds: the boxes of the clusters that are still alive
			b.ncb = ctx->hp.ncb;
			b.tiles_cap = ctx->tiles_cap;
			b.bin_count = ctx->d_bin_count;
			b.bin_rec = ctx->d_bin_rec;
			b.bin_idx = ctx->d_bin_idx;
			b.stats = stats;
			b.tile_begin = begin / kTileLights;
			b.group_alive = ctx->bin_alive;
			b.macro = ctx->bin_macro;
			b.sb_union = ctx->d_sb_union;
			b.sg = ctx->bin_sg;
			// the two-level kernel pays once there are many (group, light) pairs: >= 48 groups, or >= 2^28 pairs (measured:
			// grid A with 18 groups loses with it up to 1M lights; every z-slab of grid C at 16M lights wins, slab_sweep.py)
			const bool two_level = ctx->bin_ns && (ctx->bin2_mode > 0 || ctx->hp.ncb >= kBin2MinGroups ||
				static_cast<uint64_t>(ctx->hp.ncb) * n >= (1ull << 28));
			b.ns = two_level ? ctx->bin_ns : 0u;
			ctx->last_plan_flags |= two_level ? HVCF_PLAN_TWO_LEVEL_BIN : 0u;
			HVCF_CUDA(launch_hier_bin(b, (end_padded - begin) / kTileLights, s));
			++ctx->open_launches;
		}
		ctx->range_binned = end_padded;
		HVCF_CUDA(hvcf_kmark(ctx, frame, 2, s));
	}
	return HVCF_OK;
}

// count pass of one round (round < 0: the whole frame in one launch, or head + rest with HVCF_HIER_SPLIT)
static int hier_count(hvcf_ctx* ctx, uint32_t frame, int round, cudaStream_t s, uint32_t* launches)
{
	HierParams p = hier_params(ctx, frame);
	if (round >= 0)
	{
		const hvcf_ctx::Round& r = ctx->h_rounds[round];
		p.seg_begin = r.seg_begin, p.seg_end = r.seg_end;
		p.tile_begin = r.tile_begin, p.tiles_total = r.tile_end;
		p.tiles_per_seg = r.tiles_per_seg, p.pipe = r.pipe;
		p.macro_shift = r.macro_shift;
		p.round_chunk = static_cast<uint32_t>(round);
		p.round_alive = round > 0 ? ctx->d_group_alive + static_cast<size_t>(round) * ctx->hp.ncb : nullptr;
		HVCF_CUDA(launch_hier2(static_cast<int>(ctx->hp.K), false, p, s));
		++*launches;
		return HVCF_OK;
	}
	// the first segments run as a launch of their own when the whole grid would otherwise start at once: the
	// later segments then begin with the dense clusters already known to be full (PDL-chained, one extra launch)
	const uint32_t head = ctx->h_split_head < ctx->h_num_segs ? ctx->h_split_head : 0u;
	if (head)
	{
		p.seg_end = head;
		HVCF_CUDA(launch_hier2(static_cast<int>(ctx->hp.K), false, p, s));
		p.seg_begin = head;
		p.tile_begin = head * p.tiles_per_seg;
		p.seg_end = ctx->h_num_segs;
		++*launches;
	}
	HVCF_CUDA(launch_hier2(static_cast<int>(ctx->hp.K), false, p, s));
	++*launches;
	return HVCF_OK;
}

static int hier_finalize(hvcf_ctx* ctx, uint32_t frame, cudaStream_t s)
{
	FinalizeParams f{};
	f.seg_counts = ctx->d_hseg_counts;
	f.seg_prefix = ctx->d_hseg_prefix;
	f.num_segs = ctx->h_num_segs;
	f.cluster_count = ctx->cluster_count;
	f.cap = ctx->cap;
	f.scan_state = ctx->d_scan_state + static_cast<size_t>(frame) * ((ctx->cluster_count + 31) / 32); // per frame slot: tags are per-slot serials
	f.counts = ctx->d_counts;
	f.offsets = ctx->d_hoffsets;
	f.shard_total = ctx->d_shard_total;
	f.stats = &ctx->d_stats[frame];
	if (ctx->h_num_rounds >= 2)
	{
		f.cum = ctx->d_hscratch;
		f.num_rounds = ctx->h_num_rounds;
		for (uint32_t r = 0; r < ctx->h_num_rounds; ++r)
			f.round_seg_begin[r] = ctx->h_rounds[r].seg_begin;
		f.round_seg_begin[ctx->h_num_rounds] = ctx->h_num_segs;
	}
	HVCF_CUDA(launch_hier_finalize(f, s));
	HVCF_CUDA(hvcf_kmark(ctx, frame, 4, s));
	return HVCF_OK;
}

static cudaError_t round_mark(hvcf_ctx* ctx, uint32_t frame, uint32_t round, int i, cudaStream_t s)
{
	if (!ctx->ktimes)
		return cudaSuccess;
	cudaEvent_t& e = ctx->kev_round[frame][round][i];
	if (!e)
	{
		const cudaError_t rc = cudaEventCreate(&e);
		if (rc != cudaSuccess)
			return rc;
	}
	return cudaEventRecord(e, s);
}

// rounds: level-0 binning (against the groups that the earlier rounds left alive) + count pass, round by round, over the tiles
// that have been transformed (or, routed frames, have arrived in the inbox) already.  bin_round(round) launches the round's binning.
int hvcf_hier_rounds(hvcf_ctx* ctx, uint32_t frame, cudaStream_t s, uint32_t* launches, const std::function<int(uint32_t)>& bin_round)
{
	for (uint32_t r = 0; r < ctx->h_num_rounds; ++r)
	{
		HVCF_CUDA(round_mark(ctx, frame, r, 0, s));
		ctx->bin_alive = nullptr;
		if (r > 0)
		{
			uint32_t* flags = ctx->d_group_alive + static_cast<size_t>(r) * ctx->hp.ncb; // kept per round: the write pass asks again
			HVCF_CUDA(launch_hier_alive(ctx->d_hlane_cluster, ctx->d_aabb, ctx->cluster_begin, ctx->d_hscratch, ctx->hp.K, ctx->hp.ncb, ctx->cap, flags,
				ctx->d_alive_union, s));
			ctx->bin_alive = flags;
			++*launches;
		}
		ctx->bin_macro = 1u << ctx->h_rounds[r].macro_shift;
		const int rc = bin_round(r);
		ctx->bin_alive = nullptr;
		ctx->bin_macro = 1;
		if (rc != HVCF_OK)
			return rc;
		++*launches;
		HVCF_CUDA(round_mark(ctx, frame, r, 1, s));
		const int rc2 = hier_count(ctx, frame, static_cast<int>(r), s, launches);
		if (rc2 != HVCF_OK)
			return rc2;
		HVCF_CUDA(round_mark(ctx, frame, r, 2, s));
	}
	return HVCF_OK;
}

int hvcf_hier_close(hvcf_ctx* ctx, uint32_t frame, cudaStream_t s)
{
	if (ctx->timing)
		HVCF_CUDA(cudaEventRecord(ctx->ev[frame][1], s));
	uint32_t launches = ctx->open_launches;
	if (ctx->h_num_rounds < 2)
	{
		const int rc = hier_count(ctx, frame, -1, s, &launches);
		if (rc != HVCF_OK)
			return rc;
	}
	HVCF_CUDA(hvcf_kmark(ctx, frame, 3, s));
	if (ctx->timing)
	{
		HVCF_CUDA(cudaEventRecord(ctx->ev[frame][2], s));
		ctx->ev_valid[frame] = true;
	}
	const int rc = hier_finalize(ctx, frame, s);
	if (rc != HVCF_OK)
		return rc;
	ctx->last_launches[frame] = launches + 1;
	ctx->range_open = false;
	return HVCF_OK;
}

int hvcf_cull_begin_range(hvcf_ctx* ctx, uint32_t frame, uint32_t light_begin, uint32_t light_end, uint32_t flags, void* stream)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_begin_range: null context");
	if (!ctx->hier)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_cull_begin_range: hierarchical mode only");
	if (flags == 0)
		flags = HVCF_RANGE_TRANSFORM | HVCF_RANGE_BIN;
	if (flags & ~static_cast<uint32_t>(HVCF_RANGE_TRANSFORM | HVCF_RANGE_BIN))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_begin_range: unknown flags 0x%x", flags);
	const uint32_t n = ctx->num_lights;
	if (light_begin % kTileLights || light_begin > light_end || light_end > n || (light_end < n && light_end % kTileLights) ||
		(light_begin == light_end && n != 0))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_begin_range: [%u,%u) of %u lights: ranges are whole %u-light tiles, the last one ends at the light count",
			light_begin, light_end, n, kTileLights);
	const bool first = light_begin == 0 && (!ctx->range_open || ctx->range_frame != frame ||
		((flags & HVCF_RANGE_TRANSFORM) ? ctx->range_transformed != 0 : ctx->range_binned != 0));
	if (first)
	{
		const int rc = hvcf_cull_open(ctx, frame, stream, "hvcf_cull_begin_range", false, false);
		if (rc != HVCF_OK)
			return rc;
	}
	if (!ctx->range_open || ctx->range_frame != frame)
		return fail(HVCF_ERR_NOT_READY, "hvcf_cull_begin_range: the first range of a frame starts at light 0");
	if ((flags & HVCF_RANGE_TRANSFORM) && light_begin != ctx->range_transformed)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_begin_range: transform ranges must be consecutive (next light %u, got %u)", ctx->range_transformed, light_begin);
	if ((flags & HVCF_RANGE_BIN) && light_begin != ctx->range_binned)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_begin_range: bin ranges must be consecutive (next light %u, got %u)", ctx->range_binned, light_begin);
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	const int rc = hier_range(ctx, frame, light_begin, light_end, flags, s);
	if (rc != HVCF_OK)
		return rc;
	if ((flags & HVCF_RANGE_BIN) && light_end >= n)
		return hvcf_hier_close(ctx, frame, s);
	return HVCF_OK;
}

int hvcf_transform_lights(hvcf_ctx* ctx, uint32_t frame, uint32_t light_begin, uint32_t light_end, void* stream)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_transform_lights: null context");
	if (frame >= ctx->frame_count || !ctx->ubo_set[frame])
		return fail(HVCF_ERR_NOT_READY, "hvcf_transform_lights: hvcf_set_ubo(frame %u) has not been called", frame);
	const uint32_t n = ctx->num_lights;
	const uint32_t padded = round_up_u32(n ? n : 1, kTileLights);
	if (light_begin % kTileLights || light_begin > light_end || light_begin >= padded || (light_end < n && light_end % kTileLights) ||
		light_end > padded)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_transform_lights: [%u,%u) of %u lights: whole %u-light tiles", light_begin, light_end, n,
			kTileLights);
	HVCF_CUDA(cudaSetDevice(ctx->device));
	const uint32_t end_padded = light_end >= n ? padded : light_end;
	const hvcf_light_data* lights = ctx->bound_lights ? ctx->bound_lights : ctx->d_lights;
	LightAnim anim_storage{};
	const LightAnim* anim = nullptr;
	if (ctx->anim_len && n)
	{
		anim_storage.base_y = ctx->d_base_y + light_begin;
		anim_storage.table = ctx->d_anim_table;
		anim_storage.mask = ctx->anim_len - 1;
		anim_storage.step = ctx->anim_step[frame] + light_begin;
		anim_storage.amplitude = ctx->anim_amplitude;
		anim_storage.lights_rw = const_cast<hvcf_light_data*>(lights) + light_begin;
		anim = &anim_storage;
	}
	HVCF_CUDA(launch_light_transform(&ctx->d_ubo[frame], lights + light_begin, n > light_begin ? n - light_begin : 0u, end_padded - light_begin,
		ctx->d_lights_view + light_begin, nullptr, 0, nullptr, static_cast<cudaStream_t>(stream), anim));
	return HVCF_OK;
}

int hvcf_lights_view_ptr(hvcf_ctx* ctx, void** out_device_ptr, size_t* out_bytes)
{
	if (!ctx || !out_device_ptr)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_lights_view_ptr: null argument");
	*out_device_ptr = ctx->d_lights_view;
	if (out_bytes)
		*out_bytes = static_cast<size_t>(ctx->padded_capacity) * 16;
	return HVCF_OK;
}

int hvcf_cull_begin(hvcf_ctx* ctx, uint32_t frame, void* stream)
{
	NvtxZone zone("Light_Culling");
	int rc = hvcf_cull_open(ctx, frame, stream, "hvcf_cull_begin", false, true);
	if (rc != HVCF_OK)
		return rc;
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	uint32_t launches = ctx->open_launches;
	DeviceStats* stats = &ctx->d_stats[frame];
	const uint32_t n = ctx->num_lights;
	if (ctx->hier)
	{
		if (ctx->h_num_rounds >= 2)
		{
			rc = hier_range(ctx, frame, 0, n, HVCF_RANGE_TRANSFORM, s); // every light to view space (and the frame's counters cleared)
			if (rc != HVCF_OK)
				return rc;
			rc = hvcf_hier_rounds(ctx, frame, s, &ctx->open_launches, [&](uint32_t r) {
				const hvcf_ctx::Round& rd = ctx->h_rounds[r];
				const uint64_t le = static_cast<uint64_t>(rd.tile_end) * kTileLights;
				--ctx->open_launches; // hier_range counts the launch itself
				return hier_range(ctx, frame, rd.tile_begin * kTileLights, le < n ? static_cast<uint32_t>(le) : n, HVCF_RANGE_BIN, s);
			});
			if (rc != HVCF_OK)
				return rc;
			return hvcf_hier_close(ctx, frame, s);
		}
		rc = hier_range(ctx, frame, 0, n, HVCF_RANGE_TRANSFORM | HVCF_RANGE_BIN, s);
		if (rc != HVCF_OK)
			return rc;
		return hvcf_hier_close(ctx, frame, s);
	}
	ctx->range_open = false;
	HVCF_CUDA(cudaMemsetAsync(stats, 0, sizeof(DeviceStats), s)); // ResetGlobalIndex (PipelineLightCulling.cpp:38-43)
	const hvcf_light_data* lights = ctx->bound_lights ? ctx->bound_lights : ctx->d_lights;
	LightAnim anim_storage{};
	const LightAnim* anim = nullptr;
	if (ctx->anim_len && n)
	{
		anim_storage.base_y = ctx->d_base_y;
		anim_storage.table = ctx->d_anim_table;
		anim_storage.mask = ctx->anim_len - 1;
		anim_storage.step = ctx->anim_step[frame];
		anim_storage.amplitude = ctx->anim_amplitude;
		anim_storage.lights_rw = const_cast<hvcf_light_data*>(lights); // the moved position is written back, as UpdateLightPosition does
		anim = &anim_storage;
	}
	const uint32_t padded = round_up_u32(n, kTileLights);
	const uint32_t tiles_total = padded / kTileLights;
	uint32_t num_segs = ctx->max_segs < tiles_total ? ctx->max_segs : tiles_total;
	uint32_t tiles_per_seg = num_segs ? (tiles_total + num_segs - 1) / num_segs : 0;
	if (tiles_per_seg)
		num_segs = (tiles_total + tiles_per_seg - 1) / tiles_per_seg; // no empty segment
	if (n)
	{
		HVCF_CUDA(launch_light_transform(&ctx->d_ubo[frame], lights, n, padded, ctx->d_lights_view, nullptr, 0, nullptr, s, anim));
		CullParams p;
		p.lights_view = ctx->d_lights_view;
		p.aabbs = ctx->d_aabb;
		p.cluster_begin = ctx->cluster_begin;
		p.cluster_count = ctx->cluster_count;
		p.tiles_total = tiles_total;
		p.tiles_per_seg = tiles_per_seg;
		p.cap = ctx->cap;
		p.seg_counts = ctx->d_seg_counts;
		p.seg_lists = ctx->d_seg_lists;
		p.stats = stats;
		if (ctx->timing)
			HVCF_CUDA(cudaEventRecord(ctx->ev[frame][1], s));
		HVCF_CUDA(launch_cull(p, num_segs, s));
		if (ctx->timing)
		{
			HVCF_CUDA(cudaEventRecord(ctx->ev[frame][2], s));
			ctx->ev_valid[frame] = true;
		}
		launches += 2;
	}
	ctx->last_num_segs = num_segs;
	HVCF_CUDA(launch_count(ctx->d_seg_counts, num_segs, ctx->cluster_count, ctx->cap, ctx->d_counts, ctx->d_block_sums, stats, s));
	HVCF_CUDA(launch_scan(ctx->d_block_sums, ctx->count_blocks, ctx->d_block_offsets, ctx->d_shard_total, stats, s));
	launches += 2;
	ctx->last_launches[frame] = launches;
	return HVCF_OK;
}

int hvcf_cull_finish(hvcf_ctx* ctx, uint32_t frame, const uint32_t* d_index_base, void* stream)
{
	if (!ctx)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_finish: null context");
	if (frame >= ctx->frame_count)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_finish: frame %u >= frame_count %u", frame, ctx->frame_count);
	NvtxZone zone("Light_Culling");
	HVCF_CUDA(cudaSetDevice(ctx->device));
	if (ctx->hier)
	{
		HierParams hp = hier_params(ctx, frame);
		hp.index_base = d_index_base;
		if (ctx->h_num_rounds >= 2)
		{
			// one launch per round (their segments differ in length, and long ones take the pipelined walk).  They are independent of
			// each other and the late ones are latency-bound (few live groups, long walks), so the rounds after the first go to side
			// streams and run UNDER the first round's launch, which writes most of the frame's indices: fork after finalize, join
			// before the call returns to the caller's stream (grid C x 16M: 198 + 69 + 74 + 75 us in a row -> about the first alone)
			cudaStream_t s0 = static_cast<cudaStream_t>(stream);
			// (not for shard contexts that share this process with their peers -- the test harness of hvcf_peer_attach_local: four
			// streams per context exceed the device's hardware queues, and a write pass queued behind another shard's spinning wait
			// kernel is a false dependency that only the 4 s bound breaks)
			const bool fork = ctx->write_fork && ctx->h_num_rounds >= 2 && !(ctx->peer.world > 1 && !ctx->peer_ipc);
			if (fork)
			{
				if (!ctx->write_ev[0])
				{
					for (auto& st : ctx->write_streams)
						HVCF_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
					for (auto& e : ctx->write_ev)
						HVCF_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
				}
				HVCF_CUDA(cudaEventRecord(ctx->write_ev[0], s0));
			}
			for (uint32_t r = 0; r < ctx->h_num_rounds; ++r)
			{
				const hvcf_ctx::Round& rd = ctx->h_rounds[r];
				hp.seg_begin = rd.seg_begin, hp.seg_end = rd.seg_end;
				hp.tile_begin = rd.tile_begin, hp.tiles_total = rd.tile_end;
				hp.tiles_per_seg = rd.tiles_per_seg, hp.pipe = rd.pipe;
				hp.macro_shift = rd.macro_shift;
				hp.round_alive = r > 0 ? ctx->d_group_alive + static_cast<size_t>(r) * ctx->hp.ncb : nullptr;
				cudaStream_t sr = s0;
				if (fork && r > 0)
				{
					sr = ctx->write_streams[(r - 1) % 3];
					if (r <= 3)
						HVCF_CUDA(cudaStreamWaitEvent(sr, ctx->write_ev[0], 0));
				}
				HVCF_CUDA(launch_hier2(static_cast<int>(ctx->hp.K), true, hp, sr));
			}
			if (fork)
				for (uint32_t i = 0; i < 3 && i + 1 < ctx->h_num_rounds; ++i)
				{
					HVCF_CUDA(cudaEventRecord(ctx->write_ev[1 + i], ctx->write_streams[i]));
					HVCF_CUDA(cudaStreamWaitEvent(s0, ctx->write_ev[1 + i], 0));
				}
			ctx->last_launches[frame] += ctx->h_num_rounds - 1;
		}
		else
			HVCF_CUDA(launch_hier2(static_cast<int>(ctx->hp.K), true, hp, static_cast<cudaStream_t>(stream)));
		if (ctx->ktimes && ctx->kev[frame][4])
		{
			if (!ctx->kev[frame][5])
				HVCF_CUDA(cudaEventCreate(&ctx->kev[frame][5]));
			HVCF_CUDA(cudaEventRecord(ctx->kev[frame][5], static_cast<cudaStream_t>(stream)));
		}
		if (ctx->timing)
			HVCF_CUDA(cudaEventRecord(ctx->ev[frame][3], static_cast<cudaStream_t>(stream)));
		ctx->last_launches[frame] += 1;
		return HVCF_OK;
	}
	ScatterParams p;
	p.counts = ctx->d_counts;
	p.block_offsets = ctx->d_block_offsets;
	p.index_base = d_index_base;
	p.seg_counts = ctx->d_seg_counts;
	p.seg_lists = ctx->d_seg_lists;
	p.num_segs = ctx->last_num_segs;
	p.cluster_begin = ctx->cluster_begin;
	p.cluster_count = ctx->cluster_count;
	p.cap = ctx->cap;
	p.cells = ctx->d_cells;
	p.indices = ctx->d_indices;
	p.frame_total = &ctx->d_global_index[frame];
	HVCF_CUDA(launch_scatter(p, static_cast<cudaStream_t>(stream)));
	if (ctx->timing)
		HVCF_CUDA(cudaEventRecord(ctx->ev[frame][3], static_cast<cudaStream_t>(stream)));
	ctx->last_launches[frame] += 1;
	return HVCF_OK;
}

// frames of a few thousand lights on an unsharded context: the whole frame as one kernel (hvcf_small.cu)
static bool small_frame_ok(const hvcf_ctx* ctx)
{
	return ctx && ctx->hier && ctx->small_mode && !ctx->anim_len && !ctx->shard.on && !ctx->peer.world && ctx->cluster_count == ctx->cluster_total &&
		ctx->num_lights <= small_frame_max_lights() && ctx->d_small_state;
}

static int cull_small(hvcf_ctx* ctx, uint32_t frame, void* stream)
{
	NvtxZone zone("Light_Culling");
	const int rc = hvcf_cull_open(ctx, frame, stream, "hvcf_cull", false, false); // AABBs if dirty; frame / UBO checks
	if (rc != HVCF_OK)
		return rc;
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	SmallParams p{};
	p.ubo = &ctx->d_ubo[frame];
	p.lights = ctx->bound_lights ? ctx->bound_lights : ctx->d_lights;
	p.num_lights = ctx->num_lights;
	p.aabbs = ctx->d_aabb;
	p.cluster_count = ctx->cluster_count;
	p.cap = ctx->cap;
	p.cells = ctx->d_cells;
	p.indices = ctx->d_indices;
	p.frame_total = &ctx->d_global_index[frame];
	p.shard_total = ctx->d_shard_total;
	p.stats = &ctx->d_stats[frame];
	p.scan_state = ctx->d_small_state;
	p.over_part = ctx->d_small_over;
	p.serial = ctx->d_small_ctl;
	p.done = ctx->d_small_ctl + 1;
	p.trace = ctx->d_trace;
	for (int i = 1; i <= 2; ++i)
		HVCF_CUDA(hvcf_kmark(ctx, frame, i, s));
	if (ctx->timing)
		HVCF_CUDA(cudaEventRecord(ctx->ev[frame][1], s));
	HVCF_CUDA(launch_small_frame(p, s));
	if (ctx->timing)
	{
		HVCF_CUDA(cudaEventRecord(ctx->ev[frame][2], s));
		HVCF_CUDA(cudaEventRecord(ctx->ev[frame][3], s));
		ctx->ev_valid[frame] = true;
	}
	HVCF_CUDA(hvcf_kmark(ctx, frame, 3, s));
	HVCF_CUDA(hvcf_kmark(ctx, frame, 4, s));
	if (ctx->ktimes)
	{
		if (!ctx->kev[frame][5])
			HVCF_CUDA(cudaEventCreate(&ctx->kev[frame][5]));
		HVCF_CUDA(cudaEventRecord(ctx->kev[frame][5], s));
	}
	ctx->h_num_rounds = 0;
	ctx->h_num_segs = 1;
	ctx->last_plan_flags = HVCF_PLAN_SINGLE_KERNEL;
	ctx->last_launches[frame] = ctx->open_launches + 1;
	ctx->range_open = false;
	return HVCF_OK;
}

int hvcf_cull(hvcf_ctx* ctx, uint32_t frame, void* stream)
{
	if (small_frame_ok(ctx))
		return cull_small(ctx, frame, stream);
	const int rc = hvcf_cull_begin(ctx, frame, stream);
	if (rc != HVCF_OK)
		return rc;
	return hvcf_cull_finish(ctx, frame, nullptr, stream);
}

} // extern "C"
