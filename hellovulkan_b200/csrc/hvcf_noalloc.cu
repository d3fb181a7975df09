// hvcf_noalloc.cu -- hvcf_cull_small_noalloc (include/hvcf.h): the single-kernel culling frame of hvcf_small.cu for callers that
// cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The kernel (small_frame_kernel: LightCulling.comp:49-88 for the whole grid in one launch) needs scratch memory: the look-back
// state words of its offset scan, a per-CTA overflow count, a ticket counter, its frame serial and a block of work counters.
// hvcf_cull keeps those in cudaMalloc'ed buffers of the context.  Here they live in ONE file-scope __device__ array of a fixed,
// modest size, which every GPU that loads the library holds until the process exits:
//
//     kNoAllocMaxClusters = 131,072 clusters (4,096 CTAs; the reference's grid is 3,456), at most 2,048 lights
//     -> 512 B control + 32 KiB state words + 16 KiB overflow counts = 49,664 B
//
// The entry point takes inputs and outputs as device pointers, checks them, clears the part of the scratch that must start at
// zero (cudaMemsetAsync on the caller's stream) and launches the kernel: no cudaMalloc, no cudaFree, no allocating library call.
// In host code the array's NAME is a host-side stand-in, not a device address; the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current (scratch_for_current_device below).
//
// Every call shares the array.  Calls that overlap on the GPU (two streams, two host threads, two contexts of one device) RACE on
// it: serialise them.  Calls on different devices are independent (each device has its own copy of the array).
#include "hvcf_ctx.cuh"

#include <mutex>

using namespace hvcf;

namespace
{
constexpr uint32_t kNoAllocMaxClusters = 131072;
constexpr uint32_t kNoAllocMaxCtas = kNoAllocMaxClusters / 32;
constexpr size_t kCtlBytes = 512; // serial, ticket counter, shard total, DeviceStats
constexpr size_t kStateOffset = kCtlBytes;
constexpr size_t kScratchBytes = kStateOffset + (sizeof(unsigned long long) + sizeof(uint32_t)) * kNoAllocMaxCtas; // overflow counts follow the state words in use
constexpr size_t kStatsOffset = 256;
static_assert(kStatsOffset + sizeof(DeviceStats) <= kCtlBytes, "control block too small");

} // namespace

// file scope: one copy per device that loads the library, zero when the module is loaded, never freed
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

namespace
{
constexpr int kMaxDevices = 64;
std::mutex g_mutex;
unsigned char* g_scratch_ptr[kMaxDevices]; // device address of g_noalloc_scratch on device i, once asked for

// the device address of the scratch array on the CURRENT device (asked once per device)
int scratch_for_current_device(unsigned char** out)
{
	int dev = -1;
	HVCF_CUDA(cudaGetDevice(&dev));
	if (dev < 0 || dev >= kMaxDevices)
		return fail(HVCF_ERR_UNSUPPORTED, "hvcf_cull_small_noalloc: device ordinal %d (at most %d devices)", dev, kMaxDevices);
	std::lock_guard<std::mutex> lock(g_mutex);
	if (!g_scratch_ptr[dev])
	{
		unsigned char* ptr = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch);
		if (e != cudaSuccess || !ptr)
			return fail(HVCF_ERR_CUDA, "hvcf_cull_small_noalloc: cudaGetSymbolAddress of the scratch array on device %d: %s", dev,
				cudaGetErrorString(e));
		g_scratch_ptr[dev] = ptr;
	}
	*out = g_scratch_ptr[dev];
	return HVCF_OK;
}

bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }
} // namespace

extern "C" {

int hvcf_cull_small_noalloc(const hvcf_cluster_forward_ubo* d_ubo, const hvcf_light_data* d_lights, uint32_t num_lights,
	const hvcf_aabb* d_aabbs, uint32_t cluster_count, uint32_t max_lights_per_cluster, hvcf_light_cell* d_light_cells,
	uint32_t* d_light_indices, uint64_t light_indices_capacity, uint32_t* d_total, void* stream)
{
	if (!d_ubo || !d_aabbs || !d_light_cells || !d_light_indices || !d_total || (!d_lights && num_lights))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_small_noalloc: null argument");
	if (cluster_count == 0 || max_lights_per_cluster == 0)
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_small_noalloc: cluster_count and max_lights_per_cluster must be positive");
	if (cluster_count > kNoAllocMaxClusters)
		return fail(HVCF_ERR_CAPACITY, "hvcf_cull_small_noalloc: %u clusters, the static scratch holds at most %u (use hvcf_cull)", cluster_count,
			kNoAllocMaxClusters);
	if (num_lights > small_frame_max_lights())
		return fail(HVCF_ERR_CAPACITY, "hvcf_cull_small_noalloc: %u lights, the single-kernel frame takes at most %u (use hvcf_cull)", num_lights,
			small_frame_max_lights());
	const uint64_t worst = static_cast<uint64_t>(cluster_count) * (num_lights < max_lights_per_cluster ? num_lights : max_lights_per_cluster);
	if (light_indices_capacity < worst)
		return fail(HVCF_ERR_CAPACITY, "hvcf_cull_small_noalloc: light_indices_capacity %llu < clusters x min(lights, cap) = %llu",
			static_cast<unsigned long long>(light_indices_capacity), static_cast<unsigned long long>(worst));
	if (!aligned(d_ubo, 16) || !aligned(d_lights, 16) || !aligned(d_aabbs, 16) || !aligned(d_light_cells, 8) || !aligned(d_light_indices, 4) ||
		!aligned(d_total, 4))
		return fail(HVCF_ERR_INVALID_ARGUMENT, "hvcf_cull_small_noalloc: misaligned pointer (UBO, lights, AABBs: 16 B; light cells: 8 B)");

	unsigned char* scratch = nullptr;
	const int rc = scratch_for_current_device(&scratch);
	if (rc != HVCF_OK)
		return rc;
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	const uint32_t ctas = (cluster_count + 31) / 32;
	uint32_t* ctl = reinterpret_cast<uint32_t*>(scratch);

	// The array keeps the last call's contents.  What must start at zero: the look-back state words and overflow counts of the CTAs
	// of THIS call (an earlier call over a larger grid left words behind that a later serial could match after it wraps) and the
	// ticket counter (left non-zero only by a launch that died).  The serial (ctl[0]) is the kernel's own and is left alone.
	// One memset: the control words behind the serial, the state words, and the overflow counts packed right behind them.
	HVCF_CUDA(cudaMemsetAsync(ctl + 1, 0, kCtlBytes - sizeof(uint32_t) + (sizeof(unsigned long long) + sizeof(uint32_t)) * ctas, s));
	const size_t over_offset = kStateOffset + sizeof(unsigned long long) * ctas;

	SmallParams p{};
	p.ubo = d_ubo;
	p.lights = d_lights;
	p.num_lights = num_lights;
	p.aabbs = d_aabbs;
	p.cluster_count = cluster_count;
	p.cap = max_lights_per_cluster;
	p.cells = d_light_cells;
	p.indices = d_light_indices;
	p.frame_total = d_total;
	p.serial = ctl;
	p.done = ctl + 1;
	p.shard_total = ctl + 2;
	p.stats = reinterpret_cast<DeviceStats*>(scratch + kStatsOffset);
	p.scan_state = reinterpret_cast<unsigned long long*>(scratch + kStateOffset);
	p.over_part = reinterpret_cast<uint32_t*>(scratch + over_offset);
	p.trace = nullptr;
	HVCF_CUDA(launch_small_frame(p, s));
	return HVCF_OK;
}

} // extern "C"
