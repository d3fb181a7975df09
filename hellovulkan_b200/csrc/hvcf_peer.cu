// hvcf_peer.cu -- the exchange step of the z-slab sharded grid as kernels over NVLink peer memory.
//
// The reference has no multi-GPU path.  Sharded by z-slab, the only data that crosses GPUs after culling is
//   (1) one uint32 per rank (its index total, to turn shard-local offsets into global ones), and
//   (2) each rank's light cells and its segment of the light-index list, which every other rank needs to hold the
//       full light grid (LightCulling.comp:81-87 output, consumed by Scene.frag:73-78).
// Both are done here by the GPUs themselves: every rank maps its peers' cells / indices / mailbox allocations
// (CUDA IPC across processes, plain pointers inside one process) and
//   peer_totals_kernel   posts its total into every peer's mailbox, waits for the others', writes its index base;
//   peer_push_kernel     streams its cells and index segment into every peer (one coalesced read, world-1 remote
//                        stores over NVLink), then the last CTA raises its "arrived" flag on every peer;
//   peer_wait_kernel     waits until every peer's segment has arrived.
// No host round trip, no size exchange on the host (the segment length is read on the device), no NCCL call.
//
// Ordering.  Data stores to a peer are followed by __threadfence_system() and a release store of the flag; readers
// spin with acquire loads at system scope.  Posting the total of frame e is stream-ordered after everything rank r did
// for frame e-1 (including its consumers on the same stream), and nobody pushes frame e before it has seen every total
// of frame e -- so the totals exchange is also the "your grid buffers may be overwritten" handshake.  Mailbox slots
// are double-buffered on the frame parity.  Every spin has a time limit (kPeerTimeoutNs); on expiry the kernel sets
// the mailbox error word and leaves, so a missing peer is an error code, never a hung GPU.
#include "hvcf_kernels.cuh"

namespace hvcf
{

constexpr unsigned long long kPeerTimeoutNs = 4000000000ull; // 4 s

__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p)
{
	unsigned long long v;
	asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release_sys_u64(unsigned long long* p, unsigned long long v)
{
	asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// A frame whose wait timed out must not scatter cells and indices computed from a wrong index base into the peers' grids:
// the push kernels read their own error word first, skip every data store when it is set and tell the peers instead (their
// error word becomes kind 4), but still raise the "arrived" flags so that nobody spins for 4 s on a frame that is lost anyway.
__device__ __forceinline__ bool peer_frame_failed(const PeerParams& p)
{
	const unsigned int err = *reinterpret_cast<volatile unsigned int*>(&p.mail[p.rank]->error);
	if (err != 0u && blockIdx.x == 0 && threadIdx.x < p.world && threadIdx.x != p.rank)
		atomicCAS_system(&p.mail[threadIdx.x]->error, 0u, (4u << 28) | (p.rank << 20) | (p.epoch & 0xfffffu));
	return err != 0u;
}
__device__ __forceinline__ unsigned long long global_timer_ns()
{
	unsigned long long t;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
	return t;
}

// wait until the word's upper half (the epoch) EQUALS `epoch`; returns the word, or 0 with the error flag set.
// Equality, not >=: nobody can run ahead of a frame we still wait for (a peer needs our total of frame e + 1 before it
// pushes e + 1, and we post that after our own frame e), and after the epoch wraps from 0x7fffffff to 1 a stale word of a
// high epoch must not satisfy the wait.
// error word: kind << 28 | rank waited for << 20 | epoch & 0xfffff  (kind 1 = total, 2 = grid segment, 3 = routed lights,
// 4 = a peer reported an error of its own)
__device__ __forceinline__ unsigned long long spin_epoch(const unsigned long long* word, unsigned long long epoch, PeerMailbox* mail,
	uint32_t kind, uint32_t who)
{
	const unsigned long long t0 = global_timer_ns();
	for (;;)
	{
		const unsigned long long v = ld_acquire_sys_u64(word);
		if ((v >> 32) == epoch)
			return v;
		if (global_timer_ns() - t0 > kPeerTimeoutNs)
		{
			atomicCAS(&mail->error, 0u, (kind << 28) | (who << 20) | (static_cast<uint32_t>(epoch) & 0xfffffu));
			return 0ull;
		}
		__nanosleep(200);
	}
}

// one warp: lane r talks to rank r
// what: 1 = post this rank's total, 2 = wait for everybody's and compute the index base, 3 = both.  Posting early and
// waiting late (hvcf_cull_routed) lets the write pass run while the other ranks are still counting.
__global__ void __launch_bounds__(32) peer_totals_kernel(const PeerParams p, const uint32_t* __restrict__ shard_total,
	uint32_t* __restrict__ index_base, uint32_t* __restrict__ totals_out, uint32_t what)
{
	const uint32_t r = threadIdx.x;
	const uint32_t parity = p.epoch & 1u;
	const uint32_t mine = *shard_total;
	uint32_t total_r = 0;
	if (r < p.world && (what & 1u))
		st_release_sys_u64(&p.mail[r]->totals[parity][p.rank], (static_cast<unsigned long long>(p.epoch) << 32) | mine);
	if (!(what & 2u))
		return;
	if (r < p.world)
	{
		const unsigned long long v = spin_epoch(&p.mail[p.rank]->totals[parity][r], p.epoch, p.mail[p.rank], 1u, r);
		total_r = static_cast<uint32_t>(v);
		if (totals_out)
			totals_out[r] = total_r;
	}
	// index base = totals of the ranks below (slabs are in rank order, so this is the scan in cluster order)
	uint32_t below = r < p.rank ? total_r : 0u;
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		below += __shfl_xor_sync(0xffffffffu, below, o);
	if (r == 0)
		*index_base = below;
}

// grid-stride copy of this rank's cells [cluster_begin, cluster_begin + cluster_count) and indices
// [base, base + total) into the same positions of every peer's buffers.
__global__ void __launch_bounds__(256) peer_push_kernel(const PeerParams p, const uint32_t* __restrict__ shard_total,
	const uint32_t* __restrict__ index_base, uint32_t cluster_begin, uint32_t cluster_count)
{
	const uint32_t base = *index_base, total = *shard_total;
	const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
	const size_t nth = static_cast<size_t>(gridDim.x) * blockDim.x;
	const bool failed = peer_frame_failed(p);

	// ---- cells: 8 B records, copied as 16 B pairs where the range allows ----
	if (!failed)
	{
		const uint2* src = reinterpret_cast<const uint2*>(p.cells[p.rank]);
		const size_t b = cluster_begin, e = static_cast<size_t>(cluster_begin) + cluster_count;
		const size_t b2 = (b + 1) & ~static_cast<size_t>(1), e2 = e & ~static_cast<size_t>(1);
		for (size_t i = b2 / 2 + tid; i < e2 / 2; i += nth)
		{
			const uint4 v = reinterpret_cast<const uint4*>(src)[i];
			for (uint32_t r = 0; r < p.world; ++r)
				if (r != p.rank)
					reinterpret_cast<uint4*>(p.cells[r])[i] = v;
		}
		if (tid == 0 && b < b2 && b < e)
			for (uint32_t r = 0; r < p.world; ++r)
				if (r != p.rank)
					reinterpret_cast<uint2*>(p.cells[r])[b] = src[b];
		if (tid == 1 % nth && e2 < e && e2 >= b2)
			for (uint32_t r = 0; r < p.world; ++r)
				if (r != p.rank)
					reinterpret_cast<uint2*>(p.cells[r])[e2] = src[e2];
	}
	// ---- indices: 4 B words; head up to a 16 B boundary, 16 B body, tail ----
	if (!failed)
	{
		const uint32_t* src = p.indices[p.rank];
		const size_t b = base, e = static_cast<size_t>(base) + total;
		size_t b4 = (b + 3) & ~static_cast<size_t>(3);
		if (b4 > e)
			b4 = e;
		const size_t e4 = b4 + ((e - b4) & ~static_cast<size_t>(3));
		// four independent 16 B loads per thread per trip, then the remote stores: enough bytes in flight per SM to
		// keep the NVLink ports busy (one load per trip measured 150 GB/s at 2 GPUs, profiles/r01e)
		size_t i = b4 / 4 + tid;
		const size_t iend = e4 / 4;
		for (; i + 3 * nth < iend; i += 4 * nth)
		{
			uint4 v[4];
#pragma unroll
			for (int u = 0; u < 4; ++u)
				v[u] = __ldcs(reinterpret_cast<const uint4*>(src) + i + u * nth);
			for (uint32_t r = 0; r < p.world; ++r)
				if (r != p.rank)
				{
#pragma unroll
					for (int u = 0; u < 4; ++u)
						reinterpret_cast<uint4*>(p.indices[r])[i + u * nth] = v[u];
				}
		}
		for (; i < iend; i += nth)
		{
			const uint4 v = reinterpret_cast<const uint4*>(src)[i];
			for (uint32_t r = 0; r < p.world; ++r)
				if (r != p.rank)
					reinterpret_cast<uint4*>(p.indices[r])[i] = v;
		}
		if (tid < 8)
		{
			// lanes 0..3: head words, lanes 4..7: tail words
			const size_t i = tid < 4 ? b + tid : e4 + (tid - 4);
			const bool ok = tid < 4 ? i < b4 : i < e;
			if (ok)
			{
				const uint32_t v = src[i];
				for (uint32_t r = 0; r < p.world; ++r)
					if (r != p.rank)
						p.indices[r][i] = v;
			}
		}
	}
	// ---- last CTA out raises the flag on every peer ----
	// (one system-scope fence per CTA, after the CTA barrier: the barrier orders every thread's stores before thread 0's
	// fence, which is cumulative.  Measured alone at 2 GPUs the push moves 66 MB in 0.11 ms = 590 GB/s whatever the CTA
	// count, profiles/r01k_push_probe_N2.txt.)
	__shared__ bool last;
	__syncthreads();
	if (threadIdx.x == 0)
	{
		__threadfence_system();
		PeerMailbox* mine = p.mail[p.rank];
		const unsigned int done = atomicAdd(&mine->done, 1u);
		last = done == gridDim.x - 1;
		if (last)
			mine->done = 0u;
	}
	__syncthreads();
	if (last && threadIdx.x < p.world)
	{
		__threadfence_system();
		st_release_sys_u64(&p.mail[threadIdx.x]->arrived[p.rank], static_cast<unsigned long long>(p.epoch) << 32);
	}
}

// The same push for a frame whose write pass ran BEFORE the totals were known (hvcf_cull_routed): the rank's cells carry
// shard-local offsets and its index segment sits at seg[0, total); this kernel adds the index base to the offsets and
// places the segment at [base, base + total) on EVERY rank, its own included.  Destination-aligned 16 B stores; the source
// is read with 4 B loads (local memory, coalesced), because base is not a multiple of 4 in general.
__global__ void __launch_bounds__(256) peer_push_relocate_kernel(const PeerParams p, const uint32_t* __restrict__ shard_total,
	const uint32_t* __restrict__ index_base, const uint32_t* __restrict__ seg, uint32_t cluster_begin, uint32_t cluster_count)
{
	const uint32_t base = *index_base, total = *shard_total;
	const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
	const size_t nth = static_cast<size_t>(gridDim.x) * blockDim.x;
	const bool failed = peer_frame_failed(p);
	// ---- cells: {offset + base, count}; own cells are rewritten in place (same element, same thread) ----
	if (!failed)
	{
		const uint2* src = reinterpret_cast<const uint2*>(p.cells[p.rank]);
		for (size_t i = cluster_begin + tid; i < static_cast<size_t>(cluster_begin) + cluster_count; i += nth)
		{
			uint2 v = src[i];
			v.x += base;
			for (uint32_t r = 0; r < p.world; ++r)
				reinterpret_cast<uint2*>(p.cells[r])[i] = v;
		}
	}
	// ---- indices: destination words [base, base + total); head up to a 16 B boundary, 16 B body, tail ----
	if (!failed)
	{
		const size_t b = base, e = static_cast<size_t>(base) + total;
		size_t b4 = (b + 3) & ~static_cast<size_t>(3);
		if (b4 > e)
			b4 = e;
		const size_t e4 = b4 + ((e - b4) & ~static_cast<size_t>(3));
		size_t i = b4 / 4 + tid;
		const size_t iend = e4 / 4;
		for (; i + nth < iend; i += 2 * nth)
		{
			uint4 v[2];
#pragma unroll
			for (int u = 0; u < 2; ++u)
			{
				const uint32_t* s4 = seg + ((i + u * nth) * 4 - b);
				v[u] = make_uint4(__ldcs(s4), __ldcs(s4 + 1), __ldcs(s4 + 2), __ldcs(s4 + 3));
			}
			for (uint32_t r = 0; r < p.world; ++r)
			{
#pragma unroll
				for (int u = 0; u < 2; ++u)
					reinterpret_cast<uint4*>(p.indices[r])[i + u * nth] = v[u];
			}
		}
		for (; i < iend; i += nth)
		{
			const uint32_t* s4 = seg + (i * 4 - b);
			const uint4 v = make_uint4(s4[0], s4[1], s4[2], s4[3]);
			for (uint32_t r = 0; r < p.world; ++r)
				reinterpret_cast<uint4*>(p.indices[r])[i] = v;
		}
		if (tid < 8)
		{
			// lanes 0..3: head words, lanes 4..7: tail words
			const size_t j = tid < 4 ? b + tid : e4 + (tid - 4);
			const bool ok = tid < 4 ? j < b4 : j < e;
			if (ok)
			{
				const uint32_t v = seg[j - b];
				for (uint32_t r = 0; r < p.world; ++r)
					p.indices[r][j] = v;
			}
		}
	}
	__shared__ bool last;
	__syncthreads();
	if (threadIdx.x == 0)
	{
		__threadfence_system();
		PeerMailbox* mine = p.mail[p.rank];
		const unsigned int done = atomicAdd(&mine->done, 1u);
		last = done == gridDim.x - 1;
		if (last)
			mine->done = 0u;
	}
	__syncthreads();
	if (last && threadIdx.x < p.world)
	{
		__threadfence_system();
		st_release_sys_u64(&p.mail[threadIdx.x]->arrived[p.rank], static_cast<unsigned long long>(p.epoch) << 32);
	}
}

// also leaves the total of the WHOLE grid in globalIndexCountBuffers_[frame] (the write pass stored this shard's running
// total there), so that the counter too reads the same on every GPU as after an unsharded frame
__global__ void __launch_bounds__(32) peer_wait_kernel(const PeerParams p, const uint32_t* __restrict__ totals, uint32_t* __restrict__ frame_total)
{
	const uint32_t r = threadIdx.x;
	if (r < p.world && r != p.rank)
		spin_epoch(&p.mail[p.rank]->arrived[r], p.epoch, p.mail[p.rank], 2u, r);
	uint32_t sum = r < p.world ? totals[r] : 0u;
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		sum += __shfl_xor_sync(0xffffffffu, sum, o);
	if (r == 0 && frame_total)
		*frame_total = sum;
	__threadfence_system();
}

cudaError_t launch_peer_totals(const PeerParams& p, const uint32_t* shard_total, uint32_t* index_base, uint32_t* totals_out,
	cudaStream_t stream, uint32_t what)
{
	peer_totals_kernel<<<1, 32, 0, stream>>>(p, shard_total, index_base, totals_out, what);
	return cudaGetLastError();
}

cudaError_t launch_peer_push(const PeerParams& p, const uint32_t* shard_total, const uint32_t* index_base, uint32_t cluster_begin,
	uint32_t cluster_count, uint32_t blocks, cudaStream_t stream)
{
	peer_push_kernel<<<blocks, 256, 0, stream>>>(p, shard_total, index_base, cluster_begin, cluster_count);
	return cudaGetLastError();
}

cudaError_t launch_peer_push_relocate(const PeerParams& p, const uint32_t* shard_total, const uint32_t* index_base, const uint32_t* seg,
	uint32_t cluster_begin, uint32_t cluster_count, uint32_t blocks, cudaStream_t stream)
{
	peer_push_relocate_kernel<<<blocks, 256, 0, stream>>>(p, shard_total, index_base, seg, cluster_begin, cluster_count);
	return cudaGetLastError();
}

cudaError_t launch_peer_wait(const PeerParams& p, const uint32_t* totals, uint32_t* frame_total, cudaStream_t stream)
{
	peer_wait_kernel<<<1, 32, 0, stream>>>(p, totals, frame_total);
	return cudaGetLastError();
}


// every kernel of this file that a frame can launch, force-loaded (HVCF_PRELOAD, hvcf_device.cuh)

cudaError_t preload_peer_kernels()
{
	HVCF_PRELOAD(peer_totals_kernel);
	HVCF_PRELOAD(peer_push_kernel);
	HVCF_PRELOAD(peer_push_relocate_kernel);
	HVCF_PRELOAD(peer_wait_kernel);
	return cudaSuccess;
}

} // namespace hvcf
