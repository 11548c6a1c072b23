// ba_peer.cuh — one-shot all-reduce of the small per-iteration vectors over NVLink peer memory.
//
// The sharded PCG needs two sums over ranks per iteration (the boundary part of Es^T p, the restricted residual
// of the coarse level; a few KB to a few hundred KB each).  An ncclAllReduce of that size costs 16 us at 2 ranks and
// 35-48 us at 8 — more than the streamed product it sits next to.  Here every rank owns one exchange buffer
// (cudaMalloc, exported with cudaIpcGetMemHandle; the handles travel once through ncclAllGather) that all peers map;
// one kernel per rank then
//   1. pushes its vector into slot [parity][rank] of EVERY rank's buffer, every double as two 8-byte words
//      {low half, sequence number} {high half, sequence number} — an 8-byte store arrives whole, so the data carry
//      their own "ready" flag and no fence or separate flag round trip is needed (the LL idea of NCCL), and
//   2. sums the world slots of its own buffer in rank order, every thread polling the words it needs until they
//      carry the current sequence number — every rank gets the bitwise identical result.
// Threads only wait for data that the same thread index range of the peers pushes BEFORE it starts waiting, so
// there is no dependency between the blocks of one grid.  Two slot parities are enough: a rank can be at most one
// collective ahead of the slowest rank, because it needs that rank's data of the current one to move on, and a
// stale word of the same parity carries sequence number - 2.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vmm {

constexpr int kPeerMaxWorld = 8;
constexpr int kPeerMaxBlocks = 32;
constexpr int kPeerThreads = 256;
constexpr size_t kPeerCapDoubles = 64 * 1024;   // largest vector that goes this way (512 KB); larger ones use NCCL

struct PeerView {
    uint4* buf[kPeerMaxWorld];                  // every rank's exchange buffer as mapped into this process
    int rank, world;
};

// layout of one exchange buffer: [2 parities][world slots][kPeerCapDoubles] 16-byte words
__host__ __device__ inline size_t peer_slot_offset(int world, int parity, int r) { return ((size_t)parity * world + r) * kPeerCapDoubles; }
inline size_t peer_buffer_bytes(int world) { return (size_t)2 * world * kPeerCapDoubles * sizeof(uint4); }

__device__ __forceinline__ void st_word_pair(uint4* p, double v, uint32_t seq) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    const uint32_t lo = (uint32_t)b, hi = (uint32_t)(b >> 32);
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(lo), "r"(seq) : "memory");
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(reinterpret_cast<char*>(p) + 8), "r"(hi), "r"(seq) : "memory");
}
__device__ __forceinline__ double ld_word_pair(const uint4* p, uint32_t seq) {
    uint32_t lo, f1, hi, f2;
    do {
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(lo), "=r"(f1), "=r"(hi), "=r"(f2) : "l"(p) : "memory");
    } while (f1 != seq || f2 != seq);
    return __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
}

// data[0..n) <- sum over ranks of data[0..n); seq: this collective's sequence number (1, 2, 3, ... the same on all ranks)
__global__ void __launch_bounds__(kPeerThreads)
k_peer_allreduce(PeerView pv, int n, double* __restrict__ data, uint32_t seq) {
    const int parity = (int)(seq & 1);
    const int stride = gridDim.x * kPeerThreads;
    const int t0 = blockIdx.x * kPeerThreads + threadIdx.x;
    for (int i = t0; i < n; i += stride) {
        const double v = data[i];
        for (int r = 0; r < pv.world; ++r) st_word_pair(pv.buf[r] + peer_slot_offset(pv.world, parity, pv.rank) + i, v, seq);
    }
    const uint4* mine = pv.buf[pv.rank];
    for (int i = t0; i < n; i += stride) {
        double s = 0.0;
        for (int r = 0; r < pv.world; ++r) s += ld_word_pair(mine + peer_slot_offset(pv.world, parity, r) + i, seq);
        data[i] = s;
    }
}

}  // namespace vmm
