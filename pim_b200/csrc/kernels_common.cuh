// kernels_common.cuh — launch geometry shared by every kernel that traverses the BVH.
#pragma once

#include "integrator.h"

namespace pt {

// 128 threads per CTA keeps 16 resident CTAs per SM within the register file at
// <= 128 registers/thread; the shared traversal stack is 16 entries x 8 B x 128
// threads = 16 KB per CTA. Deeper entries spill per thread (TravStack::spill).
constexpr int kTravThreads = 128;
constexpr int kTravSharedDepth = 16;
using GpuStack = TravStack<kTravSharedDepth>;

#define PT_DECLARE_STACK(st)                                                        \
    __shared__ unsigned long long pt_stackMem[::pt::kTravSharedDepth * ::pt::kTravThreads]; \
    ::pt::GpuStack st;                                                              \
    st.init(pt_stackMem + threadIdx.x, ::pt::kTravThreads)

// Shading kernels only traverse on their cold paths (glass thickness, in-medium light sample).
// They pass this tag instead of a stack; the cold function that needs one builds a per-thread
// stack on the spot, so the hot path carries neither shared memory nor a stack pointer.
struct LazyStack {};
struct ColdStack : TravStack<24>
{
    unsigned long long mem[24];
    __device__ __forceinline__ ColdStack() { init(mem, 1); }
};
__device__ __forceinline__ RayHit intersectLocalRare(const DScene& S, V3 ro, V3 rd, float tNear, float tFar, LazyStack&)
{
    ColdStack st;
    return intersectLocal(S, ro, rd, tNear, tFar, st);
}

// Block-wide allocation of consecutive positions: per-warp ballots, one global atomic per CTA.
// Every thread of the block must call it. (A per-warp atomic whose result the warp waits for
// was 9-11 % of the shading kernels' stall samples.)
template <int kWarps>
__device__ __forceinline__ int blockAlloc(bool pred, int32_t* counter, int& blockTotal)
{
    __shared__ int sCnt[kWarps];
    __shared__ int sBase;
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) sCnt[warp] = __popc(mask);
    __syncthreads();
    int off = 0, total = 0;
    #pragma unroll
    for (int w = 0; w < kWarps; ++w)
    {
        const int c = sCnt[w];
        off += (w < warp) ? c : 0;
        total += c;
    }
    if (threadIdx.x == 0) sBase = total ? atomicAdd(counter, total) : 0;
    __syncthreads();
    blockTotal = total;
    return sBase + off + __popc(mask & ((1u << lane) - 1u));
}

} // namespace pt
