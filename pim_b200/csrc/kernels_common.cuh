// kernels_common.cuh — launch geometry shared by every kernel that traverses the BVH.
#pragma once

#include "traverse.h"

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

// Shading kernels only traverse on their cold paths (glass thickness, in-medium light sample):
// their stack is a plain per-thread array, which leaves shared memory and registers to the
// shading arithmetic.
struct ColdStack : TravStack<24>
{
    unsigned long long mem[24];
    __device__ __forceinline__ ColdStack() { init(mem, 1); }
};

} // namespace pt
