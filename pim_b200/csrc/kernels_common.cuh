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

} // namespace pt
