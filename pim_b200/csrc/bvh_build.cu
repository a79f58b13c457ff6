// bvh_build.cu — GPU build of the 8-wide compressed BVH (replaces Embree's
// rtcCommitScene at RTC_BUILD_QUALITY_HIGH, src/rendering/path_tracer.c:618-690).
//
//   K1a triBounds   per-triangle box + centroid bounds (block reduce, ordered-int atomics)
//   K1b morton      63-bit Morton key of the centroid, value = triangle index
//   K1c radix sort  LSD, 8 x 8-bit digits: histogram / scan / stable scatter with
//                   warp match-any ranking
//   K2a radixTree   Karras' binary radix tree over the sorted keys
//   K2b refit       bottom-up boxes, one atomic visit counter per internal node
//   K2c collapse    breadth-first greedy collapse to 8-wide nodes, child boxes
//                   quantised to 8 bits; triangles copied into leaf order
#include "pt_device.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace pt {

namespace {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;
constexpr int kSortTile = kSortThreads * kSortItems;   // keys per block

// up to this many triangles the tree is built top-down with binned SAH (the reference asks Embree for
// RTC_BUILD_QUALITY_HIGH, path_tracer.c:627); above it with the Morton-code radix tree (LBVH), whose
// build is one pass per level of the sort, not of the tree. With the many-CTA levels for large nodes
// (below) the SAH build of the 10 M-triangle soup is fast enough to be the default there too.
#ifndef PT_SAH_MAX_TRIS
#define PT_SAH_MAX_TRIS (64 << 20)
#endif
constexpr int kSahMaxTris = PT_SAH_MAX_TRIS;

// PIMPT_SAH_MAX_TRIS overrides the threshold at run time (the tests use it to reach the LBVH path
// with a scene that is quick to check)
int sahMaxTris()
{
    const char* e = getenv("PIMPT_SAH_MAX_TRIS");
    return e ? atoi(e) : kSahMaxTris;
}

// PIMPT_SAH_LARGE_NODE overrides kLargeNode: nodes above it take the many-CTAs-per-node levels
// (a huge value turns them off, which is how the tests compare the two paths)
#ifndef PT_SAH_LARGE_NODE
#define PT_SAH_LARGE_NODE 8192
#endif
int largeNodeMin()
{
    const char* e = getenv("PIMPT_SAH_LARGE_NODE");
    const int v = e ? atoi(e) : PT_SAH_LARGE_NODE;
    return v < 2 * 2048 ? 2 * 2048 : v;       // at least two tiles
}

// PIMPT_FLOAT_NODES overrides kFloatNodeMaxNodes (0 = never build the fp32 node form)
int floatNodeMaxNodes()
{
    const char* e = getenv("PIMPT_FLOAT_NODES");
    return e ? atoi(e) : kFloatNodeMaxNodes;
}

__device__ __forceinline__ int floatToOrdered(float f)
{
    int i = __float_as_int(f);
    return (i >= 0) ? i : (i ^ 0x7fffffff);
}
__device__ __forceinline__ float orderedToFloat(int i)
{
    return __int_as_float((i >= 0) ? i : (i ^ 0x7fffffff));
}

// bounds[0..2] = centroid min, [3..5] = centroid max, [6..8] = vertex min, [9..11] = vertex max (ordered ints)
__global__ void k_initBounds(int* bounds)
{
    int i = threadIdx.x;
    if (i < 12) bounds[i] = ((i / 3) & 1) ? floatToOrdered(-3.0e38f) : floatToOrdered(3.0e38f);
}

__global__ void __launch_bounds__(256) k_triBounds(const F4* __restrict__ pos, int n, F4* __restrict__ leafLo,
                                                   F4* __restrict__ leafHi, int* bounds)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, hi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
    float clo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, chi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
    if (t < n)
    {
        U4 a = ldg128(pos + 3 * (size_t)t), b = ldg128(pos + 3 * (size_t)t + 1), c = ldg128(pos + 3 * (size_t)t + 2);
        float ax[3] = { u2f(a.x), u2f(a.y), u2f(a.z) };
        float bx[3] = { u2f(b.x), u2f(b.y), u2f(b.z) };
        float cx[3] = { u2f(c.x), u2f(c.y), u2f(c.z) };
        for (int k = 0; k < 3; ++k)
        {
            lo[k] = f1_min(ax[k], f1_min(bx[k], cx[k]));
            hi[k] = f1_max(ax[k], f1_max(bx[k], cx[k]));
            clo[k] = chi[k] = (lo[k] + hi[k]) * 0.5f;
        }
        leafLo[t] = f4(lo[0], lo[1], lo[2], 0.0f);
        leafHi[t] = f4(hi[0], hi[1], hi[2], 0.0f);
    }
    for (int k = 0; k < 3; ++k)
    {
        for (int o = 16; o > 0; o >>= 1)
        {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
            clo[k] = fminf(clo[k], __shfl_xor_sync(0xffffffffu, clo[k], o));
            chi[k] = fmaxf(chi[k], __shfl_xor_sync(0xffffffffu, chi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0)
    {
        for (int k = 0; k < 3; ++k)
        {
            atomicMin(&bounds[k], floatToOrdered(clo[k]));
            atomicMax(&bounds[3 + k], floatToOrdered(chi[k]));
            atomicMin(&bounds[6 + k], floatToOrdered(lo[k]));
            atomicMax(&bounds[9 + k], floatToOrdered(hi[k]));
        }
    }
}

__global__ void __launch_bounds__(256) k_morton(const F4* __restrict__ leafLo, const F4* __restrict__ leafHi, int n,
                                                const int* __restrict__ bounds, uint64_t* __restrict__ keys,
                                                uint32_t* __restrict__ vals)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    V3 lo = v3(orderedToFloat(bounds[0]), orderedToFloat(bounds[1]), orderedToFloat(bounds[2]));
    V3 hi = v3(orderedToFloat(bounds[3]), orderedToFloat(bounds[4]), orderedToFloat(bounds[5]));
    V3 ext = hi - lo;
    V3 invExt = v3(ext.x > 0.0f ? 1.0f / ext.x : 0.0f, ext.y > 0.0f ? 1.0f / ext.y : 0.0f,
                   ext.z > 0.0f ? 1.0f / ext.z : 0.0f);
    F4 a = leafLo[t], b = leafHi[t];
    V3 c = v3((a.x + b.x) * 0.5f, (a.y + b.y) * 0.5f, (a.z + b.z) * 0.5f);
    keys[t] = morton63(c, lo, invExt);
    vals[t] = (uint32_t)t;
}

// ---- radix sort -------------------------------------------------------------

// hist[digit * numBlocks + block]
__global__ void __launch_bounds__(kSortThreads) k_sortHist(const uint64_t* __restrict__ keys, int n, int shift,
                                                           uint32_t* __restrict__ hist, int numBlocks)
{
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    int base = blockIdx.x * kSortTile;
    for (int r = 0; r < kSortItems; ++r)
    {
        int i = base + r * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&sh[(uint32_t)(keys[i] >> shift) & 0xffu], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * numBlocks + blockIdx.x] = sh[threadIdx.x];
}

// exclusive scan of `count` entries by one block of 1024 threads
__global__ void __launch_bounds__(1024) k_sortScan(uint32_t* __restrict__ hist, int count)
{
    __shared__ uint32_t partial[1024];
    int per = (count + 1023) / 1024;
    int b = threadIdx.x * per, e = min(b + per, count);
    uint32_t sum = 0;
    for (int i = b; i < e; ++i) sum += hist[i];
    partial[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1)
    {
        uint32_t v = (threadIdx.x >= o) ? partial[threadIdx.x - o] : 0;
        __syncthreads();
        partial[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = partial[threadIdx.x] - sum;
    for (int i = b; i < e; ++i)
    {
        uint32_t v = hist[i];
        hist[i] = run;
        run += v;
    }
}

// stable scatter: every warp ranks its 256 consecutive keys with match-any
__global__ void __launch_bounds__(kSortThreads) k_sortScatter(const uint64_t* __restrict__ keysIn,
                                                              const uint32_t* __restrict__ valsIn, int n, int shift,
                                                              const uint32_t* __restrict__ hist, int numBlocks,
                                                              uint64_t* __restrict__ keysOut,
                                                              uint32_t* __restrict__ valsOut)
{
    constexpr int kWarps = kSortThreads / 32;
    __shared__ uint32_t warpCnt[kWarps][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < kWarps * 256; i += kSortThreads) (&warpCnt[0][0])[i] = 0;
    __syncthreads();

    const int warpBase = blockIdx.x * kSortTile + warp * (kSortItems * 32);
    uint64_t key[kSortItems];
    uint32_t val[kSortItems];
    uint32_t rank[kSortItems];
    const uint32_t ltMask = (1u << lane) - 1u;
    for (int r = 0; r < kSortItems; ++r)
    {
        int i = warpBase + r * 32 + lane;
        bool valid = i < n;
        key[r] = valid ? keysIn[i] : ~0ull;
        val[r] = valid ? valsIn[i] : 0u;
        uint32_t digit = valid ? ((uint32_t)(key[r] >> shift) & 0xffu) : (0x100u + (uint32_t)lane);
        uint32_t peers = __match_any_sync(0xffffffffu, digit);
        int leader = __ffs((int)peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader)
        {
            old = warpCnt[warp][digit];
            warpCnt[warp][digit] = old + (uint32_t)__popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[r] = old + (uint32_t)__popc(peers & ltMask);
        __syncwarp();
    }
    __syncthreads();
    // per digit: exclusive prefix over warps, plus this block's global base
    {
        uint32_t d = threadIdx.x;
        uint32_t run = hist[(size_t)d * numBlocks + blockIdx.x];
        for (int w = 0; w < kWarps; ++w)
        {
            uint32_t c = warpCnt[w][d];
            warpCnt[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
    for (int r = 0; r < kSortItems; ++r)
    {
        int i = warpBase + r * 32 + lane;
        if (i < n)
        {
            uint32_t digit = (uint32_t)(key[r] >> shift) & 0xffu;
            uint32_t dst = warpCnt[warp][digit] + rank[r];
            keysOut[dst] = key[r];
            valsOut[dst] = val[r];
        }
    }
}

// ---- hierarchy --------------------------------------------------------------

__global__ void __launch_bounds__(256) k_radixTree(const uint64_t* __restrict__ keys, int n, F4* lo, F4* hi,
                                                   int* first, int* last, int* parent)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int f, l, s;
    lbvhRange(keys, n, i, f, l, s);
    int left = (s == f) ? (n - 1 + s) : s;
    int right = (s + 1 == l) ? (n - 1 + s + 1) : (s + 1);
    first[i] = f; last[i] = l;
    lo[i].w = u2f((uint32_t)left);
    hi[i].w = u2f((uint32_t)right);
    parent[left] = i;
    parent[right] = i;
}

__global__ void __launch_bounds__(256) k_leafInit(const uint32_t* __restrict__ vals, const F4* __restrict__ leafLo,
                                                  const F4* __restrict__ leafHi, int n, F4* lo, F4* hi, int* first,
                                                  int* last)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint32_t t = vals[k];
    int id = n - 1 + k;
    F4 a = leafLo[t], b = leafHi[t];
    lo[id] = f4(a.x, a.y, a.z, 0.0f);
    hi[id] = f4(b.x, b.y, b.z, 0.0f);
    first[id] = k; last[id] = k;
}

__global__ void __launch_bounds__(256) k_refit(int n, F4* lo, F4* hi, const int* __restrict__ parent, int* visits)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int node = parent[n - 1 + k];
    while (node >= 0)
    {
        __threadfence();
        if (atomicAdd(&visits[node], 1) == 0) return;   // first arrival waits for the sibling subtree
        // second arrival: both children are final
        volatile F4* vlo = lo; volatile F4* vhi = hi;
        int l = (int)f2u(vlo[node].w), r = (int)f2u(vhi[node].w);
        float lx = fminf(vlo[l].x, vlo[r].x), ly = fminf(vlo[l].y, vlo[r].y), lz = fminf(vlo[l].z, vlo[r].z);
        float hx = fmaxf(vhi[l].x, vhi[r].x), hy = fmaxf(vhi[l].y, vhi[r].y), hz = fmaxf(vhi[l].z, vhi[r].z);
        vlo[node].x = lx; vlo[node].y = ly; vlo[node].z = lz;
        vhi[node].x = hx; vhi[node].y = hy; vhi[node].z = hz;
        node = parent[node];
    }
}

// ---- binned SAH (top-down, one warp per node, one launch per tree level) --------------------
// Node 0 is the root; a node with more than kLeafMaxTris primitives is split along the best of
// 3 x 15 candidate planes (16 centroid bins per axis, cost = area(L)*n(L) + area(R)*n(R)); its
// primitives are partitioned in place in order[] and its two children get consecutive ids.
// Children boxes come from the bin boxes, so no extra pass is needed.

constexpr int kSahBins = 16;
constexpr int kSahWarps = 8;

struct SahBins
{
    int lo[3][kSahBins][3];
    int hi[3][kSahBins][3];
    int cnt[3][kSahBins];
};

__device__ __forceinline__ float warpMinF(float v)
{
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warpMaxF(float v)
{
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// counters: [0] nodes allocated, [1] next-level queue length
__global__ void __launch_bounds__(kSahWarps * 32) k_sahSplit(BinTree T, const F4* __restrict__ leafLo,
                                                              const F4* __restrict__ leafHi, int* order, int* tmp,
                                                              const int* __restrict__ qIn, int nIn, int* qOut,
                                                              uint32_t* counters)
{
    __shared__ SahBins sbins[kSahWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int w = blockIdx.x * kSahWarps + warp;
    if (w >= nIn) return;
    SahBins& B = sbins[warp];
    const int node = qIn[w];
    const int f = T.first[node], l = T.last[node], cnt = l - f + 1;

    // centroid bounds of the range
    float cmin[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, cmax[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
    for (int i = f + lane; i <= l; i += 32)
    {
        const int t = order[i];
        const F4 a = leafLo[t], b = leafHi[t];
        const float c[3] = { (a.x + b.x) * 0.5f, (a.y + b.y) * 0.5f, (a.z + b.z) * 0.5f };
        for (int k = 0; k < 3; ++k) { cmin[k] = fminf(cmin[k], c[k]); cmax[k] = fmaxf(cmax[k], c[k]); }
    }
    float scale[3];
    for (int k = 0; k < 3; ++k)
    {
        cmin[k] = warpMinF(cmin[k]); cmax[k] = warpMaxF(cmax[k]);
        const float ext = cmax[k] - cmin[k];
        scale[k] = ext > 0.0f ? (float)kSahBins / ext : 0.0f;
    }
    // bins
    for (int i = lane; i < 3 * kSahBins; i += 32)
    {
        const int a = i / kSahBins, b = i % kSahBins;
        for (int k = 0; k < 3; ++k) { B.lo[a][b][k] = floatToOrdered(3.0e38f); B.hi[a][b][k] = floatToOrdered(-3.0e38f); }
        B.cnt[a][b] = 0;
    }
    __syncwarp();
    for (int i = f + lane; i <= l; i += 32)
    {
        const int t = order[i];
        const F4 a = leafLo[t], b = leafHi[t];
        const float c[3] = { (a.x + b.x) * 0.5f, (a.y + b.y) * 0.5f, (a.z + b.z) * 0.5f };
        const int lo3[3] = { floatToOrdered(a.x), floatToOrdered(a.y), floatToOrdered(a.z) };
        const int hi3[3] = { floatToOrdered(b.x), floatToOrdered(b.y), floatToOrdered(b.z) };
        for (int ax = 0; ax < 3; ++ax)
        {
            if (scale[ax] == 0.0f) continue;
            int bin = (int)((c[ax] - cmin[ax]) * scale[ax]);
            bin = bin < 0 ? 0 : (bin > kSahBins - 1 ? kSahBins - 1 : bin);
            for (int k = 0; k < 3; ++k) { atomicMin(&B.lo[ax][bin][k], lo3[k]); atomicMax(&B.hi[ax][bin][k], hi3[k]); }
            atomicAdd(&B.cnt[ax][bin], 1);
        }
    }
    __syncwarp();
    // 45 candidate planes, two per lane
    float bestCost = 3.0e38f;
    int bestCand = -1;
    for (int cand = lane; cand < 3 * (kSahBins - 1); cand += 32)
    {
        const int ax = cand / (kSahBins - 1), sp = cand % (kSahBins - 1);
        if (scale[ax] == 0.0f) continue;
        float llo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, lhi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
        float rlo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, rhi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
        int nl = 0, nr = 0;
        for (int b = 0; b < kSahBins; ++b)
        {
            const int c = B.cnt[ax][b];
            if (c == 0) continue;
            if (b <= sp)
            {
                nl += c;
                for (int k = 0; k < 3; ++k) { llo[k] = fminf(llo[k], orderedToFloat(B.lo[ax][b][k])); lhi[k] = fmaxf(lhi[k], orderedToFloat(B.hi[ax][b][k])); }
            }
            else
            {
                nr += c;
                for (int k = 0; k < 3; ++k) { rlo[k] = fminf(rlo[k], orderedToFloat(B.lo[ax][b][k])); rhi[k] = fmaxf(rhi[k], orderedToFloat(B.hi[ax][b][k])); }
            }
        }
        if (nl == 0 || nr == 0) continue;
        const float al = boxArea(f4(llo[0], llo[1], llo[2], 0), f4(lhi[0], lhi[1], lhi[2], 0));
        const float ar = boxArea(f4(rlo[0], rlo[1], rlo[2], 0), f4(rhi[0], rhi[1], rhi[2], 0));
        const float cost = al * (float)nl + ar * (float)nr;
        if (cost < bestCost) { bestCost = cost; bestCand = cand; }
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        const float oc = __shfl_xor_sync(0xffffffffu, bestCost, o);
        const int ocand = __shfl_xor_sync(0xffffffffu, bestCand, o);
        if (oc < bestCost || (oc == bestCost && ocand >= 0 && (bestCand < 0 || ocand < bestCand))) { bestCost = oc; bestCand = ocand; }
    }
    int nLeft;
    float clo[2][3], chi[2][3];
    if (bestCand >= 0)
    {
        const int ax = bestCand / (kSahBins - 1), sp = bestCand % (kSahBins - 1);
        for (int side = 0; side < 2; ++side)
            for (int k = 0; k < 3; ++k) { clo[side][k] = 3.0e38f; chi[side][k] = -3.0e38f; }
        nLeft = 0;
        for (int b = 0; b < kSahBins; ++b)
        {
            if (B.cnt[ax][b] == 0) continue;
            const int side = b <= sp ? 0 : 1;
            if (side == 0) nLeft += B.cnt[ax][b];
            for (int k = 0; k < 3; ++k)
            {
                clo[side][k] = fminf(clo[side][k], orderedToFloat(B.lo[ax][b][k]));
                chi[side][k] = fmaxf(chi[side][k], orderedToFloat(B.hi[ax][b][k]));
            }
        }
        // stable partition of order[f..l] by bin <= sp, through tmp
        int offL = 0, offR = 0;
        for (int base = f; base <= l; base += 32)
        {
            const int i = base + lane;
            const bool valid = i <= l;
            int t = 0; bool left = false;
            if (valid)
            {
                t = order[i];
                const F4 a = leafLo[t], b = leafHi[t];
                const float c = ax == 0 ? (a.x + b.x) * 0.5f : (ax == 1 ? (a.y + b.y) * 0.5f : (a.z + b.z) * 0.5f);
                int bin = (int)((c - cmin[ax]) * scale[ax]);
                bin = bin < 0 ? 0 : (bin > kSahBins - 1 ? kSahBins - 1 : bin);
                left = bin <= sp;
            }
            const unsigned ml = __ballot_sync(0xffffffffu, valid && left);
            const unsigned mr = __ballot_sync(0xffffffffu, valid && !left);
            const unsigned lt = (1u << lane) - 1u;
            if (valid) tmp[left ? (f + offL + __popc(ml & lt)) : (f + nLeft + offR + __popc(mr & lt))] = t;
            offL += __popc(ml); offR += __popc(mr);
        }
        __syncwarp();
        for (int i = f + lane; i <= l; i += 32) order[i] = tmp[i];
        __syncwarp();
    }
    else
    {
        // all centroids in one bin on every axis: split the range in the middle
        nLeft = cnt / 2;
        for (int side = 0; side < 2; ++side)
        {
            float lo3[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, hi3[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
            const int b0 = side == 0 ? f : f + nLeft, b1 = side == 0 ? f + nLeft - 1 : l;
            for (int i = b0 + lane; i <= b1; i += 32)
            {
                const int t = order[i];
                const F4 a = leafLo[t], b = leafHi[t];
                lo3[0] = fminf(lo3[0], a.x); lo3[1] = fminf(lo3[1], a.y); lo3[2] = fminf(lo3[2], a.z);
                hi3[0] = fmaxf(hi3[0], b.x); hi3[1] = fmaxf(hi3[1], b.y); hi3[2] = fmaxf(hi3[2], b.z);
            }
            for (int k = 0; k < 3; ++k) { clo[side][k] = warpMinF(lo3[k]); chi[side][k] = warpMaxF(hi3[k]); }
        }
    }
    if (lane == 0)
    {
        const uint32_t base = atomicAdd(&counters[0], 2u);
        const int kids[2] = { (int)base, (int)base + 1 };
        T.lo[node].w = u2f((uint32_t)kids[0]);
        T.hi[node].w = u2f((uint32_t)kids[1]);
        const int kf[2] = { f, f + nLeft }, kl[2] = { f + nLeft - 1, l };
        for (int side = 0; side < 2; ++side)
        {
            const int k = kids[side];
            T.lo[k] = f4(clo[side][0], clo[side][1], clo[side][2], 0.0f);
            T.hi[k] = f4(chi[side][0], chi[side][1], chi[side][2], 0.0f);
            T.first[k] = kf[side]; T.last[k] = kl[side];
            if (kl[side] - kf[side] + 1 > kLeafMaxTris) qOut[atomicAdd(&counters[1], 1u)] = k;
        }
    }
}

__global__ void k_sahInit(BinTree T, int n, const int* __restrict__ bounds, int* order, int* qOut, uint32_t* counters)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) order[i] = i;
    if (i == 0)
    {
        T.lo[0] = f4(orderedToFloat(bounds[6]), orderedToFloat(bounds[7]), orderedToFloat(bounds[8]), 0.0f);
        T.hi[0] = f4(orderedToFloat(bounds[9]), orderedToFloat(bounds[10]), orderedToFloat(bounds[11]), 0.0f);
        T.first[0] = 0; T.last[0] = n - 1;
        counters[0] = 1u;
        counters[1] = 0u;
        if (n > kLeafMaxTris) { qOut[0] = 0; counters[1] = 1u; }
    }
}

// ---- binned SAH, large nodes: many CTAs per node -----------------------------------------------
// One warp per node is fine for the thousands of small nodes deep in the tree, but the first levels
// have a handful of nodes with millions of primitives each (the root of the 10 M-triangle soup was
// binned and partitioned by 32 lanes). Nodes above PT_SAH_LARGE_NODE primitives are therefore processed level
// by level with the work spread over tiles of kLargeTile primitives:
//   plan       tile ranges of this level's large nodes, their bin mapping, cleared bins
//   bin        every tile bins its primitives in shared memory and adds its bins to the node's
//   pick       a warp per node evaluates the 45 planes, allocates the two children
//   partition  every tile sends its primitives left or right (one atomic pair per tile for its two
//              output ranges) and accumulates the children's centroid bounds
//   copy       the partitioned ranges go back from tmp to order
// Children above that size go to the next level's list, the others to the queue of the warp-per-node
// kernel above, which finishes the tree. Tiles claim their output ranges in arrival order, so the order
// of primitives inside a node's range can differ from run to run; the split planes, and with them the
// tree and every hit, do not depend on it (only the pathological all-centroids-equal node splits by
// position, and such a node is handed to the warp-per-node kernel unsplit).

constexpr int kLargeTile = 2048;
constexpr int kLargeThreads = 256;
constexpr int kLargeItems = kLargeTile / kLargeThreads;

struct LargeNode
{
    int node, first, count;
    int tileStart;              // first tile of this node in the level's grid (plan)
    int parent, side;           // where the centroid bounds come from: parent's slot in the previous level's list (-1: scene bounds)
    float cmin[3], scale[3];    // centroid -> bin (plan); scale 0: no extent on that axis
    int axis, plane, nLeft;     // the split (pick); axis < 0: not split here
    int curL, curR;             // output cursors of the partition
    int clo[2][3], chi[2][3];   // children's centroid bounds, ordered ints (partition)
};

__device__ __forceinline__ int sahBinOf(float c, float cmin, float scale)
{
    int bin = (int)((c - cmin) * scale);
    return bin < 0 ? 0 : (bin > kSahBins - 1 ? kSahBins - 1 : bin);
}

// tile -> slot of its node in the level's list (tileStart is ascending)
__device__ __forceinline__ int largeSlotOfTile(const LargeNode* __restrict__ L, int nLarge, int tile)
{
    int lo = 0, hi = nLarge - 1;
    while (lo < hi)
    {
        const int mid = (lo + hi + 1) >> 1;
        if (L[mid].tileStart <= tile) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// one block: tile ranges and bin mappings of the level's nodes; counters[4] = tiles of the level
__global__ void __launch_bounds__(256) k_sahPlanLarge(LargeNode* L, int nLarge, const LargeNode* __restrict__ prev,
                                                     const int* __restrict__ bounds, SahBins* bins, uint32_t* counters)
{
    __shared__ int sTiles[256];
    // tiles per node, scanned in chunks of 256 nodes
    int running = 0;
    for (int base = 0; base < nLarge; base += 256)
    {
        const int i = base + threadIdx.x;
        const int tiles = i < nLarge ? (L[i].count + kLargeTile - 1) / kLargeTile : 0;
        sTiles[threadIdx.x] = tiles;
        __syncthreads();
        for (int o = 1; o < 256; o <<= 1)
        {
            const int v = threadIdx.x >= o ? sTiles[threadIdx.x - o] : 0;
            __syncthreads();
            sTiles[threadIdx.x] += v;
            __syncthreads();
        }
        if (i < nLarge) L[i].tileStart = running + sTiles[threadIdx.x] - tiles;
        running += sTiles[255];
        __syncthreads();
    }
    if (threadIdx.x == 0) counters[4] = (uint32_t)running;
    for (int i = threadIdx.x; i < nLarge; i += 256)
    {
        LargeNode& n = L[i];
        float lo[3], hi[3];
        for (int k = 0; k < 3; ++k)
        {
            if (n.parent < 0) { lo[k] = orderedToFloat(bounds[k]); hi[k] = orderedToFloat(bounds[3 + k]); }
            else { lo[k] = orderedToFloat(prev[n.parent].clo[n.side][k]); hi[k] = orderedToFloat(prev[n.parent].chi[n.side][k]); }
            const float ext = hi[k] - lo[k];
            n.cmin[k] = lo[k];
            n.scale[k] = ext > 0.0f ? (float)kSahBins / ext : 0.0f;
        }
        n.axis = -1; n.plane = 0; n.nLeft = 0; n.curL = 0; n.curR = 0;
        for (int s2 = 0; s2 < 2; ++s2)
            for (int k = 0; k < 3; ++k) { n.clo[s2][k] = floatToOrdered(3.0e38f); n.chi[s2][k] = floatToOrdered(-3.0e38f); }
    }
    // cleared bins
    int* b = reinterpret_cast<int*>(bins);
    const int words = (int)(sizeof(SahBins) / sizeof(int));
    for (int i = threadIdx.x; i < nLarge * words; i += 256)
    {
        const int w = i % words;
        b[i] = w < 3 * kSahBins * 3 ? floatToOrdered(3.0e38f) : (w < 2 * 3 * kSahBins * 3 ? floatToOrdered(-3.0e38f) : 0);
    }
}

__global__ void __launch_bounds__(kLargeThreads) k_sahBinLarge(const LargeNode* __restrict__ L, int nLarge,
                                                                const F4* __restrict__ leafLo, const F4* __restrict__ leafHi,
                                                                const int* __restrict__ order, SahBins* bins,
                                                                const uint32_t* __restrict__ counters)
{
    __shared__ SahBins B;
    const int tile = blockIdx.x;
    if (tile >= (int)counters[4]) return;
    const int slot = largeSlotOfTile(L, nLarge, tile);
    const LargeNode n = L[slot];
    for (int i = threadIdx.x; i < 3 * kSahBins; i += kLargeThreads)
    {
        const int a = i / kSahBins, b = i % kSahBins;
        for (int k = 0; k < 3; ++k) { B.lo[a][b][k] = floatToOrdered(3.0e38f); B.hi[a][b][k] = floatToOrdered(-3.0e38f); }
        B.cnt[a][b] = 0;
    }
    __syncthreads();
    const int begin = n.first + (tile - n.tileStart) * kLargeTile, end = min(begin + kLargeTile, n.first + n.count);
    for (int i = begin + threadIdx.x; i < end; i += kLargeThreads)
    {
        const int t = order[i];
        const F4 a = leafLo[t], b = leafHi[t];
        const float c[3] = { (a.x + b.x) * 0.5f, (a.y + b.y) * 0.5f, (a.z + b.z) * 0.5f };
        const int lo3[3] = { floatToOrdered(a.x), floatToOrdered(a.y), floatToOrdered(a.z) };
        const int hi3[3] = { floatToOrdered(b.x), floatToOrdered(b.y), floatToOrdered(b.z) };
        for (int ax = 0; ax < 3; ++ax)
        {
            if (n.scale[ax] == 0.0f) continue;
            const int bin = sahBinOf(c[ax], n.cmin[ax], n.scale[ax]);
            for (int k = 0; k < 3; ++k) { atomicMin(&B.lo[ax][bin][k], lo3[k]); atomicMax(&B.hi[ax][bin][k], hi3[k]); }
            atomicAdd(&B.cnt[ax][bin], 1);
        }
    }
    __syncthreads();
    SahBins& G = bins[slot];
    for (int i = threadIdx.x; i < 3 * kSahBins; i += kLargeThreads)
    {
        const int a = i / kSahBins, b = i % kSahBins;
        const int c = B.cnt[a][b];
        if (c == 0) continue;
        for (int k = 0; k < 3; ++k) { atomicMin(&G.lo[a][b][k], B.lo[a][b][k]); atomicMax(&G.hi[a][b][k], B.hi[a][b][k]); }
        atomicAdd(&G.cnt[a][b], c);
    }
}

// counters: [0] nodes allocated, [2] next level's large nodes, [3] nodes handed to the warp-per-node kernel
__global__ void __launch_bounds__(kSahWarps * 32) k_sahPickLarge(BinTree T, LargeNode* L, int nLarge, const SahBins* __restrict__ bins,
                                                                  LargeNode* next, int* smallQ, uint32_t* counters, int largeMin)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = blockIdx.x * kSahWarps + warp;
    if (slot >= nLarge) return;
    LargeNode& n = L[slot];
    const SahBins& B = bins[slot];
    float bestCost = 3.0e38f;
    int bestCand = -1;
    for (int cand = lane; cand < 3 * (kSahBins - 1); cand += 32)
    {
        const int ax = cand / (kSahBins - 1), sp = cand % (kSahBins - 1);
        if (n.scale[ax] == 0.0f) continue;
        float llo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, lhi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
        float rlo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, rhi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
        int nl = 0, nr = 0;
        for (int b = 0; b < kSahBins; ++b)
        {
            const int c = B.cnt[ax][b];
            if (c == 0) continue;
            if (b <= sp)
            {
                nl += c;
                for (int k = 0; k < 3; ++k) { llo[k] = fminf(llo[k], orderedToFloat(B.lo[ax][b][k])); lhi[k] = fmaxf(lhi[k], orderedToFloat(B.hi[ax][b][k])); }
            }
            else
            {
                nr += c;
                for (int k = 0; k < 3; ++k) { rlo[k] = fminf(rlo[k], orderedToFloat(B.lo[ax][b][k])); rhi[k] = fmaxf(rhi[k], orderedToFloat(B.hi[ax][b][k])); }
            }
        }
        if (nl == 0 || nr == 0) continue;
        const float al = boxArea(f4(llo[0], llo[1], llo[2], 0), f4(lhi[0], lhi[1], lhi[2], 0));
        const float ar = boxArea(f4(rlo[0], rlo[1], rlo[2], 0), f4(rhi[0], rhi[1], rhi[2], 0));
        const float cost = al * (float)nl + ar * (float)nr;
        if (cost < bestCost) { bestCost = cost; bestCand = cand; }
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        const float oc = __shfl_xor_sync(0xffffffffu, bestCost, o);
        const int ocand = __shfl_xor_sync(0xffffffffu, bestCand, o);
        if (oc < bestCost || (oc == bestCost && ocand >= 0 && (bestCand < 0 || ocand < bestCand))) { bestCost = oc; bestCand = ocand; }
    }
    if (lane != 0) return;
    if (bestCand < 0)
    {
        // every centroid in one bin on every axis: the warp-per-node kernel splits such a node by position
        smallQ[atomicAdd(&counters[3], 1u)] = n.node;
        return;
    }
    const int ax = bestCand / (kSahBins - 1), sp = bestCand % (kSahBins - 1);
    float clo[2][3], chi[2][3];
    for (int side = 0; side < 2; ++side)
        for (int k = 0; k < 3; ++k) { clo[side][k] = 3.0e38f; chi[side][k] = -3.0e38f; }
    int nLeft = 0;
    for (int b = 0; b < kSahBins; ++b)
    {
        if (B.cnt[ax][b] == 0) continue;
        const int side = b <= sp ? 0 : 1;
        if (side == 0) nLeft += B.cnt[ax][b];
        for (int k = 0; k < 3; ++k)
        {
            clo[side][k] = fminf(clo[side][k], orderedToFloat(B.lo[ax][b][k]));
            chi[side][k] = fmaxf(chi[side][k], orderedToFloat(B.hi[ax][b][k]));
        }
    }
    n.axis = ax; n.plane = sp; n.nLeft = nLeft;
    const uint32_t base = atomicAdd(&counters[0], 2u);
    T.lo[n.node].w = u2f(base);
    T.hi[n.node].w = u2f(base + 1u);
    const int kf[2] = { n.first, n.first + nLeft }, kc[2] = { nLeft, n.count - nLeft };
    for (int side = 0; side < 2; ++side)
    {
        const int k = (int)base + side;
        T.lo[k] = f4(clo[side][0], clo[side][1], clo[side][2], 0.0f);
        T.hi[k] = f4(chi[side][0], chi[side][1], chi[side][2], 0.0f);
        T.first[k] = kf[side]; T.last[k] = kf[side] + kc[side] - 1;
        if (kc[side] > largeMin)
        {
            LargeNode c;
            c.node = k; c.first = kf[side]; c.count = kc[side]; c.tileStart = 0; c.parent = slot; c.side = side;
            next[atomicAdd(&counters[2], 1u)] = c;
        }
        else if (kc[side] > kLeafMaxTris) smallQ[atomicAdd(&counters[3], 1u)] = k;
    }
}

__global__ void __launch_bounds__(kLargeThreads) k_sahPartitionLarge(LargeNode* L, int nLarge, const F4* __restrict__ leafLo,
                                                                      const F4* __restrict__ leafHi, const int* __restrict__ order,
                                                                      int* __restrict__ tmp, const uint32_t* __restrict__ counters)
{
    __shared__ int sCount[2], sBase[2], sFill[2];
    __shared__ int sClo[2][3], sChi[2][3];
    const int tile = blockIdx.x;
    if (tile >= (int)counters[4]) return;
    const int slot = largeSlotOfTile(L, nLarge, tile);
    LargeNode& n = L[slot];
    const int axis = n.axis;
    if (axis < 0) return;                   // not split here: its range stays as it is
    const int plane = n.plane, nodeFirst = n.first, nLeft = n.nLeft;
    const float cminA = n.cmin[axis], scaleA = n.scale[axis];
    if (threadIdx.x < 2) { sCount[threadIdx.x] = 0; sFill[threadIdx.x] = 0; }
    if (threadIdx.x < 6) { (&sClo[0][0])[threadIdx.x] = floatToOrdered(3.0e38f); (&sChi[0][0])[threadIdx.x] = floatToOrdered(-3.0e38f); }
    __syncthreads();
    const int begin = nodeFirst + (tile - n.tileStart) * kLargeTile, end = min(begin + kLargeTile, nodeFirst + n.count);
    int prim[kLargeItems];
    uint32_t leftBits = 0, validBits = 0;
    float clo[2][3], chi[2][3];
    for (int s2 = 0; s2 < 2; ++s2)
        for (int k = 0; k < 3; ++k) { clo[s2][k] = 3.0e38f; chi[s2][k] = -3.0e38f; }
    int myLeft = 0, myRight = 0;
    #pragma unroll
    for (int r = 0; r < kLargeItems; ++r)
    {
        const int i = begin + r * kLargeThreads + threadIdx.x;
        prim[r] = 0;
        if (i < end)
        {
            const int t = order[i];
            prim[r] = t;
            const F4 a = leafLo[t], b = leafHi[t];
            const float c[3] = { (a.x + b.x) * 0.5f, (a.y + b.y) * 0.5f, (a.z + b.z) * 0.5f };
            const bool left = sahBinOf(c[axis], cminA, scaleA) <= plane;
            validBits |= 1u << r;
            leftBits |= left ? (1u << r) : 0u;
            const int s2 = left ? 0 : 1;
            for (int k = 0; k < 3; ++k) { clo[s2][k] = fminf(clo[s2][k], c[k]); chi[s2][k] = fmaxf(chi[s2][k], c[k]); }
            myLeft += left ? 1 : 0; myRight += left ? 0 : 1;
        }
    }
    // tile totals -> one pair of global atomics for the tile's two output ranges
    int wl = myLeft, wr = myRight;
    for (int o = 16; o > 0; o >>= 1) { wl += __shfl_xor_sync(0xffffffffu, wl, o); wr += __shfl_xor_sync(0xffffffffu, wr, o); }
    for (int s2 = 0; s2 < 2; ++s2)
        for (int k = 0; k < 3; ++k) { clo[s2][k] = warpMinF(clo[s2][k]); chi[s2][k] = warpMaxF(chi[s2][k]); }
    if ((threadIdx.x & 31) == 0)
    {
        if (wl) atomicAdd(&sCount[0], wl);
        if (wr) atomicAdd(&sCount[1], wr);
        for (int s2 = 0; s2 < 2; ++s2)
            for (int k = 0; k < 3; ++k) { atomicMin(&sClo[s2][k], floatToOrdered(clo[s2][k])); atomicMax(&sChi[s2][k], floatToOrdered(chi[s2][k])); }
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        sBase[0] = nodeFirst + (sCount[0] ? atomicAdd(&n.curL, sCount[0]) : 0);
        sBase[1] = nodeFirst + nLeft + (sCount[1] ? atomicAdd(&n.curR, sCount[1]) : 0);
    }
    if (threadIdx.x < 6)
    {
        const int s2 = threadIdx.x / 3, k = threadIdx.x % 3;
        if (sCount[s2]) { atomicMin(&n.clo[s2][k], sClo[s2][k]); atomicMax(&n.chi[s2][k], sChi[s2][k]); }
    }
    __syncthreads();
    // positions inside the tile's ranges: one shared atomic per warp and side
    const int lane = threadIdx.x & 31;
    #pragma unroll
    for (int r = 0; r < kLargeItems; ++r)
    {
        const bool valid = (validBits >> r) & 1u, left = (leftBits >> r) & 1u;
        const unsigned ml = __ballot_sync(0xffffffffu, valid && left), mr = __ballot_sync(0xffffffffu, valid && !left);
        int bl = 0, br = 0;
        if (lane == 0)
        {
            if (ml) bl = atomicAdd(&sFill[0], __popc(ml));
            if (mr) br = atomicAdd(&sFill[1], __popc(mr));
        }
        bl = __shfl_sync(0xffffffffu, bl, 0); br = __shfl_sync(0xffffffffu, br, 0);
        const unsigned lt = (1u << lane) - 1u;
        if (valid) tmp[left ? (sBase[0] + bl + __popc(ml & lt)) : (sBase[1] + br + __popc(mr & lt))] = prim[r];
    }
}

__global__ void __launch_bounds__(kLargeThreads) k_sahCopyLarge(const LargeNode* __restrict__ L, int nLarge, int* __restrict__ order,
                                                                 const int* __restrict__ tmp, const uint32_t* __restrict__ counters)
{
    const int tile = blockIdx.x;
    if (tile >= (int)counters[4]) return;
    const int slot = largeSlotOfTile(L, nLarge, tile);
    const LargeNode& n = L[slot];
    if (n.axis < 0) return;
    const int begin = n.first + (tile - n.tileStart) * kLargeTile, end = min(begin + kLargeTile, n.first + n.count);
    for (int i = begin + threadIdx.x; i < end; i += kLargeThreads) order[i] = tmp[i];
}

struct CollapseWork { int32_t bin; uint32_t wide; };

// counters: [0] wide nodes allocated, [1] triangles allocated, [2] next-level queue length
__global__ void __launch_bounds__(128) k_collapse(BinTree T, const CollapseWork* __restrict__ qIn, int nIn,
                                                  CollapseWork* __restrict__ qOut, uint32_t* counters,
                                                  const uint32_t* __restrict__ vals, const F4* __restrict__ pos,
                                                  WideNode* nodes, PackedTri* tris)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nIn) return;
    CollapseWork w = qIn[i];
    int kids[8];
    int nk = collapseChildren(T, w.bin, kids);
    uint32_t nInt = 0, nTri = 0;
    for (int k = 0; k < nk; ++k)
    {
        if (binIsLeafCluster(T, kids[k])) nTri += (uint32_t)(T.last[kids[k]] - T.first[kids[k]] + 1);
        else nInt++;
    }
    uint32_t childBase = nInt ? atomicAdd(&counters[0], nInt) : 0u;
    uint32_t triBase = nTri ? atomicAdd(&counters[1], nTri) : 0u;
    uint32_t qBase = nInt ? atomicAdd(&counters[2], nInt) : 0u;
    WideNode node;
    encodeWideNode(node, T, kids, nk, childBase, triBase);
    nodes[w.wide] = node;
    uint32_t ci = 0, ti = 0;
    for (int k = 0; k < nk; ++k)
    {
        int kid = kids[k];
        if (binIsLeafCluster(T, kid))
        {
            for (int s = T.first[kid]; s <= T.last[kid]; ++s)
            {
                uint32_t t = vals[s];
                U4 a = ldg128(pos + 3 * (size_t)t), b = ldg128(pos + 3 * (size_t)t + 1), c = ldg128(pos + 3 * (size_t)t + 2);
                PackedTri pt;
                pt.v0x = u2f(a.x); pt.v0y = u2f(a.y); pt.v0z = u2f(a.z); pt.prim = t;
                pt.e1x = u2f(b.x) - u2f(a.x); pt.e1y = u2f(b.y) - u2f(a.y); pt.e1z = u2f(b.z) - u2f(a.z); pt.pad0 = 0;
                pt.e2x = u2f(c.x) - u2f(a.x); pt.e2y = u2f(c.y) - u2f(a.y); pt.e2z = u2f(c.z) - u2f(a.z); pt.pad1 = 0;
                tris[triBase + ti++] = pt;
            }
        }
        else
        {
            CollapseWork c; c.bin = kid; c.wide = childBase + ci;
            qOut[qBase + ci] = c;
            ++ci;
        }
    }
}

} // namespace

int radixSortBlocks(int n) { return (n + kSortTile - 1) / kSortTile; }

// LSD radix sort of (64-bit key, 32-bit value) pairs, `passes` digits of 8 bits from bit 0.
// Ping-pongs between the A and B arrays; returns true when the result is in the A arrays.
bool radixSortPairs(uint64_t* keysA, uint64_t* keysB, uint32_t* valsA, uint32_t* valsB, int n, int passes,
                    uint32_t* hist, cudaStream_t stream)
{
    const int sortBlocks = radixSortBlocks(n);
    uint64_t* kin = keysA; uint64_t* kout = keysB;
    uint32_t* vin = valsA; uint32_t* vout = valsB;
    for (int pass = 0; pass < passes; ++pass)
    {
        const int shift = pass * 8;
        k_sortHist<<<sortBlocks, kSortThreads, 0, stream>>>(kin, n, shift, hist, sortBlocks);
        k_sortScan<<<1, 1024, 0, stream>>>(hist, 256 * sortBlocks);
        k_sortScatter<<<sortBlocks, kSortThreads, 0, stream>>>(kin, vin, n, shift, hist, sortBlocks, kout, vout);
        g_counters.kernelLaunches += 3;
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    return kin == keysA;
}

namespace {

template <class T>
bool devAlloc(T*& p, size_t count)
{
    p = nullptr;
    if (count == 0) count = 1;
    return PT_CUDA(cudaMalloc((void**)&p, sizeof(T) * count));
}

// WideNode -> WideNodeF: the plane bytes as floats, scaled by the node's power-of-two cell size
__global__ void __launch_bounds__(256) k_expandNodes(const WideNode* __restrict__ nodes, int32_t n, WideNodeF* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const WideNode w = nodes[i];
    WideNodeF f;
    f.childBase = w.childBase;
    f.triBaseImask = (w.triBase & 0xffffffu) | ((w.exyzImask >> 24) << 24);
    f.meta[0] = w.meta[0]; f.meta[1] = w.meta[1];
    f.ox = w.ox; f.oy = w.oy; f.oz = w.oz; f.pad = 0.0f;
    const float sx = u2f((w.exyzImask & 0xffu) << 23), sy = u2f(((w.exyzImask >> 8) & 0xffu) << 23);
    const float sz = u2f(((w.exyzImask >> 16) & 0xffu) << 23);
    for (int c = 0; c < 8; ++c)
    {
        const int wd = c >> 2, sh = 8 * (c & 3);
        f.plane[0][c] = (float)((w.qlox[wd] >> sh) & 0xffu) * sx;       // exact: an 8-bit integer times 2^e
        f.plane[1][c] = (float)((w.qloy[wd] >> sh) & 0xffu) * sy;
        f.plane[2][c] = (float)((w.qloz[wd] >> sh) & 0xffu) * sz;
        f.plane[3][c] = (float)((w.qhix[wd] >> sh) & 0xffu) * sx;
        f.plane[4][c] = (float)((w.qhiy[wd] >> sh) & 0xffu) * sy;
        f.plane[5][c] = (float)((w.qhiz[wd] >> sh) & 0xffu) * sz;
    }
    out[i] = f;
}

// ---- refit ------------------------------------------------------------------

// the packed triangles again, from the new positions, in the leaf order the build gave them
__global__ void __launch_bounds__(256) k_refitTris(const F4* __restrict__ pos, int32_t n, PackedTri* __restrict__ tris)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t t = tris[j].prim;
    const U4 a = ldg128(pos + 3 * (size_t)t), b = ldg128(pos + 3 * (size_t)t + 1), c = ldg128(pos + 3 * (size_t)t + 2);
    PackedTri pt;
    pt.v0x = u2f(a.x); pt.v0y = u2f(a.y); pt.v0z = u2f(a.z); pt.prim = t;
    pt.e1x = u2f(b.x) - u2f(a.x); pt.e1y = u2f(b.y) - u2f(a.y); pt.e1z = u2f(b.z) - u2f(a.z); pt.pad0 = 0;
    pt.e2x = u2f(c.x) - u2f(a.x); pt.e2y = u2f(c.y) - u2f(a.y); pt.e2z = u2f(c.z) - u2f(a.z); pt.pad1 = 0;
    tris[j] = pt;
}

// one level of wide nodes, the levels below it done: child boxes from the leaves' vertices or the inner children's
// frames (boxLo / boxHi), then the node's own frame and plane bytes exactly as encodeWideNode makes them
__global__ void __launch_bounds__(128) k_refitLevel(WideNode* __restrict__ nodes, int32_t begin, int32_t end,
                                                    const PackedTri* __restrict__ tris, const F4* __restrict__ pos,
                                                    F4* __restrict__ boxLo, F4* __restrict__ boxHi)
{
    const int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    WideNode w = nodes[i];
    const uint32_t imask = w.exyzImask >> 24;
    F4 clo[8], chi[8];
    F4 lo = f4(3.0e38f, 3.0e38f, 3.0e38f, 0), hi = f4(-3.0e38f, -3.0e38f, -3.0e38f, 0);
    uint32_t present = 0;
    for (int c = 0; c < 8; ++c)
    {
        const uint32_t m = (w.meta[c >> 2] >> (8 * (c & 3))) & 0xffu;
        if (m == 0u) continue;
        F4 a = f4(3.0e38f, 3.0e38f, 3.0e38f, 0), b = f4(-3.0e38f, -3.0e38f, -3.0e38f, 0);
        if ((imask >> c) & 1u)
        {
            const uint32_t child = w.childBase + (uint32_t)__popc(imask & ((1u << c) - 1u));
            a = boxLo[child]; b = boxHi[child];
        }
        else
        {
            const uint32_t first = w.triBase + (m & 31u), count = m >> 5;
            for (uint32_t j = 0; j < count; ++j)
            {
                const uint32_t t = tris[first + j].prim;
                for (int v = 0; v < 3; ++v)
                {
                    const U4 p = ldg128(pos + 3 * (size_t)t + v);
                    a.x = fminf(a.x, u2f(p.x)); a.y = fminf(a.y, u2f(p.y)); a.z = fminf(a.z, u2f(p.z));
                    b.x = fmaxf(b.x, u2f(p.x)); b.y = fmaxf(b.y, u2f(p.y)); b.z = fmaxf(b.z, u2f(p.z));
                }
            }
        }
        clo[c] = a; chi[c] = b;
        present |= 1u << c;
        lo.x = fminf(lo.x, a.x); lo.y = fminf(lo.y, a.y); lo.z = fminf(lo.z, a.z);
        hi.x = fmaxf(hi.x, b.x); hi.y = fmaxf(hi.y, b.y); hi.z = fmaxf(hi.z, b.z);
    }
    boxLo[i] = lo; boxHi[i] = hi;
    const uint32_t ex = quantExp(hi.x - lo.x), ey = quantExp(hi.y - lo.y), ez = quantExp(hi.z - lo.z);
    const float sx = u2f(ex << 23), sy = u2f(ey << 23), sz = u2f(ez << 23);
    w.ox = lo.x; w.oy = lo.y; w.oz = lo.z;
    w.exyzImask = ex | (ey << 8) | (ez << 16) | (imask << 24);
    w.qlox[0] = w.qlox[1] = w.qloy[0] = w.qloy[1] = w.qloz[0] = w.qloz[1] = 0;
    w.qhix[0] = w.qhix[1] = w.qhiy[0] = w.qhiy[1] = w.qhiz[0] = w.qhiz[1] = 0;
    for (int c = 0; c < 8; ++c)
    {
        if (!((present >> c) & 1u))
        {
            setByte(w.qlox, c, 255); setByte(w.qloy, c, 255); setByte(w.qloz, c, 255);     // empty slot: inverted box
            continue;
        }
        int ql, qh;
        quantRange(clo[c].x, chi[c].x, w.ox, sx, ql, qh); setByte(w.qlox, c, (uint32_t)ql); setByte(w.qhix, c, (uint32_t)qh);
        quantRange(clo[c].y, chi[c].y, w.oy, sy, ql, qh); setByte(w.qloy, c, (uint32_t)ql); setByte(w.qhiy, c, (uint32_t)qh);
        quantRange(clo[c].z, chi[c].z, w.oz, sz, ql, qh); setByte(w.qloz, c, (uint32_t)ql); setByte(w.qhiz, c, (uint32_t)qh);
    }
    nodes[i] = w;
}

} // namespace

bool refitBvh(const F4* dPositions, int32_t triCount, BvhBuffers& b, cudaStream_t stream)
{
    if (triCount != b.triCount || b.nodeCount <= 0 || b.levelEnd.empty() || b.levelEnd.back() != b.nodeCount)
    {
        setError("refitBvh: the tree does not match the geometry (%d triangles against %d, %d nodes)", triCount, b.triCount, b.nodeCount);
        return false;
    }
    F4 *boxLo = nullptr, *boxHi = nullptr;
    bool ok = devAlloc(boxLo, (size_t)b.nodeCount) && devAlloc(boxHi, (size_t)b.nodeCount);
    if (ok)
    {
        k_refitTris<<<(triCount + 255) / 256, 256, 0, stream>>>(dPositions, triCount, b.tris);
        g_counters.kernelLaunches++;
        for (int l = (int)b.levelEnd.size() - 1; l >= 0; --l)
        {
            const int32_t begin = l == 0 ? 0 : b.levelEnd[(size_t)l - 1], end = b.levelEnd[(size_t)l];
            if (end <= begin) continue;
            k_refitLevel<<<(end - begin + 127) / 128, 128, 0, stream>>>(b.nodes, begin, end, b.tris, dPositions, boxLo, boxHi);
            g_counters.kernelLaunches++;
        }
        if (b.nodesF)
        {
            k_expandNodes<<<(b.nodeCount + 255) / 256, 256, 0, stream>>>(b.nodes, b.nodeCount, b.nodesF);
            g_counters.kernelLaunches++;
        }
        F4 root[2];
        ok = PT_CUDA(cudaGetLastError()) &&
             PT_CUDA(cudaMemcpyAsync(&root[0], boxLo, sizeof(F4), cudaMemcpyDeviceToHost, stream)) &&
             PT_CUDA(cudaMemcpyAsync(&root[1], boxHi, sizeof(F4), cudaMemcpyDeviceToHost, stream)) &&
             PT_CUDA(cudaStreamSynchronize(stream));
        if (ok)
        {
            b.sceneLo[0] = root[0].x; b.sceneLo[1] = root[0].y; b.sceneLo[2] = root[0].z;
            b.sceneHi[0] = root[1].x; b.sceneHi[1] = root[1].y; b.sceneHi[2] = root[1].z;
        }
    }
    cudaFree(boxLo); cudaFree(boxHi);
    return ok;
}


void freeBvh(BvhBuffers& b)
{
    if (b.nodes) cudaFree(b.nodes);
    if (b.nodesF) cudaFree(b.nodesF);
    if (b.tris) cudaFree(b.tris);
    b = BvhBuffers();
}

bool buildBvh(const F4* dPositions, int32_t n, BvhBuffers& out, cudaStream_t stream)
{
    out = BvhBuffers();
    if (n <= 0) return true;

    const int total = 2 * n - 1;
    F4 *leafLo = nullptr, *leafHi = nullptr, *lo = nullptr, *hi = nullptr;
    int *bounds = nullptr, *first = nullptr, *last = nullptr, *parent = nullptr, *visits = nullptr;
    uint64_t *keysA = nullptr, *keysB = nullptr;
    uint32_t *valsA = nullptr, *valsB = nullptr, *hist = nullptr, *counters = nullptr;
    CollapseWork *q0 = nullptr, *q1 = nullptr;
    const int sortBlocks = (n + kSortTile - 1) / kSortTile;
    bool ok = devAlloc(leafLo, n) && devAlloc(leafHi, n) && devAlloc(lo, total) && devAlloc(hi, total) &&
              devAlloc(bounds, 12) && devAlloc(first, total) && devAlloc(last, total) && devAlloc(parent, total) &&
              devAlloc(visits, n) && devAlloc(keysA, n) && devAlloc(keysB, n) && devAlloc(valsA, n) &&
              devAlloc(valsB, n) && devAlloc(hist, (size_t)256 * sortBlocks) && devAlloc(counters, 8) &&
              devAlloc(q0, n) && devAlloc(q1, n) && devAlloc(out.nodes, n) && devAlloc(out.tris, n);
    const int blocks = (n + 255) / 256;
    if (ok)
    {
        k_initBounds<<<1, 32, 0, stream>>>(bounds);
        k_triBounds<<<blocks, 256, 0, stream>>>(dPositions, n, leafLo, leafHi, bounds);
        k_morton<<<blocks, 256, 0, stream>>>(leafLo, leafHi, n, bounds, keysA, valsA);
        g_counters.kernelLaunches += 3;
        const uint32_t* sortedVals = nullptr;
        if (n <= sahMaxTris())
        {
            // binned SAH: order[] doubles as the sorted-primitive array the collapse reads
            int* order = reinterpret_cast<int*>(valsA);
            int* tmp = reinterpret_cast<int*>(valsB);
            int* sq[2] = { reinterpret_cast<int*>(q0), reinterpret_cast<int*>(q1) };
            BinTree T; T.lo = lo; T.hi = hi; T.first = first; T.last = last; T.n = n;
            k_sahInit<<<blocks, 256, 0, stream>>>(T, n, bounds, order, sq[0], counters);
            g_counters.kernelLaunches += 1;
            int cur = 0;
            int nSmall = n > kLeafMaxTris ? 1 : 0;      // k_sahInit queued the root for the warp-per-node kernel
            if (ok && n > largeNodeMin())
            {
                // the large-node levels (see above); the warp-per-node queue fills up as nodes get small enough
                const int maxLarge = n / std::max(largeNodeMin(), 1) + 2;
                const unsigned maxTiles = (unsigned)(n / kLargeTile + maxLarge + 1);
                LargeNode* dLarge[2] = { nullptr, nullptr };
                SahBins* dBins = nullptr;
                ok = devAlloc(dLarge[0], (size_t)maxLarge) && devAlloc(dLarge[1], (size_t)maxLarge) && devAlloc(dBins, (size_t)maxLarge);
                LargeNode root;
                memset(&root, 0, sizeof(root));
                root.node = 0; root.first = 0; root.count = n; root.parent = -1;
                const uint32_t init[5] = { 1u, 0u, 0u, 0u, 0u };
                ok = ok && PT_CUDA(cudaMemcpyAsync(dLarge[0], &root, sizeof(root), cudaMemcpyHostToDevice, stream));
                ok = ok && PT_CUDA(cudaMemcpyAsync(counters, init, sizeof(init), cudaMemcpyHostToDevice, stream));
                int nLarge = 1, lcur = 0;
                while (ok && nLarge > 0)
                {
                    k_sahPlanLarge<<<1, 256, 0, stream>>>(dLarge[lcur], nLarge, dLarge[lcur ^ 1], bounds, dBins, counters);
                    k_sahBinLarge<<<maxTiles, kLargeThreads, 0, stream>>>(dLarge[lcur], nLarge, leafLo, leafHi, order, dBins, counters);
                    k_sahPickLarge<<<(nLarge + kSahWarps - 1) / kSahWarps, kSahWarps * 32, 0, stream>>>(T, dLarge[lcur], nLarge, dBins,
                                                                                                          dLarge[lcur ^ 1], sq[0], counters, largeNodeMin());
                    k_sahPartitionLarge<<<maxTiles, kLargeThreads, 0, stream>>>(dLarge[lcur], nLarge, leafLo, leafHi, order, tmp, counters);
                    k_sahCopyLarge<<<maxTiles, kLargeThreads, 0, stream>>>(dLarge[lcur], nLarge, order, tmp, counters);
                    g_counters.kernelLaunches += 5;
                    uint32_t h[4];
                    ok = ok && PT_CUDA(cudaGetLastError());
                    ok = ok && PT_CUDA(cudaMemcpyAsync(h, counters, sizeof(h), cudaMemcpyDeviceToHost, stream));
                    ok = ok && PT_CUDA(cudaStreamSynchronize(stream));
                    if (!ok) break;
                    nLarge = (int)h[2];
                    nSmall = (int)h[3];
                    ok = PT_CUDA(cudaMemsetAsync(counters + 2, 0, sizeof(uint32_t), stream));
                    lcur ^= 1;
                    if (nLarge > maxLarge) { setError("buildBvh: large-node list overflow"); ok = false; }
                }
                cudaFree(dLarge[0]); cudaFree(dLarge[1]); cudaFree(dBins);
                // hand over to the warp-per-node kernel: its input queue is sq[0], its level counter counters[1]
                const uint32_t qlen = (uint32_t)nSmall;
                ok = ok && PT_CUDA(cudaMemcpyAsync(counters + 1, &qlen, sizeof(qlen), cudaMemcpyHostToDevice, stream));
            }
            for (;;)
            {
                uint32_t h[2];
                ok = ok && PT_CUDA(cudaMemcpyAsync(h, counters, sizeof(h), cudaMemcpyDeviceToHost, stream));
                ok = ok && PT_CUDA(cudaStreamSynchronize(stream));
                if (!ok || h[1] == 0) break;
                ok = ok && PT_CUDA(cudaMemsetAsync(counters + 1, 0, sizeof(uint32_t), stream));
                const int nIn = (int)h[1];
                k_sahSplit<<<(nIn + kSahWarps - 1) / kSahWarps, kSahWarps * 32, 0, stream>>>(T, leafLo, leafHi, order, tmp,
                                                                                             sq[cur], nIn, sq[cur ^ 1], counters);
                g_counters.kernelLaunches += 1;
                cur ^= 1;
            }
            sortedVals = valsA;
        }
        else
        {
        radixSortPairs(keysA, keysB, valsA, valsB, n, 8, hist, stream);
        uint64_t* kin = keysA;      // 8 passes: result is back in keysA / valsA
        uint32_t* vin = valsA;
        ok = ok && PT_CUDA(cudaMemsetAsync(parent, 0xff, sizeof(int) * total, stream));
        ok = ok && PT_CUDA(cudaMemsetAsync(visits, 0, sizeof(int) * n, stream));
        if (n > 1) k_radixTree<<<(n - 1 + 255) / 256, 256, 0, stream>>>(kin, n, lo, hi, first, last, parent);
        k_leafInit<<<blocks, 256, 0, stream>>>(vin, leafLo, leafHi, n, lo, hi, first, last);
        if (n > 1) k_refit<<<blocks, 256, 0, stream>>>(n, lo, hi, parent, visits);
        g_counters.kernelLaunches += 3;
        sortedVals = vin;
        }

        // breadth-first collapse; the root wide node is index 0
        CollapseWork root; root.bin = (n == 1) ? 0 : 0; root.wide = 0;
        uint32_t initCounters[4] = { 1u, 0u, 0u, 0u };
        ok = ok && PT_CUDA(cudaMemcpyAsync(q0, &root, sizeof(root), cudaMemcpyHostToDevice, stream));
        ok = ok && PT_CUDA(cudaMemcpyAsync(counters, initCounters, sizeof(initCounters), cudaMemcpyHostToDevice, stream));
        BinTree T; T.lo = lo; T.hi = hi; T.first = first; T.last = last; T.n = n;
        int levelCount = 1;
        CollapseWork* qin = q0; CollapseWork* qout = q1;
        while (ok && levelCount > 0)
        {
            k_collapse<<<(levelCount + 127) / 128, 128, 0, stream>>>(T, qin, levelCount, qout, counters, sortedVals,
                                                                     dPositions, out.nodes, out.tris);
            g_counters.kernelLaunches += 1;
            uint32_t h[4];
            ok = ok && PT_CUDA(cudaMemcpyAsync(h, counters, sizeof(h), cudaMemcpyDeviceToHost, stream));
            ok = ok && PT_CUDA(cudaStreamSynchronize(stream));
            levelCount = (int)h[2];
            out.depth++;
            if (out.levelEnd.empty()) out.levelEnd.push_back(1);       // the root
            if (levelCount > 0) out.levelEnd.push_back((int32_t)h[0]);
            out.nodeCount = (int32_t)h[0];
            out.triCount = (int32_t)h[1];
            ok = ok && PT_CUDA(cudaMemsetAsync(counters + 2, 0, sizeof(uint32_t), stream));
            std::swap(qin, qout);
        }
        int hb[12];
        ok = ok && PT_CUDA(cudaMemcpyAsync(hb, bounds, sizeof(hb), cudaMemcpyDeviceToHost, stream));
        ok = ok && PT_CUDA(cudaStreamSynchronize(stream));
        if (ok)
        {
            for (int k = 0; k < 3; ++k)
            {
                int a = hb[6 + k], b = hb[9 + k];
                union { int i; float f; } ua, ub;
                ua.i = (a >= 0) ? a : (a ^ 0x7fffffff);
                ub.i = (b >= 0) ? b : (b ^ 0x7fffffff);
                out.sceneLo[k] = ua.f; out.sceneHi[k] = ub.f;
            }
        }
    }
    if (ok && out.nodeCount > 0 && out.nodeCount <= floatNodeMaxNodes() && out.triCount < (1 << 24))
    {
        ok = devAlloc(out.nodesF, out.nodeCount);
        if (ok)
        {
            k_expandNodes<<<(out.nodeCount + 255) / 256, 256, 0, stream>>>(out.nodes, out.nodeCount, out.nodesF);
            g_counters.kernelLaunches++;
            ok = PT_CUDA(cudaGetLastError()) && PT_CUDA(cudaStreamSynchronize(stream));
        }
    }
    cudaFree(leafLo); cudaFree(leafHi); cudaFree(lo); cudaFree(hi); cudaFree(bounds);
    cudaFree(first); cudaFree(last); cudaFree(parent); cudaFree(visits);
    cudaFree(keysA); cudaFree(keysB); cudaFree(valsA); cudaFree(valsB); cudaFree(hist); cudaFree(counters);
    cudaFree(q0); cudaFree(q1);
    if (!ok) freeBvh(out);
    return ok;
}

} // namespace pt
