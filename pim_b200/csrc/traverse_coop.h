// traverse_coop.h — closest hit / any hit with eight lanes per ray.
//
// The one-thread-per-ray traversal (traverse.h) spends ~300 instructions per lane on every wide
// node: one lane converts and tests all eight child boxes (48 byte-to-float conversions on the
// quarter-rate XU pipe, 48 fma, the min / max chains, the key sort), and on large incoherent
// trees only ~15 of 32 lanes are in the same phase at a time. Here a wide node's eight children
// are the SIMD dimension instead: a group of eight lanes owns a ray, lane i tests child box i
// (6 conversions, 6 fma), a ballot collects the hits, every lane ranks its own child among the
// hit ones with eight shuffles and writes its own stack entry, and the triangles of hit leaf
// children are tested by the lanes that own those children. A warp carries four rays; all four
// groups run the same instruction stream (one node step per loop iteration each), so the box
// test executes with every lane of every busy group. The arithmetic per child and per triangle
// is the scalar kernel's (same expressions, same fma contraction), the result is the minimum
// over (t, primitive id) either way, so hits are bit-identical; only the order in which nodes are
// visited differs.
//
// Stack: kCoopStackDepth entries of 8 bytes per group in shared memory. A node pushes at most seven
// entries and the stack holds entries of the levels on the current path only, so the launch functions
// take this kernel only for trees with 7 * depth <= kCoopStackDepth (Bvh::depth); the overflow flag is
// a debugging aid that stays zero.
#pragma once

#include "traverse.h"

namespace pt {

constexpr int kCoopThreads = 128;                       // 16 groups of 8 lanes
constexpr int kCoopGroups = kCoopThreads / 8;
constexpr int kCoopStackDepth = 96;
constexpr int kCoopChunk = 8;                           // rays a group claims per atomic

struct CoopRay
{
    V3 ro, rd, idir;
    float tNear, tBest;
    uint32_t bestPrim;
    float bu, bv;
    bool found, negx, negy, negz;
    uint32_t cur;
    bool haveCur;
    int sp;
};

__device__ __forceinline__ void coopBegin(const Bvh& bvh, CoopRay& r, V3 ro, V3 rd, float tNear, float tFar)
{
    r.ro = ro; r.rd = rd;
    r.idir = v3(safeRcp(rd.x), safeRcp(rd.y), safeRcp(rd.z));
    r.negx = rd.x < 0.0f; r.negy = rd.y < 0.0f; r.negz = rd.z < 0.0f;
    r.tNear = tNear; r.tBest = tFar;
    r.found = false; r.bestPrim = 0xffffffffu; r.bu = r.bv = 0.0f;
    r.cur = 0; r.haveCur = bvh.nodeCount > 0;
    r.sp = 0;
}

// pops until an entry survives the distance cull; false when the stack is empty (ray finished)
__device__ __forceinline__ bool coopPop(CoopRay& r, const unsigned long long* st)
{
    while (r.sp > 0)
    {
        const unsigned long long e = st[--r.sp];
        const uint32_t key = (uint32_t)(e >> 32);
        if (u2f((key - 8u) & ~7u) > r.tBest) continue;
        r.cur = (uint32_t)e;
        r.haveCur = true;
        return true;
    }
    return false;
}

// One wide node for one group: sub = lane within the group (its child), gmask = the group's lanes,
// gshift = 8 * group index within the warp. Returns true when an occlusion ray found a hit.
template <bool kAnyHit>
__device__ __forceinline__ bool coopNodeStep(const Bvh& bvh, CoopRay& r, unsigned long long* st, int sub, unsigned gmask,
                                             int gshift, int* overflow)
{
    r.haveCur = false;
    const WideNode* np = bvh.nodes + r.cur;
    const U4 n0 = ldg128(&np->ox);                 // the same 16 bytes for all eight lanes: one broadcast each
    const U4 n1 = ldg128(&np->childBase);
    const unsigned char* qb = reinterpret_cast<const unsigned char*>(np) + 32;    // six arrays of eight plane bytes
    const uint32_t qlx = __ldg(qb + sub), qly = __ldg(qb + 8 + sub), qlz = __ldg(qb + 16 + sub);
    const uint32_t qhx = __ldg(qb + 24 + sub), qhy = __ldg(qb + 32 + sub), qhz = __ldg(qb + 40 + sub);
    const uint32_t imask = n0.w >> 24;
    const float sx = u2f((n0.w & 0xffu) << 23) * r.idir.x;
    const float sy = u2f(((n0.w >> 8) & 0xffu) << 23) * r.idir.y;
    const float sz = u2f(((n0.w >> 16) & 0xffu) << 23) * r.idir.z;
    const float bx = (u2f(n0.x) - r.ro.x) * r.idir.x;
    const float by = (u2f(n0.y) - r.ro.y) * r.idir.y;
    const float bz = (u2f(n0.z) - r.ro.z) * r.idir.z;
    const float tnx = (float)(r.negx ? qhx : qlx) * sx + bx, tfx = (float)(r.negx ? qlx : qhx) * sx + bx;
    const float tny = (float)(r.negy ? qhy : qly) * sy + by, tfy = (float)(r.negy ? qly : qhy) * sy + by;
    const float tnz = (float)(r.negz ? qhz : qlz) * sz + bz, tfz = (float)(r.negz ? qlz : qhz) * sz + bz;
    const float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, r.tNear));
    const float tmax = fminf(fminf(tfx, tfy), fminf(tfz, r.tBest));
    const uint32_t metaw = sub < 4 ? n1.z : n1.w;
    const uint32_t m = (metaw >> (8 * (sub & 3))) & 0xffu;
    const bool internal = (imask >> sub) & 1u;
    const uint32_t triCount = m >> 5;
    bool h = tmin <= tmax * 1.0000004f;
    h = h && (internal || triCount != 0u);

    // ---- triangles of the hit leaf children: each by the lane that owns the child
    const bool leafHit = h && !internal;
    if (__ballot_sync(gmask, leafHit) & gmask)
    {
        float lt = 3.0e38f, lu = 0.0f, lv = 0.0f;
        uint32_t lprim = 0xffffffffu;
        bool lfound = false;
        if (leafHit)
        {
            const uint32_t first = n1.y + (m & 31u);
            for (uint32_t j = 0; j < triCount; ++j)
            {
                float t, u, v; uint32_t prim; V3 ng;
                if (triTest(bvh.tris + first + j, r.ro, r.rd, r.tNear, r.tBest, t, u, v, prim, ng))
                {
                    if (!lfound || t < lt || (t == lt && prim < lprim)) { lfound = true; lt = t; lu = u; lv = v; lprim = prim; }
                }
            }
        }
        if (kAnyHit)
        {
            if (__ballot_sync(gmask, lfound) & gmask) { r.found = true; return true; }
        }
        else
        {
            // minimum over (t, primitive id) across the group
            #pragma unroll
            for (int o = 4; o > 0; o >>= 1)
            {
                const float ot = __shfl_xor_sync(gmask, lt, o), ou = __shfl_xor_sync(gmask, lu, o), ov = __shfl_xor_sync(gmask, lv, o);
                const uint32_t op = __shfl_xor_sync(gmask, lprim, o);
                if (ot < lt || (ot == lt && op < lprim)) { lt = ot; lu = ou; lv = ov; lprim = op; }
            }
            if (lprim != 0xffffffffu && (!r.found || lt < r.tBest || (lt == r.tBest && lprim < r.bestPrim)))
            {
                r.found = true; r.tBest = lt; r.bestPrim = lprim; r.bu = lu; r.bv = lv;
            }
        }
    }

    // ---- inner children: the nearest becomes the current node, the others go onto the stack far to near
    const bool push = h && internal && tmin <= r.tBest * 1.0000004f;
    const uint32_t key = ((f2u(tmin) & ~7u) | (uint32_t)sub) + 8u;
    const uint32_t mine = push ? key : 0u;
    const unsigned pushBits = (__ballot_sync(gmask, push) >> gshift) & 0xffu;
    if (pushBits)
    {
        const int nInt = __popc(pushBits);
        int farther = 0;
        const int leader = gshift;          // first lane of the group
        #pragma unroll
        for (int j = 0; j < 8; ++j)
        {
            const uint32_t kj = __shfl_sync(gmask, mine, leader + j);
            farther += (kj > mine) ? 1 : 0;
        }
        const uint32_t child = n1.x + (uint32_t)__popc(imask & ((1u << sub) - 1u));
        const bool nearest = push && farther == nInt - 1;
        if (push && !nearest)
        {
            const int pos = r.sp + farther;
            if (pos < kCoopStackDepth) st[pos] = ((unsigned long long)key << 32) | child;
            else *overflow = 1;
        }
        const unsigned nearBits = __ballot_sync(gmask, nearest) & gmask;
        r.cur = __shfl_sync(gmask, child, __ffs((int)nearBits) - 1);
        r.haveCur = true;
        r.sp = min(r.sp + nInt - 1, kCoopStackDepth);
        __syncwarp(gmask);                  // the entries are read by every lane of the group
    }
    return false;
}

} // namespace pt
