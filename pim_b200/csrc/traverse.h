// traverse.h — one-ray-per-thread traversal of the 8-wide compressed BVH:
// closest hit (rtcIntersect1), any hit (rtcOccluded16) and nearest triangle
// within a radius (rtcPointQuery + RtcPointQueryFn, path_tracer.c:555-616).
#pragma once

#include "bvh.h"

namespace pt {

// Traversal stack: the first kSharedDepth entries live in caller-provided memory
// addressed [depth * stride] (shared memory, stride = block size, so a warp's
// accesses hit 32 different banks whatever the per-thread depths are); deeper
// entries spill to a small per-thread array.
template <int kSharedDepth>
struct TravStack
{
    unsigned long long* base;
    int stride;
    int sp;
    // 7 pushes per wide node at most; postponed leaves (below) put leaf children here as well
    static constexpr int kSpillDepth = 72;
    unsigned long long spill[kSpillDepth];

    PT_D void init(unsigned long long* b, int s) { base = b; stride = s; sp = 0; }
    PT_D void push(uint32_t node, uint32_t key)
    {
        unsigned long long e = ((unsigned long long)key << 32) | node;
        if (sp < kSharedDepth) base[sp * stride] = e;
        else if (sp - kSharedDepth < kSpillDepth) spill[sp - kSharedDepth] = e;
        ++sp;
    }
    // An entry pushed beyond the capacity was dropped; popping it yields kDropped, whose key decodes to
    // +inf so every traversal loop culls it (the subtree is skipped instead of the ray restarting at the
    // root, which is what a zero entry would mean) and whose node id pointQuery recognises. buildBvh
    // reports the wide tree's depth (BvhBuffers::depth) so that a tree deep enough to get here is visible.
    static constexpr unsigned long long kDropped = (0x7f800008ull << 32) | 0xffffffffull;
    PT_D unsigned long long pop()
    {
        --sp;
        if (sp < kSharedDepth) return base[sp * stride];
        return (sp - kSharedDepth < kSpillDepth) ? spill[sp - kSharedDepth] : kDropped;
    }
};

PT_D float safeRcp(float d)
{
    // keep the slab arithmetic finite for axis-parallel rays
    const float tiny = 1e-20f;
    if (f1_abs(d) < tiny) d = (f2u(d) >> 31) ? -tiny : tiny;
    return 1.0f / d;
}

struct TravCounters { uint32_t nodes, tris; };


// byte `idx` (0..3) of `word` as a float: I2F.U8, 48 per wide node on the XU pipe (70 % busy in the
// traversal kernels). Every way around it measured slower on B200: splicing the byte into a float
// mantissa with PRMT + FADD -4 %; one dp4a per plane (2^23 + 128 q, folded into the fma) -4 %
// on cornell and -18 % on the 10 M-triangle soup; PRMT into the mantissa of 32768.0f with the 32768 folded
// into the fma's constant (no extra instruction, planes widened by 0.004 steps) 334 -> 320 Mrays/s on the
// soup: the ALU pipe that takes the PRMTs is as busy as the XU pipe they leave.
PT_D float byteToFloat(uint32_t word, int idx) { return (float)((word >> (8 * idx)) & 0xffu); }


PT_D uint32_t max8(const uint32_t k[8])
{
    uint32_t a = k[0] > k[1] ? k[0] : k[1];
    uint32_t b = k[2] > k[3] ? k[2] : k[3];
    uint32_t c = k[4] > k[5] ? k[4] : k[5];
    uint32_t d = k[6] > k[7] ? k[6] : k[7];
    a = a > b ? a : b;
    c = c > d ? c : d;
    return a > c ? a : c;
}

// Traversal state of one ray. travBegin sets it up, every travStep call visits one wide node
// (its child boxes and the triangles of its leaf children) and returns false once the ray is
// finished, so a kernel can either loop it to completion (traverse) or interleave it with
// fetching new rays for idle lanes (the persistent kernels of wavefront.cu).
struct TravRay
{
    V3 ro, rd, idir;
    float tNear, tBest;
    uint32_t bestPrim;
    uint32_t cur;
    bool haveCur, found, negx, negy, negz;
    uint32_t pendFirst, pendCount;      // a popped leaf waiting for its triangle tests (travNodeStepP)
    Hit hit;
};

template <class Stack>
PT_D void travBegin(const Bvh& bvh, TravRay& r, V3 ro, V3 rd, float tNear, float tFar, Stack& st)
{
    r.hit.prim = -1;
    r.hit.t = tFar;
    r.hit.u = r.hit.v = 0.0f;
    r.hit.ng = v3(0, 0, 0);
    r.ro = ro; r.rd = rd;
    r.idir = v3(safeRcp(rd.x), safeRcp(rd.y), safeRcp(rd.z));
    r.negx = rd.x < 0.0f; r.negy = rd.y < 0.0f; r.negz = rd.z < 0.0f;
    r.tNear = tNear; r.tBest = tFar;
    r.found = false;
    r.bestPrim = 0xffffffffu;
    r.cur = 0;
    r.haveCur = bvh.nodeCount > 0;
    r.pendFirst = 0u; r.pendCount = 0u;
    st.sp = 0;
}

template <bool kAnyHit, bool kCount, class Stack>
PT_D bool travStep(const Bvh& bvh, TravRay& r, Stack& st, TravCounters* cnt)
{
    while (!r.haveCur)
    {
        if (st.sp == 0) return false;
        unsigned long long e = st.pop();
        uint32_t key = (uint32_t)(e >> 32);
        if (u2f((key - 8u) & ~7u) > r.tBest) continue;
        r.cur = (uint32_t)e;
        r.haveCur = true;
    }
    r.haveCur = false;
    if (kCount) cnt->nodes++;

    const WideNode* np = bvh.nodes + r.cur;
    const U4 n0 = ldg128(&np->ox);
    const U4 n1 = ldg128(&np->childBase);
    const U4 n2 = ldg128(&np->qlox[0]);
    const U4 n3 = ldg128(&np->qloz[0]);
    const U4 n4 = ldg128(&np->qhiy[0]);

    const uint32_t imask = n0.w >> 24;
    const float sx = u2f((n0.w & 0xffu) << 23) * r.idir.x;
    const float sy = u2f(((n0.w >> 8) & 0xffu) << 23) * r.idir.y;
    const float sz = u2f(((n0.w >> 16) & 0xffu) << 23) * r.idir.z;
    const float bx = (u2f(n0.x) - r.ro.x) * r.idir.x;
    const float by = (u2f(n0.y) - r.ro.y) * r.idir.y;
    const float bz = (u2f(n0.z) - r.ro.z) * r.idir.z;

    // near / far plane bytes per axis, picked by the ray's direction sign
    uint32_t nxw[2], fxw[2], nyw[2], fyw[2], nzw[2], fzw[2];
    nxw[0] = r.negx ? n3.z : n2.x; nxw[1] = r.negx ? n3.w : n2.y;
    fxw[0] = r.negx ? n2.x : n3.z; fxw[1] = r.negx ? n2.y : n3.w;
    nyw[0] = r.negy ? n4.x : n2.z; nyw[1] = r.negy ? n4.y : n2.w;
    fyw[0] = r.negy ? n2.z : n4.x; fyw[1] = r.negy ? n2.w : n4.y;
    nzw[0] = r.negz ? n4.z : n3.x; nzw[1] = r.negz ? n4.w : n3.y;
    fzw[0] = r.negz ? n3.x : n4.z; fzw[1] = r.negz ? n3.y : n4.w;
    const uint32_t metaw[2] = { n1.z, n1.w };

    uint32_t keys[8];
    uint32_t hitMask = 0;
#if defined(__CUDA_ARCH__)
    #pragma unroll
#endif
    for (int i = 0; i < 8; ++i)
    {
        const int w = i >> 2;
        float tnx = byteToFloat(nxw[w], i & 3) * sx + bx;
        float tfx = byteToFloat(fxw[w], i & 3) * sx + bx;
        float tny = byteToFloat(nyw[w], i & 3) * sy + by;
        float tfy = byteToFloat(fyw[w], i & 3) * sy + by;
        float tnz = byteToFloat(nzw[w], i & 3) * sz + bz;
        float tfz = byteToFloat(fzw[w], i & 3) * sz + bz;
        float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, r.tNear));
        float tmax = fminf(fminf(tfx, tfy), fminf(tfz, r.tBest));
        // a few ulps of slack: the planes come out of an fma on rounded inputs
        // (empty slots hold an inverted box, lo = 255 > hi = 0, and miss by themselves; should the slack
        // ever let one through, its meta byte says "leaf with no triangles")
        bool h = tmin <= tmax * 1.0000004f;
        keys[i] = h ? (((f2u(tmin) & ~7u) | (uint32_t)i) + 8u) : 0u;
        hitMask |= h ? (1u << i) : 0u;
    }

    // leaf children: test their triangles now
    uint32_t leafMask = hitMask & ~imask;
    while (leafMask)
    {
        // nearest leaf first, so the farther ones can be skipped once something is hit
        // (cornell 1080p: extend 71.4 -> 61.7 ms, connect 101.0 -> 92.8 ms per 6 steps)
        uint32_t lk = 0xffffffffu;
#if defined(__CUDA_ARCH__)
        #pragma unroll
#endif
        for (int k = 0; k < 8; ++k) { const uint32_t kk = ((leafMask >> k) & 1u) ? keys[k] : 0xffffffffu; lk = kk < lk ? kk : lk; }
        const int i = (int)((lk - 8u) & 7u);
        leafMask &= ~(1u << i);
        if (!kAnyHit && u2f((lk - 8u) & ~7u) > r.tBest) break;
        const uint32_t m = (metaw[i >> 2] >> (8 * (i & 3))) & 0xffu;
        const uint32_t count = m >> 5, first = n1.y + (m & 31u);
        for (uint32_t j = 0; j < count; ++j)
        {
            if (kCount) cnt->tris++;
            float t, u, v; uint32_t prim; V3 ng;
            if (triTest(bvh.tris + first + j, r.ro, r.rd, r.tNear, r.tBest, t, u, v, prim, ng))
            {
                if (kAnyHit) { r.found = true; return false; }
                // equal-t hits resolve to the lowest primitive id (as the oracle does)
                if (!r.found || t < r.tBest || prim < r.bestPrim)
                {
                    r.found = true;
                    r.tBest = t; r.bestPrim = prim;
                    r.hit.t = t; r.hit.u = u; r.hit.v = v; r.hit.prim = (int32_t)prim; r.hit.ng = ng;
                }
            }
        }
    }

    // internal children: push far-to-near, keep the nearest as the next node
    uint32_t intMask = hitMask & imask;
    if (intMask)
    {
#if defined(__CUDA_ARCH__)
        #pragma unroll
#endif
        for (int i = 0; i < 8; ++i) keys[i] = ((intMask >> i) & 1u) ? keys[i] : 0u;
        int n = popc32(intMask);
        while (n > 1)
        {
            const uint32_t best = max8(keys);
            const uint32_t idx = (best - 8u) & 7u;
            st.push(n1.x + (uint32_t)popc32(imask & ((1u << idx) - 1u)), best);
#if defined(__CUDA_ARCH__)
            #pragma unroll
#endif
            for (int i = 0; i < 8; ++i) keys[i] = (keys[i] == best) ? 0u : keys[i];
            --n;
        }
        const uint32_t best = max8(keys);
        if (u2f((best - 8u) & ~7u) <= r.tBest)
        {
            const uint32_t idx = (best - 8u) & 7u;
            r.cur = n1.x + (uint32_t)popc32(imask & ((1u << idx) - 1u));
            r.haveCur = true;
        }
    }
    return r.haveCur || st.sp != 0;
}

// One wide node of the fp32 form (Bvh::nodesF). Same arithmetic as travStep from the plane onwards
// (t = plane * idir + (origin - ro) * idir), so the same nodes are visited in the same order; the
// planes arrive as floats, near / far picked by address instead of by select.
template <bool kAnyHit, class Stack>
PT_D bool travStepF(const Bvh& bvh, TravRay& r, Stack& st)
{
    while (!r.haveCur)
    {
        if (st.sp == 0) return false;
        unsigned long long e = st.pop();
        uint32_t key = (uint32_t)(e >> 32);
        if (u2f((key - 8u) & ~7u) > r.tBest) continue;
        r.cur = (uint32_t)e;
        r.haveCur = true;
    }
    r.haveCur = false;

    const WideNodeF* np = bvh.nodesF + r.cur;
    const U4 hdr = ldg128(&np->childBase);
    const U4 org = ldg128(&np->ox);
    const float bx = (u2f(org.x) - r.ro.x) * r.idir.x;
    const float by = (u2f(org.y) - r.ro.y) * r.idir.y;
    const float bz = (u2f(org.z) - r.ro.z) * r.idir.z;
    const uint32_t imask = hdr.y >> 24;
    const uint32_t triBase = hdr.y & 0xffffffu;
    const uint32_t metaw[2] = { hdr.z, hdr.w };
    // plane sets: 0 lox, 1 loy, 2 loz, 3 hix, 4 hiy, 5 hiz; each 8 floats = two 128-bit loads
    const float* pl = &np->plane[0][0];
    const float* nx = pl + (r.negx ? 24 : 0);
    const float* fx = pl + (r.negx ? 0 : 24);
    const float* ny = pl + (r.negy ? 32 : 8);
    const float* fy = pl + (r.negy ? 8 : 32);
    const float* nz = pl + (r.negz ? 40 : 16);
    const float* fz = pl + (r.negz ? 16 : 40);

    uint32_t keys[8];
    uint32_t hitMask = 0;
#if defined(__CUDA_ARCH__)
    #pragma unroll
#endif
    for (int half = 0; half < 2; ++half)
    {
        const U4 anx = ldg128(nx + 4 * half), afx = ldg128(fx + 4 * half);
        const U4 any_ = ldg128(ny + 4 * half), afy = ldg128(fy + 4 * half);
        const U4 anz = ldg128(nz + 4 * half), afz = ldg128(fz + 4 * half);
        const uint32_t vnx[4] = { anx.x, anx.y, anx.z, anx.w }, vfx[4] = { afx.x, afx.y, afx.z, afx.w };
        const uint32_t vny[4] = { any_.x, any_.y, any_.z, any_.w }, vfy[4] = { afy.x, afy.y, afy.z, afy.w };
        const uint32_t vnz[4] = { anz.x, anz.y, anz.z, anz.w }, vfz[4] = { afz.x, afz.y, afz.z, afz.w };
#if defined(__CUDA_ARCH__)
        #pragma unroll
#endif
        for (int k = 0; k < 4; ++k)
        {
            const int i = 4 * half + k;
            float tnx = u2f(vnx[k]) * r.idir.x + bx;
            float tfx = u2f(vfx[k]) * r.idir.x + bx;
            float tny = u2f(vny[k]) * r.idir.y + by;
            float tfy = u2f(vfy[k]) * r.idir.y + by;
            float tnz = u2f(vnz[k]) * r.idir.z + bz;
            float tfz = u2f(vfz[k]) * r.idir.z + bz;
            float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, r.tNear));
            float tmax = fminf(fminf(tfx, tfy), fminf(tfz, r.tBest));
            bool h = tmin <= tmax * 1.0000004f;
            keys[i] = h ? (((f2u(tmin) & ~7u) | (uint32_t)i) + 8u) : 0u;
            hitMask |= h ? (1u << i) : 0u;
        }
    }

    uint32_t leafMask = hitMask & ~imask;
    while (leafMask)
    {
        uint32_t lk = 0xffffffffu;
#if defined(__CUDA_ARCH__)
        #pragma unroll
#endif
        for (int k = 0; k < 8; ++k) { const uint32_t kk = ((leafMask >> k) & 1u) ? keys[k] : 0xffffffffu; lk = kk < lk ? kk : lk; }
        const int i = (int)((lk - 8u) & 7u);
        leafMask &= ~(1u << i);
        if (!kAnyHit && u2f((lk - 8u) & ~7u) > r.tBest) break;
        const uint32_t m = (metaw[i >> 2] >> (8 * (i & 3))) & 0xffu;
        const uint32_t count = m >> 5, first = triBase + (m & 31u);
        for (uint32_t j = 0; j < count; ++j)
        {
            float t, u, v; uint32_t prim; V3 ng;
            if (triTest(bvh.tris + first + j, r.ro, r.rd, r.tNear, r.tBest, t, u, v, prim, ng))
            {
                if (kAnyHit) { r.found = true; return false; }
                if (!r.found || t < r.tBest || prim < r.bestPrim)
                {
                    r.found = true;
                    r.tBest = t; r.bestPrim = prim;
                    r.hit.t = t; r.hit.u = u; r.hit.v = v; r.hit.prim = (int32_t)prim; r.hit.ng = ng;
                }
            }
        }
    }

    uint32_t intMask = hitMask & imask;
    if (intMask)
    {
#if defined(__CUDA_ARCH__)
        #pragma unroll
#endif
        for (int i = 0; i < 8; ++i) keys[i] = ((intMask >> i) & 1u) ? keys[i] : 0u;
        int n = popc32(intMask);
        while (n > 1)
        {
            const uint32_t best = max8(keys);
            const uint32_t idx = (best - 8u) & 7u;
            st.push(hdr.x + (uint32_t)popc32(imask & ((1u << idx) - 1u)), best);
#if defined(__CUDA_ARCH__)
            #pragma unroll
#endif
            for (int i = 0; i < 8; ++i) keys[i] = (keys[i] == best) ? 0u : keys[i];
            --n;
        }
        const uint32_t best = max8(keys);
        if (u2f((best - 8u) & ~7u) <= r.tBest)
        {
            const uint32_t idx = (best - 8u) & 7u;
            r.cur = hdr.x + (uint32_t)popc32(imask & ((1u << idx) - 1u));
            r.haveCur = true;
        }
    }
    return r.haveCur || st.sp != 0;
}

// ---- postponed leaves (the persistent kernels, large incoherent trees) ------------------------
// In travStep a lane tests the triangles of a node's leaf children on the spot, while the other
// lanes of the warp wait: on the 10 M-triangle soup 2.3 of 32 lanes were inside a triangle test
// at any time, for a third of all instructions. Here leaf children go onto the stack like inner
// children (node field: kLeafEntry | count << 28 | first triangle); a lane that pops one only
// records it (TravRay::pendCount) and stops taking node steps. The kernel lets the other lanes go
// on and runs travLeafStepP for the waiting ones once enough of them wait (Aila and Laine's
// postponed leaves), so triangle tests execute with many lanes side by side. Children are pushed
// nearest-last but otherwise unsorted: popping culls by distance anyway, and the sort loop ran with
// 2.8 lanes. Results are those of travStep: the closest hit does not depend on the visiting order,
// ties go to the lowest primitive id.
constexpr uint32_t kLeafEntry = 0x80000000u;

// the triangles of the pending leaf
template <bool kAnyHit = false>
PT_D void travLeafStepP(const Bvh& bvh, TravRay& r)
{
    const uint32_t first = r.pendFirst, count = r.pendCount;
    r.pendCount = 0u;
    for (uint32_t j = 0; j < count; ++j)
    {
        float t, u, v; uint32_t prim; V3 ng;
        if (triTest(bvh.tris + first + j, r.ro, r.rd, r.tNear, r.tBest, t, u, v, prim, ng))
        {
            if (kAnyHit) { r.found = true; return; }      // occlusion query: the caller stops the ray
            if (!r.found || t < r.tBest || prim < r.bestPrim)
            {
                r.found = true;
                r.tBest = t; r.bestPrim = prim;
                r.hit.t = t; r.hit.u = u; r.hit.v = v; r.hit.prim = (int32_t)prim; r.hit.ng = ng;
            }
        }
    }
}

// one node step; returns false when the ray is finished (stack empty, nothing pending)
template <class Stack>
PT_D bool travNodeStepP(const Bvh& bvh, TravRay& r, Stack& st)
{
    while (!r.haveCur)
    {
        if (st.sp == 0) return false;
        unsigned long long e = st.pop();
        uint32_t key = (uint32_t)(e >> 32);
        if (u2f((key - 8u) & ~7u) > r.tBest) continue;
        const uint32_t id = (uint32_t)e;
        if (id & kLeafEntry) { r.pendFirst = id & 0x0fffffffu; r.pendCount = (id >> 28) & 7u; return true; }
        r.cur = id;
        r.haveCur = true;
    }
    r.haveCur = false;

    const WideNode* np = bvh.nodes + r.cur;
    const U4 n0 = ldg128(&np->ox);
    const U4 n1 = ldg128(&np->childBase);
    const U4 n2 = ldg128(&np->qlox[0]);
    const U4 n3 = ldg128(&np->qloz[0]);
    const U4 n4 = ldg128(&np->qhiy[0]);
    const uint32_t imask = n0.w >> 24;
    const float sx = u2f((n0.w & 0xffu) << 23) * r.idir.x;
    const float sy = u2f(((n0.w >> 8) & 0xffu) << 23) * r.idir.y;
    const float sz = u2f(((n0.w >> 16) & 0xffu) << 23) * r.idir.z;
    const float bx = (u2f(n0.x) - r.ro.x) * r.idir.x;
    const float by = (u2f(n0.y) - r.ro.y) * r.idir.y;
    const float bz = (u2f(n0.z) - r.ro.z) * r.idir.z;
    uint32_t nxw[2], fxw[2], nyw[2], fyw[2], nzw[2], fzw[2];
    nxw[0] = r.negx ? n3.z : n2.x; nxw[1] = r.negx ? n3.w : n2.y;
    fxw[0] = r.negx ? n2.x : n3.z; fxw[1] = r.negx ? n2.y : n3.w;
    nyw[0] = r.negy ? n4.x : n2.z; nyw[1] = r.negy ? n4.y : n2.w;
    fyw[0] = r.negy ? n2.z : n4.x; fyw[1] = r.negy ? n2.w : n4.y;
    nzw[0] = r.negz ? n4.z : n3.x; nzw[1] = r.negz ? n4.w : n3.y;
    fzw[0] = r.negz ? n3.x : n4.z; fzw[1] = r.negz ? n3.y : n4.w;
    const uint32_t metaw[2] = { n1.z, n1.w };

    uint32_t keys[8];
    uint32_t hitMask = 0;
    uint32_t nearest = 0xffffffffu;
#if defined(__CUDA_ARCH__)
    #pragma unroll
#endif
    for (int i = 0; i < 8; ++i)
    {
        const int w = i >> 2;
        float tnx = byteToFloat(nxw[w], i & 3) * sx + bx;
        float tfx = byteToFloat(fxw[w], i & 3) * sx + bx;
        float tny = byteToFloat(nyw[w], i & 3) * sy + by;
        float tfy = byteToFloat(fyw[w], i & 3) * sy + by;
        float tnz = byteToFloat(nzw[w], i & 3) * sz + bz;
        float tfz = byteToFloat(fzw[w], i & 3) * sz + bz;
        float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, r.tNear));
        float tmax = fminf(fminf(tfx, tfy), fminf(tfz, r.tBest));
        // (an empty slot holds an inverted box and misses by itself; inside a node of no extent at all the slack can
        // let it through, and it then travels as a leaf entry of zero triangles, which the leaf step skips)
        const bool h = tmin <= tmax * 1.0000004f;
        keys[i] = h ? (((f2u(tmin) & ~7u) | (uint32_t)i) + 8u) : 0xffffffffu;
        nearest = keys[i] < nearest ? keys[i] : nearest;
        hitMask |= h ? (1u << i) : 0u;
    }
    if (hitMask == 0u) return r.haveCur || st.sp != 0;

    // stack id of child i: inner node index, or a leaf entry
    auto childId = [&](int i) -> uint32_t
    {
        if ((imask >> i) & 1u) return n1.x + (uint32_t)popc32(imask & ((1u << i) - 1u));
        const uint32_t m = (metaw[i >> 2] >> (8 * (i & 3))) & 0xffu;
        return kLeafEntry | ((m >> 5) << 28) | (n1.y + (m & 31u));
    };
    // every hit child but the nearest goes onto the stack (unrolled and predicated: all lanes in step);
    // the nearest becomes the current node, or the pending leaf
    const uint32_t rest = hitMask & ~(1u << ((nearest - 8u) & 7u));
    uint32_t idNear = 0u;
#if defined(__CUDA_ARCH__)
    #pragma unroll
#endif
    for (int i = 0; i < 8; ++i)
    {
        if ((hitMask >> i) & 1u)
        {
            const uint32_t id = childId(i);
            if ((rest >> i) & 1u) st.push(id, keys[i]);
            else idNear = id;
        }
    }
    if (idNear & kLeafEntry) { r.pendFirst = idNear & 0x0fffffffu; r.pendCount = (idNear >> 28) & 7u; }
    else { r.cur = idNear; r.haveCur = true; }
    return true;
}

// kFloatNodes: take the fp32 node form when the tree has one. Only the two wavefront traversal
// kernels ask for it: in kernels that also shade, the second traversal body costs more in
// registers and code than it saves (stage A went from 75 to 93 ms with it).
template <bool kAnyHit, bool kCount, bool kFloatNodes = false, class Stack>
PT_D bool traverse(const Bvh& bvh, V3 ro, V3 rd, float tNear, float tFar, Hit& hit, Stack& st,
                   TravCounters* cnt)
{
    TravRay r;
    travBegin(bvh, r, ro, rd, tNear, tFar, st);
    if (kFloatNodes && !kCount && bvh.nodesF)
    {
        while (travStepF<kAnyHit>(bvh, r, st)) {}
    }
    else
    {
        while (travStep<kAnyHit, kCount>(bvh, r, st, cnt)) {}
    }
    hit = r.hit;
    return r.found;
}

// Nearest triangle to p among those closer than `radius` (unsigned distance as
// sdTriangle3D returns it). Returns kRcpEpsilon when there is none, like
// RtcPointQuery's initial usr.distance (path_tracer.c:603).
template <class Stack>
PT_D float pointQuery(const Bvh& bvh, V3 p, float radius, Stack& st)
{
    float best = kRcpEpsilon;
    if (bvh.nodeCount <= 0) return best;
    st.sp = 0;
    st.push(0, 0);
    while (st.sp > 0)
    {
        const uint32_t cur = (uint32_t)st.pop();
        if (cur == 0xffffffffu) continue;       // TravStack::kDropped
        const WideNode* np = bvh.nodes + cur;
        const U4 n0 = ldg128(&np->ox);
        const U4 n1 = ldg128(&np->childBase);
        const U4 n2 = ldg128(&np->qlox[0]);
        const U4 n3 = ldg128(&np->qloz[0]);
        const U4 n4 = ldg128(&np->qhiy[0]);
        const uint32_t imask = n0.w >> 24;
        const float sx = u2f((n0.w & 0xffu) << 23);
        const float sy = u2f(((n0.w >> 8) & 0xffu) << 23);
        const float sz = u2f(((n0.w >> 16) & 0xffu) << 23);
        const uint32_t lox[2] = { n2.x, n2.y }, loy[2] = { n2.z, n2.w }, loz[2] = { n3.x, n3.y };
        const uint32_t hix[2] = { n3.z, n3.w }, hiy[2] = { n4.x, n4.y }, hiz[2] = { n4.z, n4.w };
        const uint32_t metaw[2] = { n1.z, n1.w };
        for (int i = 0; i < 8; ++i)
        {
            const int w = i >> 2, sh = 8 * (i & 3);
            const uint32_t m = (metaw[w] >> sh) & 0xffu;
            if (m == 0u) continue;
            float lx = u2f(n0.x) + (float)((lox[w] >> sh) & 0xffu) * sx;
            float ly = u2f(n0.y) + (float)((loy[w] >> sh) & 0xffu) * sy;
            float lz = u2f(n0.z) + (float)((loz[w] >> sh) & 0xffu) * sz;
            float hx = u2f(n0.x) + (float)((hix[w] >> sh) & 0xffu) * sx;
            float hy = u2f(n0.y) + (float)((hiy[w] >> sh) & 0xffu) * sy;
            float hz = u2f(n0.z) + (float)((hiz[w] >> sh) & 0xffu) * sz;
            float dx = fmaxf(fmaxf(lx - p.x, p.x - hx), 0.0f);
            float dy = fmaxf(fmaxf(ly - p.y, p.y - hy), 0.0f);
            float dz = fmaxf(fmaxf(lz - p.z, p.z - hz), 0.0f);
            if (dx * dx + dy * dy + dz * dz > radius * radius) continue;
            if ((imask >> i) & 1u)
            {
                st.push(n1.x + (uint32_t)popc32(imask & ((1u << i) - 1u)), 0);
            }
            else
            {
                const uint32_t count = m >> 5, first = n1.y + (m & 31u);
                for (uint32_t j = 0; j < count; ++j)
                {
                    const PackedTri* tp = bvh.tris + first + j;
                    U4 a = ldg128(&tp->v0x), b = ldg128(&tp->e1x), c = ldg128(&tp->e2x);
                    V3 A = v3(u2f(a.x), u2f(a.y), u2f(a.z));
                    V3 B = A + v3(u2f(b.x), u2f(b.y), u2f(b.z));
                    V3 C = A + v3(u2f(c.x), u2f(c.y), u2f(c.z));
                    float d = sdTriangle3D(A, B, C, p);
                    if (d < radius && d < best)
                    {
                        best = d;
                        radius = f1_min(radius, d);
                    }
                }
            }
        }
    }
    return best;
}

} // namespace pt
