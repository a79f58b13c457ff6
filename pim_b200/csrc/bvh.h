// bvh.h — the ray-query backend that replaces Embree in the hot path
// (reference call sites: src/rendering/path_tracer.c:484 rtcIntersect1,
// :548 rtcOccluded16, :614 rtcPointQuery, :618-690 scene build).
//
// Layout in HBM
//   WideNode  80 B, 16 B aligned, five 128-bit loads: an 8-wide node whose child
//             boxes are quantised to 8 bits per plane on a per-node power-of-two
//             grid (origin + q * 2^e). Children that are leaves address up to 3
//             triangles each, relative to the node's triangle base.
//   PackedTri 48 B, three 128-bit loads: v0, e1 = v1-v0, e2 = v2-v0 and the
//             primitive id, stored in leaf order so a leaf's triangles are adjacent.
// One thread owns one ray. The traversal stack lives in shared memory, laid out
// [depth][thread] so concurrent pushes and pops never bank-conflict; entries
// deeper than the shared part spill to a short per-thread array.
#pragma once

#include "pt_core.h"

namespace pt {

struct alignas(16) WideNode
{
    float ox, oy, oz;
    uint32_t exyzImask;     // ex | ey<<8 | ez<<16 (biased exponents) | internal-mask<<24
    uint32_t childBase;     // index of the first internal child
    uint32_t triBase;       // index of the first leaf triangle
    uint32_t meta[2];       // per child byte: 0 empty, 0xff internal, else (count<<5)|offset
    uint32_t qlox[2], qloy[2];
    uint32_t qloz[2], qhix[2];
    uint32_t qhiy[2], qhiz[2];
};
static_assert(sizeof(WideNode) == 80, "80-byte node");

struct alignas(16) PackedTri
{
    float v0x, v0y, v0z; uint32_t prim;
    float e1x, e1y, e1z; uint32_t pad0;
    float e2x, e2y, e2z; uint32_t pad1;
};
static_assert(sizeof(PackedTri) == 48, "48-byte triangle");

// The same wide node with its plane bytes already converted: plane = q * 2^e in fp32, still relative
// to the node origin, so the traversal computes the very same t = plane * idir + (origin - ro) * idir
// as it does from the bytes (scaling by a power of two commutes with the rounding) and both forms
// visit the same nodes and return the same hits. 224 bytes, fourteen 128-bit loads, no byte-to-float
// conversion in the traversal loop. Built next to the quantised nodes for trees small enough that
// node bytes do not matter (kFloatNodeMaxNodes).
struct alignas(16) WideNodeF
{
    uint32_t childBase;
    uint32_t triBaseImask;  // triangle base | internal-mask << 24
    uint32_t meta[2];
    float ox, oy, oz, pad;
    float plane[6][8];      // lox, loy, loz, hix, hiy, hiz of the eight children, relative to the origin
};
static_assert(sizeof(WideNodeF) == 224, "224-byte node");
constexpr int kFloatNodeMaxNodes = 262144;     // 59 MB at most: under half of the 126 MB L2

struct Bvh
{
    const WideNode* nodes;
    const PackedTri* tris;
    int32_t nodeCount;
    int32_t triCount;
    const WideNodeF* nodesF;    // null above kFloatNodeMaxNodes nodes (or 2^24 triangles)
};

struct Hit
{
    float t, u, v;
    int32_t prim;   // -1 = miss
    V3 ng;          // unnormalised (v1-v0)x(v2-v0), only valid on a hit
};

#ifndef PT_LEAF_MAX_TRIS
#define PT_LEAF_MAX_TRIS 3
#endif
constexpr int kLeafMaxTris = PT_LEAF_MAX_TRIS;
constexpr uint32_t kMetaInternal = 0xffu;

// ---------------------------------------------------------------- build math

PT_HD uint64_t expandBits21(uint32_t v)
{
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// 63-bit Morton code of a centroid inside [lo, lo + 1/invExt]
PT_HD uint64_t morton63(V3 c, V3 lo, V3 invExt)
{
    float fx = f1_clamp((c.x - lo.x) * invExt.x, 0.0f, 1.0f) * 2097151.0f;
    float fy = f1_clamp((c.y - lo.y) * invExt.y, 0.0f, 1.0f) * 2097151.0f;
    float fz = f1_clamp((c.z - lo.z) * invExt.z, 0.0f, 1.0f) * 2097151.0f;
    return (expandBits21((uint32_t)fx) << 2) | (expandBits21((uint32_t)fy) << 1) | expandBits21((uint32_t)fz);
}

PT_HD int clz64(uint64_t x)
{
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}

// Karras 2012: common-prefix length of keys i and j, index as tie-break
PT_HD int lbvhDelta(const uint64_t* keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + clz64((uint64_t)(uint32_t)(i ^ j) << 32);
    return clz64(a ^ b);
}

// internal node i of an n-leaf radix tree: covered range and split position
PT_HD void lbvhRange(const uint64_t* keys, int n, int i, int& first, int& last, int& split)
{
    int d = (lbvhDelta(keys, n, i, i + 1) - lbvhDelta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lbvhDelta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvhDelta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
    {
        if (lbvhDelta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    }
    int j = i + l * d;
    int dnode = lbvhDelta(keys, n, i, j);
    int s = 0;
    int t = l;
    do
    {
        t = (t + 1) >> 1;
        if (lbvhDelta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    split = i + s * d + (d < 0 ? -1 : 0);
    first = i < j ? i : j;
    last = i < j ? j : i;
}

// Binary tree in SoA form. Node ids: [0, n-2] internal, [n-1, 2n-2] leaves
// (leaf k covers sorted primitive k). lo.w / hi.w carry the child ids.
struct BinTree
{
    F4* lo;          // xyz = box min, w = left child id (bits)
    F4* hi;          // xyz = box max, w = right child id (bits)
    int32_t* first;  // first sorted primitive covered
    int32_t* last;   // last sorted primitive covered
    int32_t n;       // leaf count
};

PT_HD float boxArea(F4 lo, F4 hi)
{
    float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}
PT_HD uint32_t f2u(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
PT_HD float u2f(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

PT_HD bool binIsLeafCluster(const BinTree& T, int node)
{
    return (T.last[node] - T.first[node] + 1) <= kLeafMaxTris;
}

// Greedy collapse of the binary subtree under `root` into at most 8 children:
// repeatedly open the internal child with the largest surface area.
PT_HD int collapseChildren(const BinTree& T, int root, int kids[8])
{
    int n = 0;
    if (binIsLeafCluster(T, root))
    {
        kids[n++] = root;
        return n;
    }
    kids[n++] = (int)f2u(T.lo[root].w);
    kids[n++] = (int)f2u(T.hi[root].w);
    while (n < 8)
    {
        int best = -1;
        float bestArea = -1.0f;
        for (int i = 0; i < n; ++i)
        {
            int k = kids[i];
            if (binIsLeafCluster(T, k)) continue;
            float a = boxArea(T.lo[k], T.hi[k]);
            if (a > bestArea) { bestArea = a; best = i; }
        }
        if (best < 0) break;
        int k = kids[best];
        kids[best] = (int)f2u(T.lo[k].w);
        kids[n++] = (int)f2u(T.hi[k].w);
    }
    return n;
}

PT_HD uint32_t quantExp(float extent)
{
    // smallest power of two s with 255*s >= extent; returned as a biased exponent byte
    int e = 1;
    if (extent > 0.0f)
    {
        e = (int)((f2u(extent / 255.0f) >> 23) & 0xff);
        if (e < 1) e = 1;
        while (e < 254 && 255.0f * u2f((uint32_t)e << 23) < extent) ++e;
    }
    return (uint32_t)e;
}

// [a, b] on the grid org + q * scale: the largest q at or below a, the smallest at or above b (0 .. 255)
PT_HD void quantRange(float a, float b, float org, float scale, int& ql, int& qh)
{
    float q = floorf((a - org) / scale); ql = (int)f1_clamp(q, 0.0f, 255.0f);
    while (ql > 0 && org + (float)ql * scale > a) --ql;
    q = ceilf((b - org) / scale); qh = (int)f1_clamp(q, 0.0f, 255.0f);
    while (qh < 255 && org + (float)qh * scale < b) ++qh;
}

PT_HD void setByte(uint32_t w[2], int i, uint32_t v)
{
    w[i >> 2] |= (v & 0xffu) << (8 * (i & 3));
}

// Writes one wide node. childKind[i]: 0 internal, >0 leaf with that many triangles.
PT_HD void encodeWideNode(WideNode& out, const BinTree& T, const int kids[8], int n,
                          uint32_t childBase, uint32_t triBase)
{
    F4 lo = f4(3.0e38f, 3.0e38f, 3.0e38f, 0), hi = f4(-3.0e38f, -3.0e38f, -3.0e38f, 0);
    for (int i = 0; i < n; ++i)
    {
        F4 a = T.lo[kids[i]], b = T.hi[kids[i]];
        lo.x = f1_min(lo.x, a.x); lo.y = f1_min(lo.y, a.y); lo.z = f1_min(lo.z, a.z);
        hi.x = f1_max(hi.x, b.x); hi.y = f1_max(hi.y, b.y); hi.z = f1_max(hi.z, b.z);
    }
    uint32_t ex = quantExp(hi.x - lo.x), ey = quantExp(hi.y - lo.y), ez = quantExp(hi.z - lo.z);
    float sx = u2f(ex << 23), sy = u2f(ey << 23), sz = u2f(ez << 23);
    WideNode w;
    w.ox = lo.x; w.oy = lo.y; w.oz = lo.z;
    w.childBase = childBase; w.triBase = triBase;
    w.meta[0] = w.meta[1] = 0;
    w.qlox[0] = w.qlox[1] = w.qloy[0] = w.qloy[1] = w.qloz[0] = w.qloz[1] = 0;
    w.qhix[0] = w.qhix[1] = w.qhiy[0] = w.qhiy[1] = w.qhiz[0] = w.qhiz[1] = 0;
    uint32_t imask = 0;
    uint32_t triOff = 0;
    for (int i = 0; i < 8; ++i)
    {
        if (i >= n)
        {
            // empty slot: inverted box
            setByte(w.qlox, i, 255); setByte(w.qloy, i, 255); setByte(w.qloz, i, 255);
            continue;
        }
        int k = kids[i];
        F4 a = T.lo[k], b = T.hi[k];
        float q;
        int ql, qh;
        #define PT_QUANT(axis, org, scale, qlw, qhw)                                         \
            q = floorf((a.axis - org) / scale); ql = (int)f1_clamp(q, 0.0f, 255.0f);           \
            while (ql > 0 && org + (float)ql * scale > a.axis) --ql;                           \
            q = ceilf((b.axis - org) / scale); qh = (int)f1_clamp(q, 0.0f, 255.0f);            \
            while (qh < 255 && org + (float)qh * scale < b.axis) ++qh;                         \
            setByte(qlw, i, (uint32_t)ql); setByte(qhw, i, (uint32_t)qh);
        PT_QUANT(x, w.ox, sx, w.qlox, w.qhix)
        PT_QUANT(y, w.oy, sy, w.qloy, w.qhiy)
        PT_QUANT(z, w.oz, sz, w.qloz, w.qhiz)
        #undef PT_QUANT
        if (binIsLeafCluster(T, k))
        {
            uint32_t cnt = (uint32_t)(T.last[k] - T.first[k] + 1);
            setByte(w.meta, i, (cnt << 5) | triOff);
            triOff += cnt;
        }
        else
        {
            setByte(w.meta, i, kMetaInternal);
            imask |= 1u << i;
        }
    }
    w.exyzImask = ex | (ey << 8) | (ez << 16) | (imask << 24);
    out = w;
}

// ------------------------------------------------------------------ loading

#if defined(__CUDA_ARCH__)
PT_D U4 ldg128(const void* p)
{
    uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    U4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
}
PT_D int popc32(uint32_t x) { return __popc(x); }
PT_D int ffs32(uint32_t x) { return __ffs((int)x) - 1; }
#else
PT_D U4 ldg128(const void* p) { return *reinterpret_cast<const U4*>(p); }
PT_D int popc32(uint32_t x) { return __builtin_popcount(x); }
PT_D int ffs32(uint32_t x) { return __builtin_ffs((int)x) - 1; }
#endif

// Unfused arithmetic for the triangle test: the oracle's CPU code rounds every
// multiply and add separately, so the kernel does too (no fma contraction) and both
// sides take the same side of every edge / tFar comparison.
#if defined(__CUDA_ARCH__)
PT_D float mulr(float a, float b) { return __fmul_rn(a, b); }
PT_D float addr(float a, float b) { return __fadd_rn(a, b); }
PT_D float subr(float a, float b) { return __fsub_rn(a, b); }
PT_D float divr(float a, float b) { return __fdiv_rn(a, b); }      // IEEE even under --use_fast_math
PT_D float sqrtr(float a) { return __fsqrt_rn(a); }
#else
PT_D float mulr(float a, float b) { return a * b; }
PT_D float addr(float a, float b) { return a + b; }
PT_D float subr(float a, float b) { return a - b; }
PT_D float divr(float a, float b) { return a / b; }
PT_D float sqrtr(float a) { return sqrtf(a); }
#endif
PT_D float dotr(V3 a, V3 b) { return addr(addr(mulr(a.x, b.x), mulr(a.y, b.y)), mulr(a.z, b.z)); }
PT_D V3 crossr(V3 a, V3 b)
{
    return v3(subr(mulr(a.y, b.z), mulr(a.z, b.y)), subr(mulr(a.z, b.x), mulr(a.x, b.z)),
              subr(mulr(a.x, b.y), mulr(a.y, b.x)));
}

// two-sided Moller-Trumbore on a packed triangle; same arithmetic, in the same order,
// as oracle/cpu_bvh.c tri_test
PT_D bool triTest(const PackedTri* tp, V3 ro, V3 rd, float tNear, float tFar,
                  float& t, float& u, float& v, uint32_t& prim, V3& ng)
{
    U4 a = ldg128(&tp->v0x), b = ldg128(&tp->e1x), c = ldg128(&tp->e2x);
    V3 v0 = v3(u2f(a.x), u2f(a.y), u2f(a.z));
    V3 e1 = v3(u2f(b.x), u2f(b.y), u2f(b.z));
    V3 e2 = v3(u2f(c.x), u2f(c.y), u2f(c.z));
    V3 p = crossr(rd, e2);
    float det = dotr(e1, p);
    if (det == 0.0f) return false;
    float inv = divr(1.0f, det);
    V3 tv = v3(subr(ro.x, v0.x), subr(ro.y, v0.y), subr(ro.z, v0.z));
    u = mulr(dotr(tv, p), inv);
    if (!(u >= 0.0f && u <= 1.0f)) return false;
    V3 q = crossr(tv, e1);
    v = mulr(dotr(rd, q), inv);
    if (!(v >= 0.0f && addr(u, v) <= 1.0f)) return false;
    t = mulr(dotr(e2, q), inv);
    if (!(t >= tNear && t <= tFar)) return false;
    prim = a.w;
    ng = crossr(e1, e2);
    return true;
}

} // namespace pt
