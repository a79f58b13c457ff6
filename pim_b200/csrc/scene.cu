// scene.cu — scene-build kernels and ray-batch kernels.
//
//   K3  closestHit      Pt_Intersect over a ray batch (path_tracer.c:1466-1474)
//   K4  anyHit          RtcOccluded16 over a ray batch (path_tracer.c:524-553)
//   K0  flatten         FlattenDrawables (path_tracer.c:692-782)
//   K9  emissionPdf     CalcEmissionPdfFn / EmissionPdf (path_tracer.c:784-843)
//   K10 lightGridCells / Pairs / Bake   SetupLightGridFn (path_tracer.c:891-990): point query,
//                       16 probe rays, cells x emissives x 16 occlusion rays,
//                       then Dist1D_Bake (dist1d.c:33-73)
//   K11 distUpdate      Dist1D_Update (dist1d.c:128-163) per cell
//       brdfLutBake     BakeBrdfFn (src/math/lighting.c:31-86)
#include "pt_device.h"
#include "kernels_common.cuh"

#include <algorithm>

namespace pt {

// ---------------------------------------------------------------- ray batches

__global__ void __launch_bounds__(kTravThreads) k_closestHit(DScene S, const F4* __restrict__ ro,
                                                              const F4* __restrict__ rd, float tNear,
                                                              RayHitOut* __restrict__ out, int64_t n)
{
    PT_DECLARE_STACK(st);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    U4 o = ldg128(ro + i), d = ldg128(rd + i);
    V3 dir = v3(u2f(d.x), u2f(d.y), u2f(d.z));
    Hit th;
    traverse<false, false, true>(S.bvh, v3(u2f(o.x), u2f(o.y), u2f(o.z)), dir, tNear, u2f(d.w), th, st, (TravCounters*)nullptr);
    RayHit h = finishHit(S, dir, th);
    // PtRayHit is 48 bytes: three 128-bit stores
    U4* dst = reinterpret_cast<U4*>(out + i);
    U4 a; a.x = f2u(h.w); a.y = f2u(h.u); a.z = f2u(h.v); a.w = f2u(h.t);
    U4 b; b.x = f2u(h.normal.x); b.y = f2u(h.normal.y); b.z = f2u(h.normal.z); b.w = 0;
    U4 c; c.x = (uint32_t)h.type; c.y = (uint32_t)h.iVert; c.z = h.flags; c.w = 0;
    dst[0] = a; dst[1] = b; dst[2] = c;
}

// idle lanes at which a persistent warp fetches new rays (10 M-triangle soup, closest / any hit Mrays/s: 8 lanes
// 334 / 638, 4 lanes 345 / 661; on the final code at 9 CTAs per SM: 4 lanes 376 / 675, 3 lanes 378 / 678, 2 lanes 378 / 675)
#ifndef PT_BATCH_REFILL
#define PT_BATCH_REFILL 3
#endif
#ifndef PT_POSTPONED_LEAVES
#define PT_POSTPONED_LEAVES 1
#endif
#ifndef PT_LEAF_VOTES
#define PT_LEAF_VOTES 12
#endif
constexpr int kLeafVotes = PT_LEAF_VOTES;

// persistent form: warps refill finished lanes with the next rays of the batch (one atomic per warp)
// registers: 70 per thread when left to the compiler (43 % occupancy); capped at 64 for 8 CTAs per SM the 10 M-triangle
// soup goes from 345 to 370 Mrays/s, at 56 (9 CTAs) to 376, at 48 (10 CTAs, spills) down to 299; 72 (7 CTAs) 347
#ifndef PT_PERSIST_MIN_BLOCKS
#define PT_PERSIST_MIN_BLOCKS 9
#endif
__global__ void __launch_bounds__(kTravThreads, PT_PERSIST_MIN_BLOCKS) k_closestHitPersistent(DScene S, const F4* __restrict__ ro,
                                                                        const F4* __restrict__ rd, float tNear,
                                                                        RayHitOut* __restrict__ out, int64_t n,
                                                                        unsigned long long* cursor)
{
    PT_DECLARE_STACK(st);
    const int lane = threadIdx.x & 31;
    const uint32_t ltMask = (1u << lane) - 1u;
    TravRay r;
    bool busy = false, exhausted = false;
    long long idx = -1;
    for (;;)
    {
        const unsigned idle = __ballot_sync(0xffffffffu, !busy);
        if (!exhausted && __popc(idle) >= PT_BATCH_REFILL)
        {
            const int leader = __ffs((int)idle) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(cursor, (unsigned long long)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!busy)
            {
                idx = (long long)base + __popc(idle & ltMask);
                if (idx < n)
                {
                    U4 o = ldg128(ro + idx), d = ldg128(rd + idx);
                    travBegin(S.bvh, r, v3(u2f(o.x), u2f(o.y), u2f(o.z)), v3(u2f(d.x), u2f(d.y), u2f(d.z)), tNear, u2f(d.w), st);
                    busy = true;
                }
            }
            exhausted = (long long)base + __popc(idle) >= n;
        }
        if (__ballot_sync(0xffffffffu, busy) == 0u)
        {
            if (exhausted) break;
            continue;
        }
#if PT_POSTPONED_LEAVES
        // lanes holding a popped leaf wait; the others take a node step; once kLeafVotes lanes wait (or
        // nobody can take a node step) the waiting lanes test their triangles side by side
        const bool waiting = busy && r.pendCount != 0u;
        const unsigned waitMask = __ballot_sync(0xffffffffu, waiting);
        const unsigned nodeMask = __ballot_sync(0xffffffffu, busy && !waiting);
        if (__popc(waitMask) >= kLeafVotes || nodeMask == 0u)
        {
            if (waiting) travLeafStepP(S.bvh, r);
            continue;
        }
        if (busy && !waiting && !travNodeStepP(S.bvh, r, st))
#else
        if (busy && !travStep<false, false>(S.bvh, r, st, (TravCounters*)nullptr))
#endif
        {
            RayHit h = finishHit(S, r.rd, r.hit);
            U4* dst = reinterpret_cast<U4*>(out + idx);
            U4 a; a.x = f2u(h.w); a.y = f2u(h.u); a.z = f2u(h.v); a.w = f2u(h.t);
            U4 b; b.x = f2u(h.normal.x); b.y = f2u(h.normal.y); b.z = f2u(h.normal.z); b.w = 0;
            U4 c; c.x = (uint32_t)h.type; c.y = (uint32_t)h.iVert; c.z = h.flags; c.w = 0;
            dst[0] = a; dst[1] = b; dst[2] = c;
            busy = false;
        }
    }
}

__global__ void __launch_bounds__(kTravThreads) k_anyHit(DScene S, const F4* __restrict__ ro,
                                                          const F4* __restrict__ rd, float tNear,
                                                          uint8_t* __restrict__ out, int64_t n)
{
    PT_DECLARE_STACK(st);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    U4 o = ldg128(ro + i), d = ldg128(rd + i);
    Hit hit;
    bool occ = traverse<true, false, true>(S.bvh, v3(u2f(o.x), u2f(o.y), u2f(o.z)), v3(u2f(d.x), u2f(d.y), u2f(d.z)),
                                     tNear, u2f(d.w), hit, st, (TravCounters*)nullptr);
    out[i] = occ ? 1 : 0;
}

// any hit on large trees: persistent warps with postponed leaves; a ray stops at its first hit
__global__ void __launch_bounds__(kTravThreads) k_anyHitPersistent(DScene S, const F4* __restrict__ ro,
                                                                    const F4* __restrict__ rd, float tNear,
                                                                    uint8_t* __restrict__ out, int64_t n,
                                                                    unsigned long long* cursor)
{
    PT_DECLARE_STACK(st);
    const int lane = threadIdx.x & 31;
    const uint32_t ltMask = (1u << lane) - 1u;
    TravRay r;
    bool busy = false, exhausted = false;
    long long idx = -1;
    for (;;)
    {
        const unsigned idle = __ballot_sync(0xffffffffu, !busy);
        if (!exhausted && __popc(idle) >= PT_BATCH_REFILL)
        {
            const int leader = __ffs((int)idle) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(cursor, (unsigned long long)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!busy)
            {
                idx = (long long)base + __popc(idle & ltMask);
                if (idx < n)
                {
                    U4 o = ldg128(ro + idx), d = ldg128(rd + idx);
                    travBegin(S.bvh, r, v3(u2f(o.x), u2f(o.y), u2f(o.z)), v3(u2f(d.x), u2f(d.y), u2f(d.z)), tNear, u2f(d.w), st);
                    busy = true;
                }
            }
            exhausted = (long long)base + __popc(idle) >= n;
        }
        if (__ballot_sync(0xffffffffu, busy) == 0u)
        {
            if (exhausted) break;
            continue;
        }
        const bool waiting = busy && r.pendCount != 0u;
        const unsigned waitMask = __ballot_sync(0xffffffffu, waiting);
        const unsigned nodeMask = __ballot_sync(0xffffffffu, busy && !waiting);
        bool done = false;
        if (__popc(waitMask) >= kLeafVotes || nodeMask == 0u)
        {
            if (waiting) { travLeafStepP<true>(S.bvh, r); done = r.found; }
        }
        else if (busy && !waiting) done = !travNodeStepP(S.bvh, r, st);
        if (done)
        {
            out[idx] = r.found ? 1 : 0;
            busy = false;
        }
    }
}

__global__ void __launch_bounds__(kTravThreads) k_closestHitCount(DScene S, const F4* __restrict__ ro,
                                                                   const F4* __restrict__ rd, float tNear, int64_t n,
                                                                   unsigned long long* totals)
{
    PT_DECLARE_STACK(st);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters c; c.nodes = 0; c.tris = 0;
    if (i < n)
    {
        U4 o = ldg128(ro + i), d = ldg128(rd + i);
        Hit hit;
        traverse<false, true>(S.bvh, v3(u2f(o.x), u2f(o.y), u2f(o.z)), v3(u2f(d.x), u2f(d.y), u2f(d.z)), tNear,
                              u2f(d.w), hit, st, &c);
    }
    uint32_t nodes = c.nodes, tris = c.tris;
    for (int o = 16; o > 0; o >>= 1)
    {
        nodes += __shfl_xor_sync(0xffffffffu, nodes, o);
        tris += __shfl_xor_sync(0xffffffffu, tris, o);
    }
    if ((threadIdx.x & 31) == 0)
    {
        atomicAdd(&totals[0], (unsigned long long)nodes);
        atomicAdd(&totals[1], (unsigned long long)tris);
    }
}

static inline unsigned gridFor(int64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

bool launchIntersect(const DScene& S, const F4* dRo, const F4* dRd, float tNear, RayHitOut* dOut, int64_t n,
                     unsigned long long* dCursor, cudaStream_t stream)
{
    if (n <= 0) return true;
    if (usePersistentBatch(S.bvh.nodeCount))
    {
        // the ray cursor belongs to this call alone (two batches in flight on different streams must
        // not share one: the other batch's claims would run this one's cursor past its rays)
        if (!dCursor) { setError("launchIntersect: no cursor for the persistent kernel"); return false; }
        if (!PT_CUDA(cudaMemsetAsync(dCursor, 0, 8, stream))) return false;
        const unsigned grid = (unsigned)std::min<int64_t>(gridFor(n, kTravThreads), kSmCount * 12);
        k_closestHitPersistent<<<grid, kTravThreads, 0, stream>>>(S, dRo, dRd, tNear, dOut, n, dCursor);
    }
    else
    k_closestHit<<<gridFor(n, kTravThreads), kTravThreads, 0, stream>>>(S, dRo, dRd, tNear, dOut, n);
    g_counters.kernelLaunches++;
    g_counters.rays += (unsigned long long)n;
    return PT_CUDA(cudaGetLastError());
}

bool launchOccluded(const DScene& S, const F4* dRo, const F4* dRd, float tNear, uint8_t* dOut, int64_t n,
                    unsigned long long* dCursor, cudaStream_t stream)
{
    if (n <= 0) return true;
    if (usePersistentBatch(S.bvh.nodeCount))
    {
        if (!dCursor) { setError("launchOccluded: no cursor for the persistent kernel"); return false; }
        if (!PT_CUDA(cudaMemsetAsync(dCursor, 0, 8, stream))) return false;
        const unsigned grid = (unsigned)std::min<int64_t>(gridFor(n, kTravThreads), kSmCount * 12);
        k_anyHitPersistent<<<grid, kTravThreads, 0, stream>>>(S, dRo, dRd, tNear, dOut, n, dCursor);
    }
    else
    k_anyHit<<<gridFor(n, kTravThreads), kTravThreads, 0, stream>>>(S, dRo, dRd, tNear, dOut, n);
    g_counters.kernelLaunches++;
    g_counters.occlusionRays += (unsigned long long)n;
    return PT_CUDA(cudaGetLastError());
}

bool launchIntersectCount(const DScene& S, const F4* dRo, const F4* dRd, float tNear, int64_t n,
                          unsigned long long* dTotals, unsigned long long* nodesOut, unsigned long long* trisOut,
                          cudaStream_t stream)
{
    bool ok = PT_CUDA(cudaMemsetAsync(dTotals, 0, 16, stream));
    if (ok && n > 0)
    {
        k_closestHitCount<<<gridFor(n, kTravThreads), kTravThreads, 0, stream>>>(S, dRo, dRd, tNear, n, dTotals);
        g_counters.kernelLaunches++;
    }
    unsigned long long h[2] = { 0, 0 };
    ok = ok && PT_CUDA(cudaMemcpyAsync(h, dTotals, 16, cudaMemcpyDeviceToHost, stream));
    ok = ok && PT_CUDA(cudaStreamSynchronize(stream));
    *nodesOut = h[0]; *trisOut = h[1];
    return ok;
}

// RtcPointQuery (path_tracer.c:597-616) over a batch: pts = (x, y, z, radius); dist = distance to the
// nearest triangle within the radius, kRcpEpsilon when there is none
__global__ void __launch_bounds__(kTravThreads) k_pointQuery(DScene S, const F4* __restrict__ pts, float* __restrict__ dist, int64_t n)
{
    PT_DECLARE_STACK(st);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    U4 p = ldg128(pts + i);
    dist[i] = pointQuery(S.bvh, v3(u2f(p.x), u2f(p.y), u2f(p.z)), u2f(p.w), st);
}

bool launchPointQuery(const DScene& S, const F4* dPts, float* dDist, int64_t n, cudaStream_t stream)
{
    if (n <= 0) return true;
    k_pointQuery<<<gridFor(n, kTravThreads), kTravThreads, 0, stream>>>(S, dPts, dDist, n);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// -------------------------------------------------------------------- flatten

__global__ void __launch_bounds__(256) k_flatten(const EntityUpload* __restrict__ ents, int entityCount,
                                                 const F4* __restrict__ srcPos, const F4* __restrict__ srcNrm,
                                                 const F4* __restrict__ srcUv, F4* __restrict__ pos,
                                                 F4* __restrict__ nrm, V2* __restrict__ uv,
                                                 int32_t* __restrict__ matIds, int vertCount)
{
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= vertCount) return;
    // entity whose [vertOffset, vertOffset+length) holds v
    int lo = 0, hi = entityCount - 1;
    while (lo < hi)
    {
        int mid = (lo + hi + 1) >> 1;
        if (ents[mid].vertOffset <= v) lo = mid; else hi = mid - 1;
    }
    const EntityUpload& e = ents[lo];
    int j = v - e.vertOffset;
    U4 p = ldg128(srcPos + e.srcOffset + j), nn = ldg128(srcNrm + e.srcOffset + j), t = ldg128(srcUv + e.srcOffset + j);
    const float* M = e.M;
    // f4x4_mul_pt: (c0*x + c1*y) + (c2*z + c3), every product and sum rounded on its own
    // (the reference's CPU code has no fused multiply-add), so vertices come out bit-equal
    float px = u2f(p.x), py = u2f(p.y), pz = u2f(p.z);
    F4 wp;
    wp.x = addr(addr(mulr(M[0], px), mulr(M[4], py)), addr(mulr(M[8], pz), M[12]));
    wp.y = addr(addr(mulr(M[1], px), mulr(M[5], py)), addr(mulr(M[9], pz), M[13]));
    wp.z = addr(addr(mulr(M[2], px), mulr(M[6], py)), addr(mulr(M[10], pz), M[14]));
    wp.w = addr(addr(mulr(M[3], px), mulr(M[7], py)), addr(mulr(M[11], pz), M[15]));
    pos[v] = wp;
    // f3x3_mul_col: c0*x + (c1*y + c2*z), then f4_normalize3
    const float* I = e.IM;
    float nx = u2f(nn.x), ny = u2f(nn.y), nz = u2f(nn.z);
    V3 wn = v3(addr(mulr(I[0], nx), addr(mulr(I[3], ny), mulr(I[6], nz))),
               addr(mulr(I[1], nx), addr(mulr(I[4], ny), mulr(I[7], nz))),
               addr(mulr(I[2], nx), addr(mulr(I[5], ny), mulr(I[8], nz))));
    {
        const float len = sqrtr(dotr(wn, wn));      // f4_normalize3 = f4_divvs(x, length): IEEE here, the
        wn = v3(divr(wn.x, len), divr(wn.y, len), divr(wn.z, len));   // library is built with --use_fast_math
    }
    nrm[v] = f4(wn, 0.0f);
    // uvs are copied per whole triangle (path_tracer.c:755-766)
    int tri = j / 3;
    V2 tuv = v2(0.0f, 0.0f);
    if (tri * 3 + 3 <= e.length) tuv = v2(u2f(t.x), u2f(t.y));
    uv[v] = tuv;
    matIds[v] = lo;
}

bool launchFlatten(const EntityUpload* dEntities, int32_t entityCount, const F4* dSrcPos, const F4* dSrcNrm,
                   const F4* dSrcUv, F4* dPos, F4* dNrm, V2* dUv, int32_t* dMatIds, int32_t vertCount,
                   cudaStream_t stream)
{
    if (vertCount <= 0) return true;
    k_flatten<<<gridFor(vertCount, 256), 256, 0, stream>>>(dEntities, entityCount, dSrcPos, dSrcNrm, dSrcUv, dPos,
                                                           dNrm, dUv, dMatIds, vertCount);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// --------------------------------------------------------------- emission pdf

__global__ void __launch_bounds__(128) k_emissionPdf(DScene S, int triCount, uint32_t seed, float* __restrict__ pdfs)
{
    int tri = blockIdx.x * blockDim.x + threadIdx.x;
    if (tri >= triCount) return;
    const int iVert = tri * 3;
    const DMaterial& mat = S.materials[S.matIds[iVert]];
    if (mat.flags & MatFlag_Sky) { pdfs[tri] = 1.0f; return; }
    if (mat.rome < 0) { pdfs[tri] = 0.0f; return; }
    const DTexture tex = S.textures[mat.rome];
    if (tex.w == 1 && tex.h == 1)
    {
        // a flat texture: all 1000 probes of EmissionPdf (path_tracer.c:784-822) read the same texel
        pdfs[tri] = (tex.texels[0] >> 24) ? 1.0f : 0.0f;
        return;
    }
    const V2 UA = S.uvs[iVert], UB = S.uvs[iVert + 1], UC = S.uvs[iVert + 2];
    Rng rng = seed_prng((uint32_t)tri, 0u, seed, Salt_Emission);
    const int attempts = 1000;
    int hits = 0;
    for (int i = 0; i < attempts; ++i)
    {
        V3 wuv = SampleBaryCoord(rng_f2(rng));
        V2 uv = v2((UA.x * wuv.x + UB.x * wuv.y) + UC.x * wuv.z, (UA.y * wuv.x + UB.y * wuv.y) + UC.y * wuv.z);
        int cx = PointWrap1D(tex.w, uv.x), cy = PointWrap1D(tex.h, uv.y);
        uint32_t sample = tex.texels[tex.w * cy + cx] >> 24;
        hits += sample ? 1 : 0;
    }
    pdfs[tri] = (float)hits / (float)attempts;
}

bool launchEmissionPdf(const DScene& S, int32_t triCount, uint32_t seed, float* dPdfOut, cudaStream_t stream)
{
    if (triCount <= 0) return true;
    k_emissionPdf<<<gridFor(triCount, 128), 128, 0, stream>>>(S, triCount, seed, dPdfOut);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// ----------------------------------------------------------------- light grid

__device__ void distBake(const DScene& S, int cell)      // Dist1D_Bake, dist1d.c:33-73
{
    const int pdfLen = S.emissiveCount;
    const int cdfLen = pdfLen + 1;
    float* pdf = S.distPdf + (size_t)cell * pdfLen;
    float* cdf = S.distCdf + (size_t)cell * cdfLen;
    const float rcpLen = 1.0f / pdfLen;
    cdf[0] = 0.0f;
    for (int i = 1; i < cdfLen; ++i) cdf[i] = cdf[i - 1] + pdf[i - 1] * rcpLen;
    float integral = cdf[cdfLen - 1];
    if (integral == 0.0f)
    {
        for (int i = 1; i < cdfLen; ++i) cdf[i] = i * rcpLen;
    }
    else
    {
        float rcpIntegral = 1.0f / integral;
        for (int i = 1; i < cdfLen; ++i) cdf[i] = cdf[i] * rcpIntegral;
        for (int i = 0; i < pdfLen; ++i) pdf[i] = pdf[i] * rcpIntegral;
    }
}

// Grid_Position (grid.h:24-39) and SampleBaryCoord (sampling.h:120-128) with IEEE division / square root and
// unfused products, as the reference's CPU build rounds them: the light-grid rays start and end where the
// reference's do, to the bit (the library's --use_fast_math forms differ by an ulp, which flips grazing rays)
__device__ __forceinline__ V3 gridPositionR(const Grid& g, int index)
{
    const float metersPerCell = divr(1.0f, g.cellsPerMeter);
    const int ix = index % g.sx, iy = (index / g.sx) % g.sy, iz = index / (g.sx * g.sy);
    return v3(addr(g.lo.x, mulr((float)ix + 0.5f, metersPerCell)), addr(g.lo.y, mulr((float)iy + 0.5f, metersPerCell)),
              addr(g.lo.z, mulr((float)iz + 0.5f, metersPerCell)));
}
__device__ __forceinline__ V3 sampleBaryCoordR(V2 Xi)
{
    const float r1 = sqrtr(f1_max(Xi.x, kEpsilonSq)), r2 = Xi.y;
    const float u = mulr(r1, subr(1.0f, r2)), v = mulr(r2, r1);
    return v3(subr(1.0f, addr(u, v)), u, v);
}

// SetupLightGridFn (path_tracer.c:891-990) in three kernels. The reference walks a cell's emissive triangles in
// order with ONE random stream per cell (a hash chain: no skipping ahead), which is why a thread per cell was the
// first form — and the slow one: the 1 M-triangle interior has 1875 cells x 648 emissives x 16 rays, and 15 CTAs
// of cells ran for 317 ms at 5.6 % occupancy. Now
//   cells   a thread per cell: the surface test (point query, probe rays), and for a cell that is kept the random
//           stream stepped through all its emissives with the state at the start of each one saved (16 B per pair);
//   pairs   a thread per (cell, emissive): its 16 visibility rays from the saved state -> pdf;
//   bake    a thread per cell: Dist1D_Bake.
// Same draws, same ray set-up, same any-hit decisions as before: the pdfs are unchanged.
__global__ void __launch_bounds__(kTravThreads) k_lightGridCells(DScene S, float metersPerCell, uint32_t seed, Rng* __restrict__ states,
                                                                  uint8_t* __restrict__ keep)
{
    PT_DECLARE_STACK(st);
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= Grid_Len(S.grid)) return;
    const float radius = metersPerCell * 0.666f;
    const V3 position = Grid_Position(S.grid, cell);
    S.distLength[cell] = 0;
    S.distSum[cell] = 0;
    keep[cell] = 0;
    {
        float dist = pointQuery(S.bvh, position, radius + 0.01f * kMilli, st);
        if (dist > radius)
        {
            // far from any surface: keep the cell only if most probe rays see front faces
            float hitcount = 0.0f;
            for (int j = 0; j < 16; ++j)
            {
                V3 dir = SampleUnitSphere(Hammersley2D((uint32_t)j, 16u));
                RayHit hit = intersectLocal(S, position, dir, 0.0f, kRcpEpsilon, st);
                if (hit.type == Hit_Triangle) hitcount += 1.0f;
            }
            if (hitcount / 16.0f < 0.5f) return;
        }
    }
    const int E = S.emissiveCount;
    if (E <= 0) return;
    keep[cell] = 1;
    Rng rng = seed_prng((uint32_t)cell, 0u, seed, Salt_LightGrid);
    for (int iEmit = 0; iEmit < E; ++iEmit)
    {
        states[(size_t)cell * E + iEmit] = rng;
        for (int k = 0; k < 16; ++k) { rng_f4(rng); rng_f2(rng); }      // the draws of this emissive's 16 rays
    }
}

__global__ void __launch_bounds__(kTravThreads) k_lightGridPairs(DScene S, float metersPerCell, const Rng* __restrict__ states,
                                                                  const uint8_t* __restrict__ keep, long long pairs)
{
    PT_DECLARE_STACK(st);
    const long long pair = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= pairs) return;
    const int E = S.emissiveCount;
    const int cell = (int)(pair / E), iEmit = (int)(pair % E);
    if (!keep[cell]) return;
    const float radius = metersPerCell * 0.666f;
    const V3 position = gridPositionR(S.grid, cell);
    Rng rng = states[pair];
    const int iVert = S.emitToVert[iEmit];
    const V3 A = ld3(S.positions, iVert), B = ld3(S.positions, iVert + 1), C = ld3(S.positions, iVert + 2);
    int hits = 0;
    for (int k = 0; k < 16; ++k)
    {
        // ray setup without fused multiply-adds: same roundings as the reference's CPU code,
        // so the any-hit decisions (and with them the k/16 pdfs) come out the same
        F4 r4 = rng_f4(rng);
        V3 t = v3(mulr(addr(-1.5f, mulr(r4.x, 3.0f)), radius), mulr(addr(-1.5f, mulr(r4.y, 3.0f)), radius),
                  mulr(addr(-1.5f, mulr(r4.z, 3.0f)), radius));
        V3 ro = v3(addr(position.x, t.x), addr(position.y, t.y), addr(position.z, t.z));
        V3 wuv = sampleBaryCoordR(rng_f2(rng));
        V3 at = v3(addr(addr(mulr(A.x, wuv.x), mulr(B.x, wuv.y)), mulr(C.x, wuv.z)),
                   addr(addr(mulr(A.y, wuv.x), mulr(B.y, wuv.y)), mulr(C.y, wuv.z)),
                   addr(addr(mulr(A.z, wuv.x), mulr(B.z, wuv.y)), mulr(C.z, wuv.z)));
        V3 rd = v3(subr(at.x, ro.x), subr(at.y, ro.y), subr(at.z, ro.z));
        float dist = sqrtr(dotr(rd, rd));
        // f4_divvs under the reference's -ffast-math: one reciprocal, three products
        const float rcpDist = divr(1.0f, dist);
        rd = v3(mulr(rd.x, rcpDist), mulr(rd.y, rcpDist), mulr(rd.z, rcpDist));
        float tFar = dist - 0.01f * kMilli;
        bool visible = false;
        if (tFar > 0.0f)
        {
            Hit hit;
            visible = !traverse<true, false>(S.bvh, ro, rd, 0.0f, tFar, hit, st, (TravCounters*)nullptr);
        }
        hits += visible ? 1 : 0;
    }
    S.distPdf[pair] = (float)hits / 16.0f;
    S.distLive[pair] = 0;
}

__global__ void __launch_bounds__(128) k_lightGridBake(DScene S, const uint8_t* __restrict__ keep)
{
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= Grid_Len(S.grid) || !keep[cell]) return;
    distBake(S, cell);
    S.distLength[cell] = S.emissiveCount;
}

bool launchLightGrid(const DScene& S, float ptDistMeters, uint32_t seed, cudaStream_t stream)
{
    int cells = Grid_Len(S.grid);
    if (cells <= 0) return true;
    const long long pairs = (long long)cells * (long long)std::max(S.emissiveCount, 0);
    Rng* dStates = nullptr; uint8_t* dKeep = nullptr;
    bool ok = PT_CUDA(cudaMalloc((void**)&dStates, sizeof(Rng) * (size_t)std::max<long long>(pairs, 1))) &&
              PT_CUDA(cudaMalloc((void**)&dKeep, (size_t)cells));
    if (ok)
    {
        k_lightGridCells<<<gridFor(cells, kTravThreads), kTravThreads, 0, stream>>>(S, ptDistMeters, seed, dStates, dKeep);
        if (pairs > 0)
        {
            k_lightGridPairs<<<gridFor(pairs, kTravThreads), kTravThreads, 0, stream>>>(S, ptDistMeters, dStates, dKeep, pairs);
            k_lightGridBake<<<gridFor(cells, 128), 128, 0, stream>>>(S, dKeep);
        }
        g_counters.kernelLaunches += 3;
        g_counters.occlusionRays += (unsigned long long)pairs * 16ull;
        ok = PT_CUDA(cudaGetLastError()) && PT_CUDA(cudaStreamSynchronize(stream));
    }
    if (dStates) cudaFree(dStates);
    if (dKeep) cudaFree(dKeep);
    return ok;
}

__global__ void __launch_bounds__(128) k_distUpdate(DScene S)     // Dist1D_Update, dist1d.c:128-163
{
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= Grid_Len(S.grid)) return;
    const int pdfLen = S.distLength[cell];
    if (pdfLen <= 0) return;
    float* pdf = S.distPdf + (size_t)cell * pdfLen;
    uint32_t* live = S.distLive + (size_t)cell * pdfLen;
    uint32_t sum = 0;
    for (int i = 0; i < pdfLen; ++i) sum += live[i];
    if (sum < 30) return;
    const float scale = 1.0f / sum;
    const uint32_t prevSum = S.distSum[cell];
    S.distSum[cell] = sum;
    float alpha = 0.5f;
    if (prevSum > 0)
    {
        double ratio = (double)sum / (double)prevSum;
        alpha = f1_sat((float)ratio) * 0.9f;
        alpha = alpha * alpha;
    }
    for (int i = 0; i < pdfLen; ++i)
    {
        uint32_t ct = live[i];
        pdf[i] = f1_lerp(pdf[i], ct * scale, alpha);
        live[i] = ct >> 1;
    }
    distBake(S, cell);
}

bool launchDistUpdate(const DScene& S, cudaStream_t stream)
{
    int cells = Grid_Len(S.grid);
    if (cells <= 0 || S.emissiveCount <= 0) return true;
    k_distUpdate<<<gridFor(cells, 128), 128, 0, stream>>>(S);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------- BRDF LUT

__global__ void __launch_bounds__(128) k_brdfLut(float* __restrict__ texels, int w, int h, int numSamples, uint32_t seed)
{
    const int iTexel = blockIdx.x * blockDim.x + threadIdx.x;
    if (iTexel >= w * h) return;
    const int cx = iTexel % w, cy = iTexel / w;
    const float dx = 1.0f / numSamples;
    const float rcpWidth = 1.0f / w, rcpHeight = 1.0f / h;
    Rng rng = seed_prng((uint32_t)iTexel, 0u, seed, Salt_Lut);
    float rx = 0.0f, ry = 0.0f;
    for (int iSample = 0; iSample < numSamples; ++iSample)
    {
        V2 aa = SampleGaussPixelFilter(rng_f2(rng), 0.5f);
        const float NoV = f1_clamp((cx + aa.x) * rcpWidth, kEpsilon, 1.0f - kEpsilon);
        const float alpha = f1_clamp((cy + aa.y) * rcpHeight, kMinAlpha, 1.0f);
        const V3 V = SphericalToCartesian(NoV, rng_f32(rng) * kTau);
        V3 H = SampleGGXMicrofacet(rng_f2(rng), alpha);
        V3 L = reflect(-V, H);
        float NoL = L.z, NoH = H.z, HoV = dot(H, V);
        float pdf = GGXPdf(NoH, HoV, alpha);
        if ((NoL > kEpsilon) && (pdf > kEpsilon))
        {
            float D = D_GTR(NoH, alpha) / pdf;
            float G = V_SmithCorrelated(NoL, NoV, alpha);
            float Fc = F_Dielectric(HoV, 1.000293f, 1.52f);
            float DG_NoL_dx = D * G * NoL * dx;
            rx += DG_NoL_dx * Fc;
            ry += DG_NoL_dx;
        }
    }
    texels[2 * iTexel] = rx;
    texels[2 * iTexel + 1] = ry;
}

bool launchBrdfLutBake(float* dLut, int32_t w, int32_t h, int32_t samples, uint32_t seed, cudaStream_t stream)
{
    k_brdfLut<<<gridFor((int64_t)w * h, 128), 128, 0, stream>>>(dLut, w, h, samples, seed);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// -------------------------------------------------------------------- images

__global__ void __launch_bounds__(256) k_resolve(float* __restrict__ image, const float* __restrict__ sum, int64_t n,
                                                 float prevCount, float addCount)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    image[i] = (image[i] * prevCount + sum[i]) / (prevCount + addCount);
}
__global__ void __launch_bounds__(256) k_lerp(float* __restrict__ image, const float* __restrict__ sample, int64_t n,
                                              float weight)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float a = image[i];
    image[i] = a + (sample[i] - a) * weight;      // f3_lerpvs, path_tracer.c:2550-2552
}

// element-wise helpers of the in-process multi-GPU mode: partial images and light-learning counters of the
// replicas are brought to the primary device with peer copies and folded in there
__global__ void __launch_bounds__(256) k_accumulate(float* __restrict__ dst, const float* __restrict__ src, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] += src[i];
}
__global__ void __launch_bounds__(256) k_u32Sub(uint32_t* __restrict__ dst, const uint32_t* __restrict__ a, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = a[i] - dst[i];          // dst held the values before the call: now what the call added
}
__global__ void __launch_bounds__(256) k_u32Add(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] += src[i];
}
bool launchAccumulate(float* dDst, const float* dSrc, int64_t n, cudaStream_t stream)
{
    if (n <= 0) return true;
    k_accumulate<<<gridFor(n, 256), 256, 0, stream>>>(dDst, dSrc, n);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}
bool launchU32Sub(uint32_t* dDst, const uint32_t* dA, int64_t n, cudaStream_t stream)
{
    if (n <= 0) return true;
    k_u32Sub<<<gridFor(n, 256), 256, 0, stream>>>(dDst, dA, n);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}
bool launchU32Add(uint32_t* dDst, const uint32_t* dSrc, int64_t n, cudaStream_t stream)
{
    if (n <= 0) return true;
    k_u32Add<<<gridFor(n, 256), 256, 0, stream>>>(dDst, dSrc, n);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// read-only sweep of a buffer that fits in L2, for the L2 roofline of the cache-resident configs: every thread
// streams 16-byte loads through the whole buffer `reps` times; the xor keeps the loads alive
__global__ void __launch_bounds__(256) k_l2Read(const uint4* __restrict__ buf, size_t n16, int reps, unsigned* sink)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned acc = 0;
    for (int r = 0; r < reps; ++r)
    {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride)
        {
            uint4 v;
            asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}
bool launchL2Read(const void* dBuf, size_t bytes, int reps, unsigned* dSink, cudaStream_t stream)
{
    k_l2Read<<<kSmCount * 8, 256, 0, stream>>>((const uint4*)dBuf, bytes / 16, reps, dSink);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

bool launchResolve(float* dImage, const float* dSum, int64_t n, float prevCount, float addCount, cudaStream_t stream)
{
    if (n <= 0) return true;
    k_resolve<<<gridFor(n, 256), 256, 0, stream>>>(dImage, dSum, n, prevCount, addCount);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}
bool launchLerp(float* dImage, const float* dSample, int64_t n, float weight, cudaStream_t stream)
{
    if (n <= 0) return true;
    k_lerp<<<gridFor(n, 256), 256, 0, stream>>>(dImage, dSample, n, weight);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

} // namespace pt
