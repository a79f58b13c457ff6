// wavefront.cu — the pixel loop of Pt_Trace (src/rendering/path_tracer.c:2514-2585)
// as a wavefront over a resident pool of path slots.
//
//   K5  regen     TraceFn's head: pixel filter, proj_dir, CalculateDof (or RayGenFn /
//                 explicit rays). A slot whose path ended claims the next work item of the
//                 call, so the pool stays full until the call's work runs out.
//   K3' extend    closest hit for every live path
//   K6a shadeA    Pt_TraceRay from the hit up to EstimateDirect's ray
//   K4' connect   the NEE / MIS ray (closest hit + primitive id, as the reference does)
//   K6b shadeB    rest of EstimateDirect, BSDF scatter, roulette for the next bounce
//   K6f finish    once the work is used up and few paths are left: one thread per path to its end
//   K7  (in shade) finished paths add their result to the per-pixel sums
//   K8  (in regen/shade) work claims and live counts: ballot/popc per warp, one atomic per CTA
//
// Compaction moves state, not indices: the six persistent state fields are double buffered,
// stage B writes each surviving path to the next free position of the other buffer (one atomic
// per CTA, so a CTA's survivors land contiguously) and regen appends the new paths behind them.
// Every kernel therefore runs over a dense range [0, N) with full warps, fully coalesced 16-byte
// fields (512 contiguous bytes per warp and array) and the freshly generated, coherent camera
// rays side by side. Index queues over fixed slots (an earlier version) paid a 32-byte sector
// per 16-byte field once slot order had decayed; regenerating in place instead scattered the
// camera rays among old paths and slowed the traversal kernels (profiles/README.md has the
// three measurements).
#include "pt_device.h"
#include "kernels_common.cuh"

#include <algorithm>
#include <cstring>
#include <mutex>

namespace pt {

bool g_kernelTiming = false;
thread_local bool t_workerThread = false;
int g_persistentTraversal = -1;   // -1: by scene size (see usePersistent)
constexpr int kSmallCallPaths = 131072;   // a call with no more paths than this skips the wavefront altogether
// live paths at which the tail is finished in one launch; PIMPT_FINISH=n overrides, 0 disables. Msamples/s at
// 128 K / 512 K / 2 M: cornell boxes 1080p 2454 / 2455 / 2466, 320x180 x 64 spp 1454 / 1503 / 1569, its lightmap bake
// 1231 / 1289 / 1326 (M texel-samples/s), 1 M-triangle interior 1266 / 1273 / 1277, but cornell spheres (glass: long,
// uneven paths) 740 / 739 / 676: 512 K is the largest that costs no scene anything
int g_finishThreshold = 524288;
KernelStats g_kernelStats = {};

// Refilling idle lanes pays off only when rays take very many steps: +22 % on the 10 M-triangle
// soup (1.8 M nodes, 58 nodes per ray), but -14 % on cornell boxes (19 nodes, 2-4 per ray) and
// -18 % on the 1 M-triangle interior (149 K nodes): there the refill's ballot + atomic cost more
// than the idle lanes. The switch is the BVH size.
bool usePersistent(int32_t bvhNodeCount)
{
    if (g_persistentTraversal >= 0) return g_persistentTraversal != 0;
    return bvhNodeCount >= (1 << 20);
}

// The batch entry points (Pt_IntersectN / Pt_OccludedN) carry rays in the caller's order, usually
// unrelated ones: there the persistent kernels with postponed leaves win from small trees on
// (closest / any hit, Mrays/s: cornell spheres, 5 K nodes, 4259 -> 6739 / 5222 -> 9987; 1 M-triangle soup
// 323 -> 569; only the 19-node cornell boxes lose, 16.9 -> 13.3 G). The wavefront's rays are the paths of
// neighbouring pixels and keep the higher switch above.
bool usePersistentBatch(int32_t bvhNodeCount)
{
    if (g_persistentTraversal >= 0) return g_persistentTraversal != 0;
    return bvhNodeCount >= 2048;
}

namespace {

// shading kernels are latency bound (dependent loads through hit -> vertex -> material ->
// texel); PT_SHADE_BLOCKS trades registers per thread for resident warps
#ifndef PT_SHADE_THREADS
#define PT_SHADE_THREADS 128
#endif
constexpr int kShadeThreads = PT_SHADE_THREADS;
#ifndef PT_SHADE_BLOCKS_A
#define PT_SHADE_BLOCKS_A 8
#endif
// stage B, cornell 1080p, us per full-pool launch at 4 / 5 / 6 / 7 / 8 CTAs per SM (128 / 96 / 80 / 72 / 64 registers):
// 583 / 509 / 490 / 496 / 523; stage A at 6 / 7 / 8 / 9 / 10: 402 / 378 / 364 / 414 / 424
#ifndef PT_SHADE_BLOCKS_B
#define PT_SHADE_BLOCKS_B 6
#endif

// What stage A leaves for stage B. rd2.w holds flag bits: the low two are Pend_None (shade
// without a second ray), Pend_Light / Pend_Bsdf (the connect kernel traced one) or kPendDead;
// then the surface's back-face, refractive and "shading normal differs from the interpolated
// one" bits. The surface itself travels as surfN = (N, roughness), surfA = (albedo, metallic)
// and, only when the normal map changed it, surfM = (M, 0); re-deriving it in stage B cost a
// second walk through vertex, material and texture gathers. pend = (selectPdf, distSq,
// light vertex bits, -) for a light sample, (attenuation, pdf) for a BSDF sample, (ior, ...)
// when there is no second ray (the only case that can be refractive).
constexpr uint32_t kPendDead = 3u;
constexpr uint32_t kSurfBackface = 4u, kSurfRefractive = 8u, kSurfHasM = 16u;

struct PathArrays { F4* ro; F4* rd; F4* atten; F4* lum; F4* albedo; F4* normal; };

struct PoolView
{
    PathArrays cur, next;         // persistent path state, ping-pong
    U4* hit1; F4* ro2; F4* rd2; F4* pend; U4* hit2;   // per-iteration temporaries
    F4* surfN; F4* surfA; F4* surfM;                  // stage A's surface interaction for stage B
    int32_t* connectList;
    int32_t* counters;
    unsigned long long* cursor;   // [0] next work item (may run past the total), [1] total
    int32_t capacity;
};

__device__ __forceinline__ F4 ldF4(const F4* p) { U4 a = ldg128(p); return f4(u2f(a.x), u2f(a.y), u2f(a.z), u2f(a.w)); }

// warp-aggregated counter bump: one atomic per warp
__device__ __forceinline__ void warpCount(bool pred, int32_t* counter)
{
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask && (threadIdx.x & 31) == (__ffs((int)mask) - 1)) atomicAdd(counter, __popc(mask));
}

// warp-aggregated allocation of consecutive positions: one atomic per warp (K8)
__device__ __forceinline__ int warpAlloc(bool pred, int32_t* counter)
{
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0u) return 0;
    const int lane = threadIdx.x & 31, leader = __ffs((int)mask) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(mask & ((1u << lane) - 1u));
}

// Path state in HBM: six 16-byte fields per slot. Only two of the RNG's four words ever advance
// on a path (rng_u32 / rng_f32 / rng_f2), so they ride in the spare lanes of atten and lum — the
// fields stage A rewrites anyway — and the result index rides with the denoiser accumulators.
//   ro     = origin.xyz, 0               rd     = direction.xyz, bounce | prevRefractive << 16
//   atten  = attenuation.xyz, rng.x      lum    = luminance.xyz, rng.y
//   albedo = albedo.xyz, result index    normal = normal.xyz, AOV weight
// Stage A needs the first four (plus the last two only when a path ends or scatters in a medium);
// stage B needs all but the origin, which it replaces.
__device__ __forceinline__ void loadPathHot(const PathArrays& P, int slot, Path& p)
{
    F4 a = P.ro[slot], b = P.rd[slot], c = P.atten[slot], d = P.lum[slot];
    p.ro = v3(a);
    p.rd = v3(b);
    uint32_t bits = f2u(b.w);
    p.bounce = (int32_t)(bits & 0xffffu);
    p.prevFlags = (bits >> 16) ? MatFlag_Refractive : 0u;
    p.atten = v3(c); p.rng.x = f2u(c.w);
    p.lum = v3(d); p.rng.y = f2u(d.w);
    p.rng.z = p.rng.w = 0u;
}

__device__ __forceinline__ void loadPathNoOrigin(const PathArrays& P, int slot, Path& p)
{
    F4 b = P.rd[slot], c = P.atten[slot], d = P.lum[slot];
    p.ro = v3s(0.0f);
    p.rd = v3(b);
    uint32_t bits = f2u(b.w);
    p.bounce = (int32_t)(bits & 0xffffu);
    p.prevFlags = (bits >> 16) ? MatFlag_Refractive : 0u;
    p.atten = v3(c); p.rng.x = f2u(c.w);
    p.lum = v3(d); p.rng.y = f2u(d.w);
    p.rng.z = p.rng.w = 0u;
}

__device__ __forceinline__ void loadPathCold(const PathArrays& P, int slot, Path& p, uint32_t& outIndex)
{
    F4 e = P.albedo[slot], f = P.normal[slot];
    p.albedo = v3(e); outIndex = f2u(e.w);
    p.normal = v3(f); p.weight = f.w;
}

__device__ __forceinline__ void storePath(const PathArrays& P, int slot, const Path& p, uint32_t outIndex)
{
    P.ro[slot] = f4(p.ro, 0.0f);
    uint32_t bits = (uint32_t)p.bounce | ((p.prevFlags & MatFlag_Refractive) ? 0x10000u : 0u);
    P.rd[slot] = f4(p.rd, u2f(bits));
    P.atten[slot] = f4(p.atten, u2f(p.rng.x));
    P.lum[slot] = f4(p.lum, u2f(p.rng.y));
    P.albedo[slot] = f4(p.albedo, u2f(outIndex));
    P.normal[slot] = f4(p.normal, p.weight);
}

__device__ __forceinline__ RayHit unpackHit(const DScene& S, V3 rd, U4 h)
{
    Hit hit;
    hit.t = u2f(h.x); hit.u = u2f(h.y); hit.v = u2f(h.z); hit.prim = (int32_t)h.w;
    hit.ng = v3s(0.0f);
    if (hit.prim >= 0)
    {
        V3 A = ld3(S.positions, hit.prim * 3), B = ld3(S.positions, hit.prim * 3 + 1), C = ld3(S.positions, hit.prim * 3 + 2);
        hit.ng = cross(B - A, C - A);
    }
    return finishHit(S, rd, hit);
}

// K7: a finished path adds its result
__device__ __forceinline__ void emitResult(const WorkDesc& W, const Path& p, uint32_t outIndex)
{
    V3 color, albedo, normal;
    pathFinish(p, color, albedo, normal);
    const size_t o = (size_t)outIndex * 3;
    if (W.accumulate)
    {
        atomicAdd(W.sumColor + o, color.x); atomicAdd(W.sumColor + o + 1, color.y); atomicAdd(W.sumColor + o + 2, color.z);
        if (W.sumAlbedo)
        {
            atomicAdd(W.sumAlbedo + o, albedo.x); atomicAdd(W.sumAlbedo + o + 1, albedo.y); atomicAdd(W.sumAlbedo + o + 2, albedo.z);
        }
        if (W.sumNormal)
        {
            atomicAdd(W.sumNormal + o, normal.x); atomicAdd(W.sumNormal + o + 1, normal.y); atomicAdd(W.sumNormal + o + 2, normal.z);
        }
    }
    else
    {
        W.sumColor[o] = color.x; W.sumColor[o + 1] = color.y; W.sumColor[o + 2] = color.z;
        if (W.sumAlbedo) { W.sumAlbedo[o] = albedo.x; W.sumAlbedo[o + 1] = albedo.y; W.sumAlbedo[o + 2] = albedo.z; }
        if (W.sumNormal) { W.sumNormal[o] = normal.x; W.sumNormal[o + 1] = normal.y; W.sumNormal[o + 2] = normal.z; }
    }
}

__global__ void k_wfInit(PoolView P, unsigned long long total)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0)
    {
        for (int k = 0; k < Wf_COUNT; ++k) P.counters[k] = 0;
        P.cursor[0] = 0ull;
        P.cursor[1] = total;
    }
}

// K5: new paths go behind the survivors, in work-item order (coherent camera rays)
__global__ void __launch_bounds__(128) k_wfRegen(DScene S, PoolView P, WorkDesc W)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = P.counters[Wf_NumA];
    const unsigned long long cursor = P.cursor[0], total = P.cursor[1];
    const unsigned long long left = total - cursor;
    const int room = P.capacity - n;
    const int nNew = (left < (unsigned long long)room) ? (int)left : room;
    if (i >= nNew) return;
    const int slot = n + i;
    const unsigned long long item = cursor + (unsigned long long)i;

    Path p;
    uint32_t outIndex;
    V3 ro, rd;
    Rng rng;
    if (W.kind == Work_Camera)
    {
        const unsigned long long pixels = W.pixelCount > 0 ? (unsigned long long)W.pixelCount
                                                           : (unsigned long long)W.cam.width * (unsigned long long)W.cam.height;
        const uint32_t pixel = (uint32_t)W.pixelBegin + (uint32_t)(item % pixels);
        const uint32_t sample = (uint32_t)(item / pixels);
        rng = seed_ctx(pixel, W.passBase + sample, W.seed, W.salt);
        cameraRay(W.cam, rng, (int)pixel, ro, rd);
        outIndex = pixel;
    }
    else if (W.kind == Work_Rays)
    {
        rng = seed_ctx(W.rngIds ? W.rngIds[item] : (uint32_t)item, W.passBase, W.seed, W.salt);
        F4 o = ldF4(W.ro + item), d = ldF4(W.rd + item);
        ro = v3(o); rd = v3(d);
        outIndex = (uint32_t)item;
    }
    else
    {
        rng = seed_ctx((uint32_t)item, W.passBase, W.seed, W.salt);
        rd = SampleUnitSphere(rng_f2(rng));            // RayGenFn, path_tracer.c:2607
        ro = W.origin;
        W.directions[item] = f4(rd, 0.0f);
        outIndex = (uint32_t)item;
    }
    pathInit(p, rng, ro, rd);
    pathRoulette(p);    // bounce 0: probability 1, consumes one word like the reference
    storePath(P.cur, slot, p, outIndex);
}

__global__ void k_wfAfterRegen(PoolView P)
{
    const int n = P.counters[Wf_NumA];
    const unsigned long long left = P.cursor[1] - P.cursor[0];
    const int room = P.capacity - n;
    const int nNew = (left < (unsigned long long)room) ? (int)left : room;
    P.counters[Wf_NumA] = n + nNew;
    P.cursor[0] += (unsigned long long)nNew;
    P.counters[Wf_NumNext] = 0;
    P.counters[Wf_ExtendCursor] = 0;
    P.counters[Wf_ConnectCursor] = 0;
    P.counters[Wf_NumConnect] = 0;
}

// K3'
#ifndef PT_TRAV_BLOCKS
#define PT_TRAV_BLOCKS 8
#endif
__global__ void __launch_bounds__(kTravThreads, PT_TRAV_BLOCKS) k_wfExtend(DScene S, PoolView P)
{
    PT_DECLARE_STACK(st);
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= P.counters[Wf_NumA]) return;
    F4 o = P.cur.ro[slot], d = P.cur.rd[slot];
    Hit hit;
    traverse<false, false, true>(S.bvh, v3(o), v3(d), 0.0f, kRcpEpsilon, hit, st, (TravCounters*)nullptr);
    U4 h; h.x = f2u(hit.t); h.y = f2u(hit.u); h.z = f2u(hit.v); h.w = (uint32_t)hit.prim;
    P.hit1[slot] = h;
}

// K3'/K4' persistent form: a warp keeps stepping its rays one wide node at a time and, when
// at least kRefillLanes lanes have finished, fetches the next rays for them with one atomic
// per warp. Rays take very different numbers of steps (17 of 32 lanes were active on average
// in the one-thread-per-ray kernels), so refilling keeps the SIMD lanes busy.
#ifndef PT_REFILL_LANES
#define PT_REFILL_LANES 8
#endif
constexpr int kRefillLanes = PT_REFILL_LANES;
constexpr int kPersistentBlocksPerSm = 12;
#ifndef PT_LEAF_VOTES
#define PT_LEAF_VOTES 12
#endif
constexpr int kLeafVotes = PT_LEAF_VOTES;

template <bool kConnect>
__global__ void __launch_bounds__(kTravThreads) k_wfTracePersistent(DScene S, PoolView P)
{
    PT_DECLARE_STACK(st);
    const int lane = threadIdx.x & 31;
    const uint32_t ltMask = (1u << lane) - 1u;
    int32_t* cursor = &P.counters[kConnect ? Wf_ConnectCursor : Wf_ExtendCursor];
    const int n = P.counters[kConnect ? Wf_NumConnect : Wf_NumA];
    TravRay r;
    bool busy = false, exhausted = false;
    int slot = -1;
    for (;;)
    {
        const unsigned idle = __ballot_sync(0xffffffffu, !busy);
        if (!exhausted && (__popc(idle) >= kRefillLanes))
        {
            const int leader = __ffs((int)idle) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(cursor, __popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!busy)
            {
                const int idx = base + __popc(idle & ltMask);
                if (idx < n)
                {
                    slot = kConnect ? P.connectList[idx] : idx;
                    F4 o = kConnect ? P.ro2[slot] : P.cur.ro[slot];
                    F4 d = kConnect ? P.rd2[slot] : P.cur.rd[slot];
                    travBegin(S.bvh, r, v3(o), v3(d), 0.0f, kConnect ? o.w : kRcpEpsilon, st);
                    busy = true;
                }
            }
            exhausted = base + __popc(idle) >= n;
        }
        if (__ballot_sync(0xffffffffu, busy) == 0u)
        {
            if (exhausted) break;
            continue;
        }
        // postponed leaves (traverse.h): lanes holding a popped leaf wait while the others take node steps;
        // the triangles are tested once kLeafVotes lanes wait or nobody can take a node step
        const bool waiting = busy && r.pendCount != 0u;
        const unsigned waitMask = __ballot_sync(0xffffffffu, waiting);
        const unsigned nodeMask = __ballot_sync(0xffffffffu, busy && !waiting);
        if (__popc(waitMask) >= kLeafVotes || nodeMask == 0u)
        {
            if (waiting) travLeafStepP(S.bvh, r);
            continue;
        }
        if (busy && !waiting && !travNodeStepP(S.bvh, r, st))
        {
            U4 h; h.x = f2u(r.hit.t); h.y = f2u(r.hit.u); h.z = f2u(r.hit.v); h.w = (uint32_t)r.hit.prim;
            if (kConnect) P.hit2[slot] = h; else P.hit1[slot] = h;
            busy = false;
        }
    }
}

// K6a
__global__ void __launch_bounds__(kShadeThreads, PT_SHADE_BLOCKS_A) k_wfShadeA(DScene S, PoolView P, WorkDesc W)
{
    LazyStack st;
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;     // always below the pool capacity
    // issue the state loads before the live count is known: they do not depend on it
    Path p;
    loadPathHot(P.cur, slot, p);
    p.albedo = p.normal = v3s(0.0f);    // stage A only ever adds to the AOV accumulators (medium scatter):
    p.weight = 0.0f;                    // collect the additions, fold them in if the cold fields get loaded
    const U4 h1 = P.hit1[slot];
    const bool active = slot < P.counters[Wf_NumA];
    bool ray2 = false, moved = false;
    uint32_t outIndex = 0;
    if (active)
    {
        RayHit hit = unpackHit(S, p.rd, h1);
        Pending pend;
        SurfHit surf;
        int r = pathStageA(S, p, hit, pend, surf, st);
        if (r == Stage_Done || r == Stage_Next)
        {
            // the path ends here or scattered inside the medium: both need the rest of the state
            Path cold;
            loadPathCold(P.cur, slot, cold, outIndex);
            p.albedo = cold.albedo + p.albedo; p.normal = cold.normal + p.normal; p.weight = cold.weight + p.weight;
            // scattered: the next bounce starts without a surface interaction
            if (r == Stage_Next && pathAdvance(p)) moved = true;
            else emitResult(W, p, outIndex);
            P.rd2[slot] = f4(0.0f, 0.0f, 0.0f, u2f(kPendDead));
        }
        else
        {
            ray2 = (r == Stage_Ray2);
            // stage A only changes rng, atten, lum
            P.cur.atten[slot] = f4(p.atten, u2f(p.rng.x));
            P.cur.lum[slot] = f4(p.lum, u2f(p.rng.y));
            const bool hasM = (surf.M.x != surf.N.x) | (surf.M.y != surf.N.y) | (surf.M.z != surf.N.z);
            const uint32_t bits = (uint32_t)pend.kind | (surf.type == Hit_Backface ? kSurfBackface : 0u) |
                                  ((surf.flags & MatFlag_Refractive) ? kSurfRefractive : 0u) | (hasM ? kSurfHasM : 0u);
            if (pend.kind == Pend_Light) pend.a2 = u2f((uint32_t)pend.iVertLight);
            else if (pend.kind == Pend_None) pend.a0 = surf.ior;
            P.ro2[slot] = f4(pend.ro2, pend.tFar2);
            P.rd2[slot] = f4(pend.rd2, u2f(bits));
            P.pend[slot] = f4(pend.a0, pend.a1, pend.a2, pend.a3);
            P.surfN[slot] = f4(surf.N, surf.roughness);
            P.surfA[slot] = f4(surf.albedo, surf.metallic);
            if (hasM) P.surfM[slot] = f4(surf.M, 0.0f);
        }
    }
    {
        const int dst = warpAlloc(moved, &P.counters[Wf_NumNext]);      // rare: no atomic unless a lane scattered
        if (moved) storePath(P.next, dst, p, outIndex);
    }
    int nRay2 = 0;
    const int idx = blockAlloc<kShadeThreads / 32>(ray2, &P.counters[Wf_NumConnect], nRay2);
    if (ray2) P.connectList[idx] = slot;
}

// K4'
__global__ void __launch_bounds__(kTravThreads, PT_TRAV_BLOCKS) k_wfConnect(DScene S, PoolView P)
{
    PT_DECLARE_STACK(st);
    // dense over the slots that asked for a ray (shadeA's list is ascending within each CTA's chunk)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.counters[Wf_NumConnect]) return;
    const int slot = P.connectList[i];
    F4 o = P.ro2[slot], d = P.rd2[slot];
    Hit hit;
    traverse<false, false, true>(S.bvh, v3(o), v3(d), 0.0f, o.w, hit, st, (TravCounters*)nullptr);
    U4 h; h.x = f2u(hit.t); h.y = f2u(hit.u); h.z = f2u(hit.v); h.w = (uint32_t)hit.prim;
    P.hit2[slot] = h;
}

// K6b
__global__ void __launch_bounds__(kShadeThreads, PT_SHADE_BLOCKS_B) k_wfShadeB(DScene S, PoolView P, WorkDesc W)
{
    LazyStack st;
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;     // always below the pool capacity
    // loads that do not depend on the live count go first
    Path p; uint32_t outIndex = 0;
    loadPathNoOrigin(P.cur, slot, p);           // stage B replaces the origin and never reads the old one
    loadPathCold(P.cur, slot, p, outIndex);
    const U4 h2 = P.hit2[slot];
    const F4 o2 = P.ro2[slot], d2 = P.rd2[slot], pa = P.pend[slot];
    const F4 sn = P.surfN[slot], sa = P.surfA[slot];
    const uint32_t bits = f2u(d2.w);
    const uint32_t kind = (slot < P.counters[Wf_NumA]) ? (bits & 3u) : kPendDead;
    bool alive = false;
    if (kind != kPendDead)
    {
        Pending pend;
        pend.kind = (int32_t)kind;
        pend.iVertLight = (kind == Pend_Light) ? (int32_t)f2u(pa.z) : -1;
        pend.ro2 = v3(o2); pend.tFar2 = o2.w; pend.rd2 = v3(d2);
        pend.a0 = pa.x; pend.a1 = pa.y; pend.a2 = pa.z; pend.a3 = pa.w;
        SurfHit surf;
        surf.P = pend.ro2;
        surf.N = v3(sn); surf.roughness = sn.w;
        surf.albedo = v3(sa); surf.metallic = sa.w;
        surf.M = surf.N;
        if (bits & kSurfHasM) surf.M = v3(P.surfM[slot]);
        surf.emission = v3s(0.0f); surf.occlusion = 0.0f;
        surf.ior = (kind == Pend_None) ? pa.x : 1.0f;
        surf.flags = (bits & kSurfRefractive) ? MatFlag_Refractive : 0u;
        surf.type = (bits & kSurfBackface) ? Hit_Backface : Hit_Triangle;
        RayHit hit2;
        hit2.type = Hit_Nothing; hit2.iVert = -1; hit2.t = -1.0f; hit2.flags = 0;
        hit2.normal = v3s(0.0f); hit2.w = hit2.u = hit2.v = 0.0f;
        if (kind != Pend_None) hit2 = unpackHit(S, pend.rd2, h2);
        int r = pathStageB(S, p, surf, pend, hit2, st);
        if (r == Stage_Next && pathAdvance(p)) alive = true;
        else emitResult(W, p, outIndex);
    }
    // survivors move to the other buffer, packed (K8): one atomic per CTA
    int nAlive = 0;
    const int dst = blockAlloc<kShadeThreads / 32>(alive, &P.counters[Wf_NumNext], nAlive);
    if (alive) storePath(P.next, dst, p, outIndex);
}

// end of an iteration: publish progress for the host loop
__global__ void k_wfEndIteration(PoolView P, int32_t* status)
{
    const unsigned long long cur = P.cursor[0], total = P.cursor[1];
    const unsigned long long left = cur < total ? total - cur : 0ull;
    status[0] = P.counters[Wf_NumNext];
    status[1] = (left > 0x7fffffffull) ? 0x7fffffff : (int32_t)left;
    status[2] = P.counters[Wf_NumA];          // rays the extend kernel traced
    status[3] = P.counters[Wf_NumConnect];    // rays the connect kernel traced
    P.counters[Wf_NumA] = P.counters[Wf_NumNext];
    P.counters[Wf_NumNext] = 0;
}

// The tail of a call: once the work items are used up the live paths thin out geometrically, and
// an iteration of seven launches over a few thousand paths is all launch latency. Below
// g_finishThreshold live paths one launch runs each remaining path to its end in its own thread
// (Pt_TraceRay's loop, as k_traceRays does), from the state the last iteration left.

__global__ void __launch_bounds__(kTravThreads) k_wfFinish(DScene S, PoolView P, WorkDesc W)
{
    PT_DECLARE_STACK(st);
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t nExtend = 0, nConnect = 0;
    if (slot < P.counters[Wf_NumA])
    {
        Path p; uint32_t outIndex = 0;
        loadPathHot(P.cur, slot, p);
        loadPathCold(P.cur, slot, p, outIndex);
        bool alive = true;
        while (alive)
        {
            RayHit hit = intersectLocal(S, p.ro, p.rd, 0.0f, kRcpEpsilon, st);
            ++nExtend;
            Pending pend; SurfHit surf;
            int r = pathStageA(S, p, hit, pend, surf, st);
            if (r == Stage_Done) break;
            if (r != Stage_Next)
            {
                RayHit hit2;
                hit2.type = Hit_Nothing; hit2.iVert = -1; hit2.t = -1.0f; hit2.flags = 0;
                hit2.normal = v3s(0.0f); hit2.w = hit2.u = hit2.v = 0.0f;
                if (r == Stage_Ray2)
                {
                    hit2 = intersectLocal(S, pend.ro2, pend.rd2, 0.0f, pend.tFar2, st);
                    ++nConnect;
                }
                r = pathStageB(S, p, surf, pend, hit2, st);
                if (r == Stage_Done) break;
            }
            alive = pathAdvance(p);
        }
        emitResult(W, p, outIndex);
    }
    // ray statistics: one atomic pair per warp
    for (int o = 16; o > 0; o >>= 1)
    {
        nExtend += __shfl_down_sync(0xffffffffu, nExtend, o);
        nConnect += __shfl_down_sync(0xffffffffu, nConnect, o);
    }
    if ((threadIdx.x & 31) == 0 && (nExtend | nConnect))
    {
        atomicAdd(&P.counters[Wf_FinishExtend], (int32_t)nExtend);
        atomicAdd(&P.counters[Wf_FinishConnect], (int32_t)nConnect);
    }
}

__global__ void k_wfEndFinish(PoolView P, int32_t* status)
{
    status[0] = 0;
    status[1] = 0;
    status[2] = P.counters[Wf_FinishExtend];
    status[3] = P.counters[Wf_FinishConnect];
    P.counters[Wf_NumA] = 0;
    P.counters[Wf_NumNext] = 0;
}

PoolView makeView(const PathPool& pool)
{
    PoolView v;
    for (int b = 0; b < 2; ++b)
    {
        PathArrays& a = b ? v.next : v.cur;
        a.ro = pool.ro[b]; a.rd = pool.rd[b]; a.atten = pool.atten[b]; a.lum = pool.lum[b];
        a.albedo = pool.albedo[b]; a.normal = pool.normal[b];
    }
    v.hit1 = pool.hit1; v.ro2 = pool.ro2; v.rd2 = pool.rd2; v.pend = pool.pend; v.hit2 = pool.hit2;
    v.surfN = pool.surfN; v.surfA = pool.surfA; v.surfM = pool.surfM;
    v.connectList = pool.connectList;
    v.counters = pool.counters;
    v.cursor = reinterpret_cast<unsigned long long*>(pool.counters + Wf_CursorLo);
    v.capacity = pool.capacity;
    return v;
}

} // namespace

bool poolAlloc(PathPool& pool, int32_t capacity)
{
    poolFree(pool);
    if (capacity < 1024) capacity = 1024;
    capacity = (capacity + 255) & ~255;
    const size_t cap = (size_t)capacity;
    const size_t stateBytes = cap * 16 * 20;      // 2 x 6 state fields + 8 temporaries
    const size_t listBytes = cap * sizeof(int32_t);
    const size_t counterBytes = kWfCounterBytes;
    char* base = nullptr;
    if (!PT_CUDA(cudaMalloc((void**)&base, stateBytes + listBytes + counterBytes))) return false;
    pool.all = base;
    pool.capacity = capacity;
    char* p = base;
    auto take = [&](size_t bytes) { char* r = p; p += bytes; return r; };
    for (int b = 0; b < 2; ++b)
    {
        pool.ro[b] = (F4*)take(cap * 16); pool.rd[b] = (F4*)take(cap * 16);
        pool.atten[b] = (F4*)take(cap * 16); pool.lum[b] = (F4*)take(cap * 16); pool.albedo[b] = (F4*)take(cap * 16);
        pool.normal[b] = (F4*)take(cap * 16);
    }
    pool.hit1 = (U4*)take(cap * 16); pool.ro2 = (F4*)take(cap * 16);
    pool.rd2 = (F4*)take(cap * 16); pool.pend = (F4*)take(cap * 16); pool.hit2 = (U4*)take(cap * 16);
    pool.surfN = (F4*)take(cap * 16); pool.surfA = (F4*)take(cap * 16); pool.surfM = (F4*)take(cap * 16);
    pool.connectList = (int32_t*)take(cap * sizeof(int32_t));
    pool.counters = (int32_t*)take(counterBytes);
    if (!PT_CUDA(cudaMallocHost((void**)&pool.hostCounters, 64 * sizeof(int32_t)))) { poolFree(pool); return false; }
    bool ok = PT_CUDA(cudaEventCreate(&pool.evBegin)) && PT_CUDA(cudaEventCreate(&pool.evEnd));
    for (int r = 0; ok && r < PathPool::kRing; ++r) ok = PT_CUDA(cudaEventCreateWithFlags(&pool.evDone[r], cudaEventDisableTiming));
    if (!ok) { poolFree(pool); return false; }
    return true;
}

bool poolAux(PathPool& pool, size_t bytes, size_t keepBytes, cudaStream_t stream)
{
    if (bytes <= pool.auxBytes) return true;
    const size_t cap = (bytes + bytes / 4 + (1u << 20)) & ~(size_t)((1u << 20) - 1);
    void* grown = nullptr;
    if (!PT_CUDA(cudaMalloc(&grown, cap))) return false;
    bool ok = true;
    if (pool.aux)
    {
        if (keepBytes) ok = PT_CUDA(cudaMemcpyAsync(grown, pool.aux, std::min(keepBytes, pool.auxBytes), cudaMemcpyDeviceToDevice, stream));
        ok = PT_CUDA(cudaStreamSynchronize(stream)) && ok;      // nothing in flight may still use the old block
        cudaFree(pool.aux);
    }
    pool.aux = grown; pool.auxBytes = cap;
    return ok;
}

void poolFree(PathPool& pool)
{
    if (pool.all) cudaFree(pool.all);
    if (pool.aux) cudaFree(pool.aux);
    if (pool.hostCounters) cudaFreeHost(pool.hostCounters);
    if (pool.evBegin) cudaEventDestroy(pool.evBegin);
    if (pool.evEnd) cudaEventDestroy(pool.evEnd);
    for (int r = 0; r < PathPool::kRing; ++r)
    {
        if (pool.evDone[r]) cudaEventDestroy(pool.evDone[r]);
        for (int k = 0; k < 6; ++k) if (pool.evKernel[r][k]) cudaEventDestroy(pool.evKernel[r][k]);
    }
    pool = PathPool();
}

bool runWavefront(const DScene& S, PathPool& pool, const WorkDesc& work, cudaStream_t stream, float* kernelMsOut)
{
    if (kernelMsOut) *kernelMsOut = 0.0f;
    if (work.count <= 0) return true;
    if (pool.capacity <= 0) { setError("runWavefront: path pool not allocated"); return false; }
    PoolView P = makeView(pool);
    // no more slots than work items: small calls then run small grids
    if ((long long)P.capacity > (long long)work.count) P.capacity = (int32_t)((work.count + 127) & ~127ll);
    const int cap = P.capacity;
    const unsigned grid = (unsigned)((cap + kTravThreads - 1) / kTravThreads);
    // the events live in the pool (created with it): a call creates and destroys none
    cudaEvent_t e0 = pool.evBegin, e1 = pool.evEnd;
    bool ok = e0 && e1;
    if (!ok) setError("runWavefront: pool has no events");
    int32_t* dStatus = pool.counters + Wf_Status;     // 4 ints inside the counter block
    k_wfInit<<<(cap + 255) / 256, 256, 0, stream>>>(P, (unsigned long long)work.count);
    g_counters.kernelLaunches++;
    ok = ok && PT_CUDA(cudaEventRecord(e0, stream));
    // The host runs kLag iterations ahead of the device: iteration i's status (live paths, work
    // left) is copied to a pinned ring and read kLag iterations later, so the stream never
    // drains between iterations. The few iterations launched after the last path ended find
    // empty ranges and cost a few microseconds each.
    constexpr int kRing = PathPool::kRing, kLag = 3;
    const bool persistent = usePersistent(S.bvh.nodeCount);
    const unsigned gridP = std::min<unsigned>(grid, (unsigned)(kSmCount * kPersistentBlocksPerSm));
    unsigned long long iterations = 0, consumed = 0, extendRays = 0, connectRays = 0;
    const bool timing = g_kernelTiming && !t_workerThread;
    cudaEvent_t* done = pool.evDone;
    cudaEvent_t (*ev)[6] = pool.evKernel;
    if (timing && !ev[0][0])
    {
        for (int r = 0; ok && r < kRing; ++r)
            for (int k = 0; ok && k < 6; ++k) ok = PT_CUDA(cudaEventCreate(&ev[r][k]));
    }
    bool finished = false;
    // Upper bound on the live range of the iterations still to be launched. While work items
    // remain the pool refills to capacity; once they are used up the live count only shrinks, so
    // the (lagged) status gives a bound and the drain iterations launch correspondingly fewer CTAs.
    int bound = cap;
    bool workLeft = true;
    auto consume = [&](unsigned long long it) -> bool
    {
        const int r = (int)(it % kRing);
        if (!PT_CUDA(cudaEventSynchronize(done[r]))) return false;
        const int32_t* st = pool.hostCounters + 4 * r;
        extendRays += (unsigned long long)(uint32_t)st[2];
        connectRays += (unsigned long long)(uint32_t)st[3];
        if (timing)
        {
            for (int k = 0; k < 5; ++k)
            {
                float ms = 0.0f;
                cudaEventElapsedTime(&ms, ev[r][k], ev[r][k + 1]);
                g_kernelStats.ms[k] += ms;
                if (st[2] != 0 || st[3] != 0) g_kernelStats.launches[k] += 1;
            }
            g_kernelStats.extendRays += (unsigned long long)(uint32_t)st[2];
            g_kernelStats.connectRays += (unsigned long long)(uint32_t)st[3];
        }
        if (st[1] == 0)
        {
            workLeft = false;
            bound = std::min(bound, std::max(st[0], 1));
        }
        if (st[0] == 0 && st[1] == 0) finished = true;
        return true;
    };
    while (ok && !finished)
    {
        const int r = (int)(iterations % kRing);
        const bool smallCall = iterations == 0 && work.count <= (int64_t)std::min(g_finishThreshold, kSmallCallPaths) && work.count <= (int64_t)cap;
        if ((!workLeft && bound <= g_finishThreshold) || smallCall)
        {
            // the few paths still alive (after the iterations already queued) finish in one launch;
            // a call that small to begin with generates its paths and goes straight there
            if (timing) cudaEventRecord(ev[r][0], stream);
            if (smallCall)
            {
                k_wfRegen<<<grid, 128, 0, stream>>>(S, P, work);
                k_wfAfterRegen<<<1, 1, 0, stream>>>(P);
                g_counters.kernelLaunches += 2;
                bound = (int)work.count;
            }
            if (timing) cudaEventRecord(ev[r][1], stream);
            k_wfFinish<<<(unsigned)((bound + kTravThreads - 1) / kTravThreads), kTravThreads, 0, stream>>>(S, P, work);
            if (timing) for (int k = 2; k < 6; ++k) cudaEventRecord(ev[r][k], stream);
            k_wfEndFinish<<<1, 1, 0, stream>>>(P, dStatus);
            g_counters.kernelLaunches += 2;
            ok = ok && PT_CUDA(cudaGetLastError());
            ok = ok && PT_CUDA(cudaMemcpyAsync(pool.hostCounters + 4 * r, dStatus, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
            ok = ok && PT_CUDA(cudaEventRecord(done[r], stream));
            ++iterations;
            break;
        }
        const unsigned gridLive = (unsigned)((bound + kTravThreads - 1) / kTravThreads);
        const unsigned gridRegen = workLeft ? grid : 1u;
        const unsigned gridPersist = std::min<unsigned>(gridLive, gridP);
        const unsigned gridShade = (unsigned)((bound + kShadeThreads - 1) / kShadeThreads);
        if (timing) cudaEventRecord(ev[r][0], stream);
        k_wfRegen<<<gridRegen, 128, 0, stream>>>(S, P, work);
        k_wfAfterRegen<<<1, 1, 0, stream>>>(P);
        if (timing) cudaEventRecord(ev[r][1], stream);
        if (persistent) k_wfTracePersistent<false><<<gridPersist, kTravThreads, 0, stream>>>(S, P);
        else k_wfExtend<<<gridLive, kTravThreads, 0, stream>>>(S, P);
        if (timing) cudaEventRecord(ev[r][2], stream);
        k_wfShadeA<<<gridShade, kShadeThreads, 0, stream>>>(S, P, work);
        if (timing) cudaEventRecord(ev[r][3], stream);
        if (persistent) k_wfTracePersistent<true><<<gridPersist, kTravThreads, 0, stream>>>(S, P);
        else k_wfConnect<<<gridLive, kTravThreads, 0, stream>>>(S, P);
        if (timing) cudaEventRecord(ev[r][4], stream);
        k_wfShadeB<<<gridShade, kShadeThreads, 0, stream>>>(S, P, work);
        if (timing) cudaEventRecord(ev[r][5], stream);
        k_wfEndIteration<<<1, 1, 0, stream>>>(P, dStatus);
        g_counters.kernelLaunches += 7;
        ok = ok && PT_CUDA(cudaGetLastError());     // a failed launch of any of the seven shows up here
        ok = ok && PT_CUDA(cudaMemcpyAsync(pool.hostCounters + 4 * r, dStatus, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
        ok = ok && PT_CUDA(cudaEventRecord(done[r], stream));
        ++iterations;
        std::swap(P.cur, P.next);
        if (ok && iterations - consumed > (unsigned long long)kLag) ok = consume(consumed++);
        if (iterations > 1000000ull) { setError("runWavefront: iteration limit"); ok = false; }
    }
    while (ok && consumed < iterations) ok = consume(consumed++);
    ok = ok && PT_CUDA(cudaEventRecord(e1, stream));
    ok = ok && PT_CUDA(cudaEventSynchronize(e1));
    if (ok && kernelMsOut) cudaEventElapsedTime(kernelMsOut, e0, e1);
    if (!ok) cudaStreamSynchronize(stream);      // leave nothing in flight that still uses the pool
    g_counters.iterations += iterations;
    g_counters.paths += (unsigned long long)work.count;
    g_counters.rays += extendRays + connectRays;
    return ok;
}

// ------------------------------------------------------------------------------------
// Pt_TraceRayN without allocations.
//
// Who needs it: code that may not call cudaMalloc at the point where it traces — a frame
// loop that forbids allocation after start-up so its latency stays flat, a host thread that
// holds a lock other allocating CUDA calls would contend on, or an embedding that budgets
// device memory once when the library is loaded. For those callers the wavefront's scratch
// (double-buffered path state, per-iteration ray/hit temporaries, connect list, counters) lives
// in the file-scope array below instead of a cudaMalloc'ed pool. Every GPU that loads the
// library carries the array until the process exits, hence the modest fixed size.
constexpr int kNoAllocMaxPaths = 32768;                                   // paths (= scratch slots) per call
constexpr size_t kNoAllocSlotBytes = 20 * 16 + sizeof(int32_t);       // 12 state + 8 temporary fields + connect list
constexpr size_t kNoAllocBytes = (size_t)kNoAllocMaxPaths * kNoAllocSlotBytes + kWfCounterBytes;   // + counter block

static __device__ __align__(256) unsigned char g_noAllocScratch[kNoAllocBytes];

// In host code `g_noAllocScratch` is only a handle for the runtime, not a device address;
// the address a kernel can use comes from cudaGetSymbolAddress, once per device.
static unsigned char* g_noAllocPtr[64] = {};
static std::mutex g_noAllocMutex;       // guards the first-call fill of g_noAllocPtr

int32_t noAllocMaxPaths() { return kNoAllocMaxPaths; }

bool runWavefrontNoAlloc(const DScene& S, const F4* dRo, const F4* dRd, int64_t n, uint32_t pass, uint32_t seed,
                         float* dColor3, float* dAlbedo3, float* dNormal3, cudaStream_t stream)
{
    if (n <= 0) return true;
    if (n > kNoAllocMaxPaths)
    {
        setError("Pt_TraceRayN_NoAlloc: %lld paths exceed the static scratch (at most %d per call)", (long long)n,
                 kNoAllocMaxPaths);
        return false;
    }
    int dev = 0;
    if (!PT_CUDA(cudaGetDevice(&dev))) return false;
    if (dev < 0 || dev >= 64) { setError("Pt_TraceRayN_NoAlloc: device index %d out of range", dev); return false; }
    {
        // first call on this device (it is the current one): resolve the array's device address.
        // The lock only makes the look-up safe when first calls arrive from several host threads;
        // the calls themselves still have to be serialised by the caller (one shared array).
        std::lock_guard<std::mutex> lock(g_noAllocMutex);
        if (!g_noAllocPtr[dev])
        {
            void* p = nullptr;
            if (!PT_CUDA(cudaGetSymbolAddress(&p, g_noAllocScratch))) return false;
            g_noAllocPtr[dev] = (unsigned char*)p;
        }
    }
    // carve the same layout poolAlloc uses out of the static array
    const size_t cap = (size_t)kNoAllocMaxPaths;
    unsigned char* p = g_noAllocPtr[dev];
    auto take = [&](size_t bytes) { unsigned char* r = p; p += bytes; return r; };
    PoolView P;
    for (int b = 0; b < 2; ++b)
    {
        PathArrays& a = b ? P.next : P.cur;
        a.ro = (F4*)take(cap * 16); a.rd = (F4*)take(cap * 16);
        a.atten = (F4*)take(cap * 16); a.lum = (F4*)take(cap * 16); a.albedo = (F4*)take(cap * 16);
        a.normal = (F4*)take(cap * 16);
    }
    P.hit1 = (U4*)take(cap * 16); P.ro2 = (F4*)take(cap * 16); P.rd2 = (F4*)take(cap * 16);
    P.pend = (F4*)take(cap * 16); P.hit2 = (U4*)take(cap * 16);
    P.surfN = (F4*)take(cap * 16); P.surfA = (F4*)take(cap * 16); P.surfM = (F4*)take(cap * 16);
    P.connectList = (int32_t*)take(cap * sizeof(int32_t));
    P.counters = (int32_t*)take(kWfCounterBytes);
    P.cursor = reinterpret_cast<unsigned long long*>(P.counters + Wf_CursorLo);
    P.capacity = (int32_t)((n + 127) & ~127ll);
    int32_t* dStatus = P.counters + Wf_Status;

    WorkDesc W;
    memset(&W, 0, sizeof(W));
    W.kind = Work_Rays; W.ro = dRo; W.rd = dRd; W.count = n;
    W.passBase = pass; W.seed = seed; W.salt = Salt_Rays;
    W.sumColor = dColor3; W.sumAlbedo = dAlbedo3; W.sumNormal = dNormal3; W.accumulate = 0;

    // The array keeps whatever the previous call left: k_wfInit zeroes the counter block and
    // the work cursor; state and temporaries are written before they are read.
    const int cap1 = P.capacity;
    const unsigned grid = (unsigned)((cap1 + kTravThreads - 1) / kTravThreads);
    const unsigned gridShade = (unsigned)((cap1 + kShadeThreads - 1) / kShadeThreads);
    k_wfInit<<<1, 32, 0, stream>>>(P, (unsigned long long)n);
    g_counters.kernelLaunches++;
    bool ok = PT_CUDA(cudaGetLastError());
    unsigned long long iterations = 0, rays = 0;
    while (ok)
    {
        k_wfRegen<<<grid, 128, 0, stream>>>(S, P, W);
        k_wfAfterRegen<<<1, 1, 0, stream>>>(P);
        k_wfExtend<<<grid, kTravThreads, 0, stream>>>(S, P);
        k_wfShadeA<<<gridShade, kShadeThreads, 0, stream>>>(S, P, W);
        k_wfConnect<<<grid, kTravThreads, 0, stream>>>(S, P);
        k_wfShadeB<<<gridShade, kShadeThreads, 0, stream>>>(S, P, W);
        k_wfEndIteration<<<1, 1, 0, stream>>>(P, dStatus);
        g_counters.kernelLaunches += 7;
        ++iterations;
        int32_t status[4] = { 0, 0, 0, 0 };
        ok = ok && PT_CUDA(cudaMemcpyAsync(status, dStatus, sizeof(status), cudaMemcpyDeviceToHost, stream));
        ok = ok && PT_CUDA(cudaStreamSynchronize(stream));
        if (!ok) break;
        rays += (unsigned long long)(uint32_t)status[2] + (unsigned long long)(uint32_t)status[3];
        if (status[0] == 0 && status[1] == 0) break;
        std::swap(P.cur, P.next);
        if (iterations > 1000000ull) { setError("Pt_TraceRayN_NoAlloc: iteration limit"); ok = false; }
    }
    g_counters.iterations += iterations;
    g_counters.paths += (unsigned long long)n;
    g_counters.rays += rays;
    return ok;
}

} // namespace pt
