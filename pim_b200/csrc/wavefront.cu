// wavefront.cu — the pixel loop of Pt_Trace (src/rendering/path_tracer.c:2514-2585)
// as a wavefront over a resident pool of path slots.
//
//   K5  regen     TraceFn's head: pixel filter, proj_dir, CalculateDof (or RayGenFn /
//                 explicit rays) into free slots; keeps the pool full until the work
//                 items of the call run out
//   K3' extend    closest hit for every live path
//   K6a shadeA    Pt_TraceRay from the hit up to EstimateDirect's ray
//   K4' connect   the NEE / MIS ray (closest hit + primitive id, as the reference does)
//   K6b shadeB    rest of EstimateDirect, BSDF scatter, roulette for the next bounce
//   K7  (in shade) finished paths add their result to the per-pixel sums
//   K8  (in shade) survivors are compacted into the next queue with warp ballot/popc
//                 and one atomic per warp
//
// The pool (capacity ~ a few hundred thousand slots, ~200 B each) is sized so the
// whole path state stays in the 126 MB L2 between kernels.
#include "pt_device.h"
#include "kernels_common.cuh"

#include <algorithm>

namespace pt {

bool g_kernelTiming = false;
KernelStats g_kernelStats = {};

namespace {

struct PoolView
{
    U4* rng; F4* ro; F4* rd; F4* atten; F4* lum; F4* albedo; F4* normal;
    U4* hit1; F4* ro2; F4* rd2; F4* pend; U4* hit2;
    int32_t* freeList;
    int32_t* counters;
    unsigned long long* cursor;   // [0] next work item, [1] total work items
    int32_t capacity;
};

__device__ __forceinline__ F4 ldF4(const F4* p) { U4 a = ldg128(p); return f4(u2f(a.x), u2f(a.y), u2f(a.z), u2f(a.w)); }
__device__ __forceinline__ F4 ldF4v(const F4* p) { return *p; }

// warp-aggregated append: one atomic per warp (K8)
__device__ __forceinline__ void warpAppend(bool pred, int32_t value, int32_t* queue, int32_t* counter)
{
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0u) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs((int)mask) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) queue[base + __popc(mask & ((1u << lane) - 1u))] = value;
}

__device__ __forceinline__ void loadPath(const PoolView& P, int slot, Path& p, uint32_t& outIndex)
{
    U4 r = P.rng[slot];
    p.rng.x = r.x; p.rng.y = r.y; p.rng.z = r.z; p.rng.w = r.w;
    F4 a = P.ro[slot], b = P.rd[slot], c = P.atten[slot], d = P.lum[slot], e = P.albedo[slot], f = P.normal[slot];
    p.ro = v3(a); p.weight = a.w;
    p.rd = v3(b);
    uint32_t bits = f2u(b.w);
    p.bounce = (int32_t)(bits & 0xffffu);
    p.prevFlags = (bits >> 16) ? MatFlag_Refractive : 0u;
    p.atten = v3(c); outIndex = f2u(c.w);
    p.lum = v3(d);
    p.albedo = v3(e);
    p.normal = v3(f);
}

__device__ __forceinline__ void storePath(const PoolView& P, int slot, const Path& p, uint32_t outIndex)
{
    U4 r; r.x = p.rng.x; r.y = p.rng.y; r.z = p.rng.z; r.w = p.rng.w;
    P.rng[slot] = r;
    P.ro[slot] = f4(p.ro, p.weight);
    uint32_t bits = (uint32_t)p.bounce | ((p.prevFlags & MatFlag_Refractive) ? 0x10000u : 0u);
    P.rd[slot] = f4(p.rd, u2f(bits));
    P.atten[slot] = f4(p.atten, u2f(outIndex));
    P.lum[slot] = f4(p.lum, 0.0f);
    P.albedo[slot] = f4(p.albedo, 0.0f);
    P.normal[slot] = f4(p.normal, 0.0f);
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
    if (i < P.capacity) P.freeList[i] = P.capacity - 1 - i;
    if (i == 0)
    {
        for (int k = 0; k < Wf_COUNT; ++k) P.counters[k] = 0;
        P.counters[Wf_NumFree] = P.capacity;
        P.cursor[0] = 0ull;
        P.cursor[1] = total;
    }
}

// K5
__global__ void __launch_bounds__(128) k_wfRegen(DScene S, PoolView P, WorkDesc W, int32_t* queueA)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int nFree = P.counters[Wf_NumFree];
    const unsigned long long cursor = P.cursor[0], total = P.cursor[1];
    const unsigned long long left = total - cursor;
    const int nNew = (left < (unsigned long long)nFree) ? (int)left : nFree;
    if (i >= nNew) return;
    const int slot = P.freeList[nFree - 1 - i];
    const unsigned long long item = cursor + (unsigned long long)i;

    Path p;
    uint32_t outIndex;
    V3 ro, rd;
    Rng rng;
    if (W.kind == Work_Camera)
    {
        const unsigned long long pixels = (unsigned long long)W.cam.width * (unsigned long long)W.cam.height;
        const uint32_t pixel = (uint32_t)(item % pixels);
        const uint32_t sample = (uint32_t)(item / pixels);
        rng = seed_ctx(pixel, W.passBase + sample, W.seed, W.salt);
        cameraRay(W.cam, rng, (int)pixel, ro, rd);
        outIndex = pixel;
    }
    else if (W.kind == Work_Rays)
    {
        rng = seed_ctx((uint32_t)item, W.passBase, W.seed, W.salt);
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
    storePath(P, slot, p, outIndex);
    queueA[P.counters[Wf_NumA] + i] = slot;
}

__global__ void k_wfAfterRegen(PoolView P)
{
    const int nFree = P.counters[Wf_NumFree];
    const unsigned long long left = P.cursor[1] - P.cursor[0];
    const int nNew = (left < (unsigned long long)nFree) ? (int)left : nFree;
    P.counters[Wf_NumA] += nNew;
    P.counters[Wf_NumFree] = nFree - nNew;
    P.cursor[0] += (unsigned long long)nNew;
    P.counters[Wf_NumB] = 0;
    P.counters[Wf_NumNext] = 0;
}

// K3'
__global__ void __launch_bounds__(kTravThreads) k_wfExtend(DScene S, PoolView P, const int32_t* __restrict__ queueA)
{
    PT_DECLARE_STACK(st);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.counters[Wf_NumA]) return;
    const int slot = queueA[i];
    F4 o = P.ro[slot], d = P.rd[slot];
    Hit hit;
    traverse<false, false>(S.bvh, v3(o), v3(d), 0.0f, kRcpEpsilon, hit, st, (TravCounters*)nullptr);
    U4 h; h.x = f2u(hit.t); h.y = f2u(hit.u); h.z = f2u(hit.v); h.w = (uint32_t)hit.prim;
    P.hit1[slot] = h;
}

// K6a
__global__ void __launch_bounds__(kTravThreads) k_wfShadeA(DScene S, PoolView P, WorkDesc W,
                                                            const int32_t* __restrict__ queueA, int32_t* queueB,
                                                            int32_t* queueNext)
{
    PT_DECLARE_STACK(st);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < P.counters[Wf_NumA];
    int slot = -1;
    bool toB = false, toNext = false, dead = false, ray2 = false;
    if (active)
    {
        slot = queueA[i];
        Path p; uint32_t outIndex;
        loadPath(P, slot, p, outIndex);
        RayHit hit = unpackHit(S, p.rd, P.hit1[slot]);
        Pending pend;
        int r = pathStageA(S, p, hit, pend, st);
        if (r == Stage_Next)
        {
            if (pathAdvance(p)) { toNext = true; storePath(P, slot, p, outIndex); }
            else r = Stage_Done;
        }
        if (r == Stage_Done)
        {
            emitResult(W, p, outIndex);
            dead = true;
        }
        else if (r == Stage_Ray2 || r == Stage_NoRay2)
        {
            toB = true;
            ray2 = (r == Stage_Ray2);
            // stage A only changes rng, atten, lum
            U4 rr; rr.x = p.rng.x; rr.y = p.rng.y; rr.z = p.rng.z; rr.w = p.rng.w;
            P.rng[slot] = rr;
            P.atten[slot] = f4(p.atten, u2f(outIndex));
            P.lum[slot] = f4(p.lum, 0.0f);
            P.ro2[slot] = f4(pend.ro2, pend.tFar2);
            P.rd2[slot] = f4(pend.rd2, u2f((uint32_t)pend.kind | ((uint32_t)(pend.iVertLight + 1) << 2)));
            P.pend[slot] = f4(pend.a0, pend.a1, pend.a2, pend.a3);
        }
    }
    warpAppend(toB, slot, queueB, &P.counters[Wf_NumB]);
    warpAppend(toNext, slot, queueNext, &P.counters[Wf_NumNext]);
    warpAppend(dead, slot, P.freeList, &P.counters[Wf_NumFree]);
    const unsigned rmask = __ballot_sync(0xffffffffu, ray2);
    if ((threadIdx.x & 31) == 0 && rmask)
    {
        atomicAdd(&P.counters[Wf_Rays], __popc(rmask));
        atomicAdd(&P.counters[Wf_Paths], __popc(rmask));
    }
}

// K4'
__global__ void __launch_bounds__(kTravThreads) k_wfConnect(DScene S, PoolView P, const int32_t* __restrict__ queueB)
{
    PT_DECLARE_STACK(st);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.counters[Wf_NumB]) return;
    const int slot = queueB[i];
    F4 d = P.rd2[slot];
    if ((f2u(d.w) & 3u) == (uint32_t)Pend_None) return;
    F4 o = P.ro2[slot];
    Hit hit;
    traverse<false, false>(S.bvh, v3(o), v3(d), 0.0f, o.w, hit, st, (TravCounters*)nullptr);
    U4 h; h.x = f2u(hit.t); h.y = f2u(hit.u); h.z = f2u(hit.v); h.w = (uint32_t)hit.prim;
    P.hit2[slot] = h;
}

// K6b
__global__ void __launch_bounds__(kTravThreads) k_wfShadeB(DScene S, PoolView P, WorkDesc W,
                                                            const int32_t* __restrict__ queueB, int32_t* queueNext)
{
    PT_DECLARE_STACK(st);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < P.counters[Wf_NumB];
    int slot = -1;
    bool toNext = false, dead = false;
    if (active)
    {
        slot = queueB[i];
        Path p; uint32_t outIndex;
        loadPath(P, slot, p, outIndex);
        RayHit hit = unpackHit(S, p.rd, P.hit1[slot]);
        Pending pend;
        F4 o2 = P.ro2[slot], d2 = P.rd2[slot], pa = P.pend[slot];
        const uint32_t bits = f2u(d2.w);
        pend.kind = (int32_t)(bits & 3u);
        pend.iVertLight = (int32_t)(bits >> 2) - 1;
        pend.ro2 = v3(o2); pend.tFar2 = o2.w; pend.rd2 = v3(d2);
        pend.a0 = pa.x; pend.a1 = pa.y; pend.a2 = pa.z; pend.a3 = pa.w;
        RayHit hit2;
        hit2.type = Hit_Nothing; hit2.iVert = -1; hit2.t = -1.0f; hit2.flags = 0;
        hit2.normal = v3s(0.0f); hit2.w = hit2.u = hit2.v = 0.0f;
        if (pend.kind != Pend_None) hit2 = unpackHit(S, pend.rd2, P.hit2[slot]);
        int r = pathStageB(S, p, hit, pend, hit2, st);
        if (r == Stage_Next && pathAdvance(p))
        {
            toNext = true;
            storePath(P, slot, p, outIndex);
        }
        else
        {
            emitResult(W, p, outIndex);
            dead = true;
        }
    }
    warpAppend(toNext, slot, queueNext, &P.counters[Wf_NumNext]);
    warpAppend(dead, slot, P.freeList, &P.counters[Wf_NumFree]);
}

// end of an iteration: the next queue becomes the current one; publish progress
__global__ void k_wfEndIteration(PoolView P, int32_t* status)
{
    const int n = P.counters[Wf_NumNext];
    P.counters[Wf_Rays] += P.counters[Wf_NumA];
    P.counters[Wf_NumA] = n;
    P.counters[Wf_NumNext] = 0;
    P.counters[Wf_NumB] = 0;
    const unsigned long long left = P.cursor[1] - P.cursor[0];
    status[0] = n;
    status[1] = (left > 0x7fffffffull) ? 0x7fffffff : (int32_t)left;
    status[2] = P.counters[Wf_Rays];
    status[3] = P.counters[Wf_Paths];      // connect (NEE / MIS) rays so far
}

PoolView makeView(const PathPool& pool)
{
    PoolView v;
    v.rng = pool.rng; v.ro = pool.ro; v.rd = pool.rd; v.atten = pool.atten; v.lum = pool.lum;
    v.albedo = pool.albedo; v.normal = pool.normal; v.hit1 = pool.hit1; v.ro2 = pool.ro2; v.rd2 = pool.rd2;
    v.pend = pool.pend; v.hit2 = pool.hit2;
    v.freeList = pool.freeList;
    v.counters = pool.counters;
    v.cursor = reinterpret_cast<unsigned long long*>(pool.counters + Wf_COUNT);
    v.capacity = pool.capacity;
    return v;
}

} // namespace

bool poolAlloc(PathPool& pool, int32_t capacity)
{
    poolFree(pool);
    if (capacity < 1024) capacity = 1024;
    const size_t cap = (size_t)capacity;
    const size_t stateBytes = cap * 16 * 12;
    const size_t queueBytes = cap * sizeof(int32_t) * 4;
    const size_t counterBytes = 256;
    char* base = nullptr;
    if (!PT_CUDA(cudaMalloc((void**)&base, stateBytes + queueBytes + counterBytes))) return false;
    pool.all = base;
    pool.capacity = capacity;
    char* p = base;
    auto take = [&](size_t bytes) { char* r = p; p += bytes; return r; };
    pool.rng = (U4*)take(cap * 16); pool.ro = (F4*)take(cap * 16); pool.rd = (F4*)take(cap * 16);
    pool.atten = (F4*)take(cap * 16); pool.lum = (F4*)take(cap * 16); pool.albedo = (F4*)take(cap * 16);
    pool.normal = (F4*)take(cap * 16); pool.hit1 = (U4*)take(cap * 16); pool.ro2 = (F4*)take(cap * 16);
    pool.rd2 = (F4*)take(cap * 16); pool.pend = (F4*)take(cap * 16); pool.hit2 = (U4*)take(cap * 16);
    pool.queueA = (int32_t*)take(cap * 4); pool.queueB = (int32_t*)take(cap * 4);
    pool.queueNext = (int32_t*)take(cap * 4); pool.freeList = (int32_t*)take(cap * 4);
    pool.counters = (int32_t*)take(counterBytes);
    if (!PT_CUDA(cudaMallocHost((void**)&pool.hostCounters, 64 * sizeof(int32_t)))) { poolFree(pool); return false; }
    return true;
}

void poolFree(PathPool& pool)
{
    if (pool.all) cudaFree(pool.all);
    if (pool.hostCounters) cudaFreeHost(pool.hostCounters);
    pool = PathPool();
}

bool runWavefront(const DScene& S, PathPool& pool, const WorkDesc& work, cudaStream_t stream, float* kernelMsOut)
{
    if (kernelMsOut) *kernelMsOut = 0.0f;
    if (work.count <= 0) return true;
    if (pool.capacity <= 0) { setError("runWavefront: path pool not allocated"); return false; }
    PoolView P = makeView(pool);
    const int cap = pool.capacity;
    const unsigned gridCap = (unsigned)((cap + kTravThreads - 1) / kTravThreads);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    bool ok = PT_CUDA(cudaEventCreate(&e0)) && PT_CUDA(cudaEventCreate(&e1));
    int32_t* dStatus = pool.counters + Wf_COUNT + 8;     // 3 ints inside the counter block
    k_wfInit<<<(cap + 255) / 256, 256, 0, stream>>>(P, (unsigned long long)work.count);
    g_counters.kernelLaunches++;
    ok = ok && PT_CUDA(cudaEventRecord(e0, stream));
    int32_t* q[2] = { pool.queueA, pool.queueNext };
    int parity = 0;
    // An upper bound on live paths per iteration is the pool capacity; grids are sized
    // for it and threads past the device-side counts exit at once.
    unsigned long long iterations = 0;
    int32_t lastRays = 0, lastConnect = 0;
    const bool timing = g_kernelTiming;
    cudaEvent_t ev[6] = { nullptr, nullptr, nullptr, nullptr, nullptr, nullptr };
    if (timing) for (int k = 0; ok && k < 6; ++k) ok = PT_CUDA(cudaEventCreate(&ev[k]));
    while (ok)
    {
        int32_t* queueA = q[parity];
        int32_t* queueNext = q[parity ^ 1];
        if (timing) cudaEventRecord(ev[0], stream);
        k_wfRegen<<<gridCap, 128, 0, stream>>>(S, P, work, queueA);
        k_wfAfterRegen<<<1, 1, 0, stream>>>(P);
        if (timing) cudaEventRecord(ev[1], stream);
        k_wfExtend<<<gridCap, kTravThreads, 0, stream>>>(S, P, queueA);
        if (timing) cudaEventRecord(ev[2], stream);
        k_wfShadeA<<<gridCap, kTravThreads, 0, stream>>>(S, P, work, queueA, pool.queueB, queueNext);
        if (timing) cudaEventRecord(ev[3], stream);
        k_wfConnect<<<gridCap, kTravThreads, 0, stream>>>(S, P, pool.queueB);
        if (timing) cudaEventRecord(ev[4], stream);
        k_wfShadeB<<<gridCap, kTravThreads, 0, stream>>>(S, P, work, pool.queueB, queueNext);
        if (timing) cudaEventRecord(ev[5], stream);
        k_wfEndIteration<<<1, 1, 0, stream>>>(P, dStatus);
        g_counters.kernelLaunches += 7;
        ++iterations;
        ok = ok && PT_CUDA(cudaMemcpyAsync(pool.hostCounters, dStatus, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
        ok = ok && PT_CUDA(cudaStreamSynchronize(stream));
        if (!ok) break;
        if (timing)
        {
            for (int k = 0; k < 5; ++k)
            {
                float ms = 0.0f;
                cudaEventElapsedTime(&ms, ev[k], ev[k + 1]);
                g_kernelStats.ms[k] += ms;
                g_kernelStats.launches[k] += 1;
            }
            g_kernelStats.extendRays += (unsigned long long)(uint32_t)(pool.hostCounters[2] - lastRays) -
                                        (unsigned long long)(uint32_t)(pool.hostCounters[3] - lastConnect);
            g_kernelStats.connectRays += (unsigned long long)(uint32_t)(pool.hostCounters[3] - lastConnect);
        }
        lastRays = pool.hostCounters[2];
        lastConnect = pool.hostCounters[3];
        if (pool.hostCounters[0] == 0 && pool.hostCounters[1] == 0) break;
        parity ^= 1;
        if (iterations > 100000ull) { setError("runWavefront: iteration limit"); ok = false; }
    }
    for (int k = 0; k < 6; ++k) if (ev[k]) cudaEventDestroy(ev[k]);
    ok = ok && PT_CUDA(cudaEventRecord(e1, stream));
    ok = ok && PT_CUDA(cudaEventSynchronize(e1));
    if (ok && kernelMsOut) cudaEventElapsedTime(kernelMsOut, e0, e1);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    g_counters.iterations += iterations;
    g_counters.paths += (unsigned long long)work.count;
    g_counters.rays += (unsigned long long)(uint32_t)lastRays;
    return ok;
}

// ------------------------------------------------------- one thread per path

__global__ void __launch_bounds__(kTravThreads) k_traceRays(DScene S, const F4* __restrict__ ro,
                                                             const F4* __restrict__ rd, int64_t n, uint32_t pass,
                                                             uint32_t seed, float* __restrict__ out9)
{
    PT_DECLARE_STACK(st);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Rng rng = seed_ctx((uint32_t)i, pass, seed, Salt_Rays);
    V3 c, a, nr;
    traceRay(S, rng, v3(ldF4(ro + i)), v3(ldF4(rd + i)), c, a, nr, st, (uint32_t*)nullptr);
    float* o = out9 + 9 * i;
    o[0] = c.x; o[1] = c.y; o[2] = c.z; o[3] = a.x; o[4] = a.y; o[5] = a.z; o[6] = nr.x; o[7] = nr.y; o[8] = nr.z;
}

bool launchTraceRays(const DScene& S, const F4* dRo, const F4* dRd, int64_t n, uint32_t pass, uint32_t seed,
                     float* dOut9, cudaStream_t stream)
{
    if (n <= 0) return true;
    k_traceRays<<<(unsigned)((n + kTravThreads - 1) / kTravThreads), kTravThreads, 0, stream>>>(S, dRo, dRd, n, pass,
                                                                                                 seed, dOut9);
    g_counters.kernelLaunches++;
    g_counters.paths += (unsigned long long)n;
    return PT_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------ lightmap bake
// K12: BakeFn (src/rendering/lightmap.c:1125-1201). The SG fit is a sequential,
// non-linear running fit per texel, so one thread owns a texel and applies its
// spp samples in order with the five lobes in registers.
__global__ void __launch_bounds__(kTravThreads) k_lightmapBake(DScene S, LightmapView lm, float metersPerTexel,
                                                                float timeSlice, int spp, uint32_t pass, uint32_t seed)
{
    PT_DECLARE_STACK(st);
    const int iTexel = blockIdx.x * blockDim.x + threadIdx.x;
    if (iTexel >= lm.texelCount) return;
    float sampleCount = lm.sampleCounts[iTexel];
    if (sampleCount == 0.0f) return;
    const uint32_t item = (uint32_t)(lm.texelBase + iTexel);
    Rng rng = seed_prng(item, pass, seed, Salt_Bake);       // BakeFn's Prng_Get() stream
    Rng ctx = seed_ctx(item, pass, seed, Salt_Bake);        // Pt_TraceRay's context stream
    if (rng_f32(rng) > timeSlice) return;

    const V3 N = normalize(v3(lm.normal[3 * iTexel], lm.normal[3 * iTexel + 1], lm.normal[3 * iTexel + 2]));
    const V3 P = v3(lm.position[3 * iTexel], lm.position[3 * iTexel + 1], lm.position[3 * iTexel + 2]) + N * kMilli;
    const Tbn tbn = NormalToTBN(N);
    // kGiAxii, src/rendering/lightmap.h:14-21
    const F4 kAxii[5] = {
        f4(0.0f, 0.0f, 1.0f, 4.999773f), f4(0.577350f, 0.577350f, 0.577350f, 4.999773f),
        f4(-0.577350f, 0.577350f, 0.577350f, 4.999773f), f4(0.577350f, -0.577350f, 0.577350f, 4.999773f),
        f4(-0.577350f, -0.577350f, 0.577350f, 4.999773f) };
    F4 probes[5], axii[5];
    for (int i = 0; i < 5; ++i)
    {
        probes[i] = lm.probes[i][iTexel];
        V3 ax = TbnToWorld(tbn, v3(kAxii[i]));
        axii[i] = f4(ax, kAxii[i].w);
    }
    for (int i = 0; i < spp; ++i)
    {
        V3 Lts = SampleUnitHemisphere(rng_f2(rng));
        V3 rd = TbnToWorld(tbn, Lts);
        float dt = (rng_f32(rng) - 0.5f) * metersPerTexel;
        float db = (rng_f32(rng) - 0.5f) * metersPerTexel;
        V3 ro = (P + tbn.c0 * dt) + tbn.c1 * db;
        V3 c, a, nr;
        traceRay(S, ctx, ro, rd, c, a, nr, st, (uint32_t*)nullptr);
        float weight = 1.0f / sampleCount;
        sampleCount += 1.0f;
        SG_Accumulate(weight, rd, c, axii, probes, 5);
    }
    for (int i = 0; i < 5; ++i) lm.probes[i][iTexel] = probes[i];
    lm.sampleCounts[iTexel] = sampleCount;
}

bool launchLightmapBake(const DScene& S, const LightmapView& lm, float metersPerTexel, float timeSlice, int32_t spp,
                        uint32_t pass, uint32_t seed, cudaStream_t stream)
{
    if (lm.texelCount <= 0) return true;
    k_lightmapBake<<<(unsigned)((lm.texelCount + kTravThreads - 1) / kTravThreads), kTravThreads, 0, stream>>>(
        S, lm, metersPerTexel, timeSlice, spp < 1 ? 1 : spp, pass, seed);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

} // namespace pt
