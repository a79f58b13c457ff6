// pathkernels.cu — kernels that run a whole path in one thread: Pt_TraceRayN for small
// batches and the lightmap bake (K12), where one thread owns a texel because the
// spherical-gaussian fit is sequential per texel.
#include "pt_device.h"
#include "kernels_common.cuh"

#include <algorithm>
#include <cstring>

namespace pt {

namespace {
__device__ __forceinline__ F4 ldF4(const F4* p) { U4 a = ldg128(p); return f4(u2f(a.x), u2f(a.y), u2f(a.z), u2f(a.w)); }
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

// ---------------------------------------------------------------- sparse page transfer
// LmPack_BakeDesc moves only the texels a chart covers (sampleCounts != 0), one 112-byte LmTexelRecord each
// (five lobes, position, normal, sample count, texel index), so that any range of them is one copy each way.

__global__ void __launch_bounds__(256) k_lmScatter(LightmapView lm, const LmTexelRecord* __restrict__ rec, int count)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    const LmTexelRecord r = rec[j];
    const int t = r.idx;
    for (int k = 0; k < 5; ++k) lm.probes[k][t] = r.probes[k];
    float* p = const_cast<float*>(lm.position) + 3 * (size_t)t;
    float* n = const_cast<float*>(lm.normal) + 3 * (size_t)t;
    p[0] = r.pos[0]; p[1] = r.pos[1]; p[2] = r.pos[2];
    n[0] = r.nrm[0]; n[1] = r.nrm[1]; n[2] = r.nrm[2];
    lm.sampleCounts[t] = r.cnt;
}

__global__ void __launch_bounds__(256) k_lmGather(LightmapView lm, LmTexelRecord* __restrict__ rec, int count)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    const int t = rec[j].idx;
    for (int k = 0; k < 5; ++k) rec[j].probes[k] = lm.probes[k][t];
    rec[j].cnt = lm.sampleCounts[t];
}

bool launchLightmapScatter(const LightmapView& lm, const LmTexelRecord* dRecords, int32_t count, cudaStream_t stream)
{
    if (count <= 0) return true;
    k_lmScatter<<<(count + 255) / 256, 256, 0, stream>>>(lm, dRecords, count);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

bool launchLightmapGather(const LightmapView& lm, LmTexelRecord* dRecords, int32_t count, cudaStream_t stream)
{
    if (count <= 0) return true;
    k_lmGather<<<(count + 255) / 256, 256, 0, stream>>>(lm, dRecords, count);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------- wavefront bake
// The same BakeFn, arranged for the GPU: (1) pick the texels that bake this pass (covered,
// time-slice draw) and compact them, (2) one thread per picked texel writes the hemisphere
// rays of all its samples (BakeFn's own RNG stream, so the rays are those of the sequential
// kernel), (3) the rays of all texels and samples go through the path wavefront at once,
// (4) one thread per texel folds its samples, in order, into the five lobes. Only the
// integrator's RNG stream differs from k_lightmapBake: it is seeded per ray instead of running
// on from the texel's previous sample, which is what lets the samples trace in parallel.

__global__ void __launch_bounds__(256) k_bakeSelect(LightmapView lm, float timeSlice, uint32_t pass, uint32_t seed,
                                                    int32_t* active, int32_t* counter)
{
    const int iTexel = blockIdx.x * blockDim.x + threadIdx.x;
    bool pick = false;
    if (iTexel < lm.texelCount && lm.sampleCounts[iTexel] != 0.0f)
    {
        Rng rng = seed_prng((uint32_t)(lm.texelBase + iTexel), pass, seed, Salt_Bake);
        pick = !(rng_f32(rng) > timeSlice);
    }
    int total = 0;
    const int idx = blockAlloc<8>(pick, counter, total);
    if (pick) active[idx] = iTexel;
}

__global__ void __launch_bounds__(128) k_bakeRays(LightmapView lm, const int32_t* __restrict__ active, int first,
                                                  int count, float metersPerTexel, int spp, uint32_t pass,
                                                  uint32_t seed, F4* __restrict__ ro, F4* __restrict__ rd,
                                                  uint32_t* __restrict__ ids)
{
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= count) return;
    const int iTexel = active[first + a];
    Rng rng = seed_prng((uint32_t)(lm.texelBase + iTexel), pass, seed, Salt_Bake);
    rng_f32(rng);       // the time-slice draw
    const V3 N = normalize(v3(lm.normal[3 * iTexel], lm.normal[3 * iTexel + 1], lm.normal[3 * iTexel + 2]));
    const V3 P = v3(lm.position[3 * iTexel], lm.position[3 * iTexel + 1], lm.position[3 * iTexel + 2]) + N * kMilli;
    const Tbn tbn = NormalToTBN(N);
    for (int i = 0; i < spp; ++i)
    {
        V3 Lts = SampleUnitHemisphere(rng_f2(rng));
        V3 dir = TbnToWorld(tbn, Lts);
        float dt = (rng_f32(rng) - 0.5f) * metersPerTexel;
        float db = (rng_f32(rng) - 0.5f) * metersPerTexel;
        V3 org = (P + tbn.c0 * dt) + tbn.c1 * db;
        ro[(size_t)a * spp + i] = f4(org, 0.0f);
        rd[(size_t)a * spp + i] = f4(dir, 0.0f);
        // the integrator's seed follows the texel and sample, not the (unordered) compaction slot
        ids[(size_t)a * spp + i] = (uint32_t)(lm.texelBase + iTexel) * (uint32_t)spp + (uint32_t)i;
    }
}

__global__ void __launch_bounds__(128) k_bakeFit(LightmapView lm, const int32_t* __restrict__ active, int first,
                                                 int count, int spp, const F4* __restrict__ rd,
                                                 const float* __restrict__ color3)
{
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= count) return;
    const int iTexel = active[first + a];
    const V3 N = normalize(v3(lm.normal[3 * iTexel], lm.normal[3 * iTexel + 1], lm.normal[3 * iTexel + 2]));
    const Tbn tbn = NormalToTBN(N);
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
    float sampleCount = lm.sampleCounts[iTexel];
    for (int i = 0; i < spp; ++i)
    {
        const size_t j = (size_t)a * spp + i;
        const V3 dir = v3(ldF4(rd + j));
        const V3 c = v3(color3[3 * j], color3[3 * j + 1], color3[3 * j + 2]);
        const float weight = 1.0f / sampleCount;
        sampleCount += 1.0f;
        SG_Accumulate(weight, dir, c, axii, probes, 5);
    }
    for (int i = 0; i < 5; ++i) lm.probes[i][iTexel] = probes[i];
    lm.sampleCounts[iTexel] = sampleCount;
}

bool launchLightmapBakeWavefront(const DScene& S, PathPool& pool, const LightmapView& lm, float metersPerTexel,
                                 float timeSlice, int32_t spp, uint32_t pass, uint32_t seed, cudaStream_t stream)
{
    if (lm.texelCount <= 0) return true;
    if (spp < 1) spp = 1;
    // scratch comes from the pool's growable block: [picked texels][counter][ray origins][directions]
    // [colours][rng ids]; a steady-state pass allocates nothing
    auto align = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t activeBytes = align(sizeof(int32_t) * (size_t)lm.texelCount), headBytes = activeBytes + 256;
    bool ok = poolAux(pool, headBytes, 0, stream);
    int32_t* dActive = (int32_t*)pool.aux;
    int32_t* dCount = (int32_t*)((char*)pool.aux + activeBytes);
    ok = ok && PT_CUDA(cudaMemsetAsync(dCount, 0, sizeof(int32_t), stream));
    int32_t nActive = 0;
    if (ok)
    {
        k_bakeSelect<<<(lm.texelCount + 255) / 256, 256, 0, stream>>>(lm, timeSlice, pass, seed, dActive, dCount);
        g_counters.kernelLaunches++;
        ok = PT_CUDA(cudaMemcpyAsync(&nActive, dCount, sizeof(int32_t), cudaMemcpyDeviceToHost, stream)) &&
             PT_CUDA(cudaStreamSynchronize(stream));
    }
    // texels per chunk: at most ~8 M rays (76 B each) in flight
    const int chunkTexels = std::max(1, (8 << 20) / spp);
    const size_t chunkRays = (size_t)std::min(chunkTexels, std::max(nActive, 1)) * (size_t)spp;
    const size_t roBytes = align(sizeof(F4) * chunkRays), colBytes = align(sizeof(float) * 3 * chunkRays),
                 idBytes = align(sizeof(uint32_t) * chunkRays);
    ok = ok && poolAux(pool, headBytes + 2 * roBytes + colBytes + idBytes, headBytes, stream);
    dActive = (int32_t*)pool.aux;       // the block may have moved (the picked texels moved with it)
    char* q = (char*)pool.aux + headBytes;
    F4* dRo = (F4*)q; q += roBytes;
    F4* dRd = (F4*)q; q += roBytes;
    float* dColor = (float*)q; q += colBytes;
    uint32_t* dIds = (uint32_t*)q;
    for (int first = 0; ok && first < nActive; first += chunkTexels)
    {
        const int count = std::min(chunkTexels, nActive - first);
        k_bakeRays<<<(count + 127) / 128, 128, 0, stream>>>(lm, dActive, first, count, metersPerTexel, spp, pass, seed, dRo, dRd, dIds);
        g_counters.kernelLaunches++;
        WorkDesc W;
        memset(&W, 0, sizeof(W));
        W.kind = Work_Rays; W.ro = dRo; W.rd = dRd; W.rngIds = dIds; W.count = (int64_t)count * spp;
        W.passBase = pass;
        W.seed = seed; W.salt = Salt_Bake;
        W.sumColor = dColor; W.accumulate = 0;
        float ms = 0.0f;
        ok = runWavefront(S, pool, W, stream, &ms);
        if (ok)
        {
            k_bakeFit<<<(count + 127) / 128, 128, 0, stream>>>(lm, dActive, first, count, spp, dRd, dColor);
            g_counters.kernelLaunches++;
            ok = PT_CUDA(cudaGetLastError());
        }
    }
    ok = PT_CUDA(cudaStreamSynchronize(stream)) && ok;
    return ok;
}

} // namespace pt
