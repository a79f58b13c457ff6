// pathkernels.cu — kernels that run a whole path in one thread: Pt_TraceRayN for small
// batches and the lightmap bake (K12), where one thread owns a texel because the
// spherical-gaussian fit is sequential per texel.
#include "pt_device.h"
#include "kernels_common.cuh"

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

} // namespace pt
