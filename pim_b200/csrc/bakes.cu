// bakes.cu — the callers either side of the path tracer:
//   Cubemap_Bake      src/rendering/cubemap.c:138-187   one jittered ray per face texel through the
//                                                       path wavefront, lerped into the face
//   Cubemap_Convolve  src/rendering/cubemap.c:189-303   GGX prefilter of every mip texel
//   AmbCube_Bake      src/math/ambcube.c:5-33           ordered fit of Pt_RayGen's samples
//   BakeSkyFn         src/rendering/render_system.c:403-425 + src/math/atmosphere.h:79-181
//   screenshot loop   src/rendering/render_system.c:707-727 (exposure, GT tonemap, sRGB, rgba8)
// Cubemaps live on the device as 6 contiguous faces: colour float3[size*size] per face,
// convolved float4[elemCount] per face (mip chain, cubemap.c:102-113).
#include "pt_device.h"
#include "kernels_common.cuh"

namespace pt {

namespace {

inline unsigned gridFor(int64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

// CalcMipOffset (src/rendering/sampler.h:42-51) for a square image
__device__ __host__ inline int mipOffset(int size, int m)
{
    int y = 0;
    for (int i = 0; i < m; ++i) { int s = size >> i; if (s < 1) s = 1; y += s * s; }
    return y;
}

// cubemap.c BakeFn's ray: one per face texel
__global__ void __launch_bounds__(256) k_cubeRays(int size, V3 origin, uint32_t pass, uint32_t seed, F4* __restrict__ ro,
                                                  F4* __restrict__ rd)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int flen = size * size;
    if (i >= 6 * flen) return;
    const int face = i / flen, fi = i % flen;
    Rng rng = seed_prng((uint32_t)i, pass, seed, Salt_CubeBake);
    const V2 Xi = f2_tent(rng_f2(rng));
    const V3 dir = Cubemap_CalcDir(size, face, fi % size, fi / size, Xi);
    ro[i] = f4(origin, 0.0f);
    rd[i] = f4(dir, 0.0f);
}

// PrefilterFn + PrefilterEnvMap (cubemap.c:189-260): split-sum N = V approximation.
// kLanes lanes share a mip texel (a 64^2 cubemap has 32 766 mip texels in all, a tenth of the GPU's
// threads): every lane runs the texel's random stream — a hash chain, so nobody can skip ahead — but
// evaluates only samples s = lane (mod kLanes); the partial sums are added across the lanes.
template <int kLanes>
__global__ void __launch_bounds__(128) k_cubeConvolve(const float* __restrict__ color, float* __restrict__ convolved, int size,
                                                      int elemCount, int mipCount, uint32_t sampleCount, float weight,
                                                      uint32_t pass, uint32_t seed)
{
    // all mips in one launch: find this lane group's mip (the reference runs one task per mip, and the
    // texel index inside the mip seeds the stream, cubemap.c:271-287)
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t rest = tid / kLanes;
    const int lane = (int)(tid % kLanes);
    int mip = 0, mSize = size;
    while (mip < mipCount && rest >= 6ll * mSize * mSize) { rest -= 6ll * mSize * mSize; ++mip; mSize >>= 1; }
    if (mip >= mipCount) return;        // whole lane groups leave together
    const int len = mSize * mSize;
    const int i = (int)rest;
    const int face = i / len, fi = i % len;
    const float roughness = (float)mip / 6.0f;          // MipToRoughness, CUBEMAP_MAX_MIP (cubemap.h:12,66-69)
    Rng rng = seed_prng((uint32_t)i, pass, seed, Salt_CubeConvolve | (uint32_t)(mip & 0xf));
    const V2 Xi = f2_tent(rng_f2(rng));
    const V3 N = Cubemap_CalcDir(mSize, face, fi % mSize, fi / mSize, Xi);
    const Tbn tbn = NormalToTBN(N);
    const V3 I = -tbn.c2;
    const float alpha = BrdfAlpha(roughness);
    float basisIntegral = 0.0f;
    F4 light = f4(0, 0, 0, 0);
    for (uint32_t s0 = 0; s0 < sampleCount; s0 += kLanes)
    {
        // kLanes draws of the stream in lockstep, lane j keeps draw j; then all lanes evaluate side by side
        V2 u = v2(0.0f, 0.0f);
        bool mine = false;
        for (int j = 0; j < kLanes; ++j)
        {
            if (s0 + (uint32_t)j >= sampleCount) break;
            const V2 d = rng_f2(rng);
            if (j == lane) { u = d; mine = true; }
        }
        if (!mine) continue;
        const V3 H = TbnToWorld(tbn, SampleGGXMicrofacet(u, alpha));
        const V3 L = reflect(I, H);
        const float NoL = dot(L, tbn.c2);
        if (NoL > 0.0f)
        {
            V2 uv;
            const int f = Cubemap_CalcUv(L, uv);                                   // Cubemap_ReadColor
            const V3 c = UvBilinearClamp_f3(color + (size_t)f * size * size * 3, size, size, uv);
            light.x += c.x * NoL; light.y += c.y * NoL; light.z += c.z * NoL; light.w += NoL;
            basisIntegral += NoL;
        }
    }
    if (kLanes > 1)
    {
        // lane groups are aligned sub-warps: xor shuffles below kLanes stay inside the group
        for (int o = kLanes >> 1; o > 0; o >>= 1)
        {
            light.x += __shfl_xor_sync(0xffffffffu, light.x, o); light.y += __shfl_xor_sync(0xffffffffu, light.y, o);
            light.z += __shfl_xor_sync(0xffffffffu, light.z, o); light.w += __shfl_xor_sync(0xffffffffu, light.w, o);
            basisIntegral += __shfl_xor_sync(0xffffffffu, basisIntegral, o);
        }
        if (lane != 0) return;
    }
    if (basisIntegral > 0.0f)
    {
        const float r = 1.0f / basisIntegral;
        light.x *= r; light.y *= r; light.z *= r; light.w *= r;
    }
    // Cubemap_BlendMip (cubemap.h:169-188)
    F4* dst = reinterpret_cast<F4*>(convolved) + (size_t)face * elemCount + mipOffset(size, mip) + fi;
    *dst = f4_lerp(*dst, light, weight);
}

// AmbCube_Bake's loop (ambcube.c:22-29) and AmbCube_Fit (ambcube.h:43-66). The fit is a chain of
// lerps, so the samples are folded in order by one thread.
__global__ void k_ambcubeFit(F4* __restrict__ cube, const F4* __restrict__ directions, const float* __restrict__ colors3,
                             int samples, int prevSampleCount)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    F4 v[6];
    for (int k = 0; k < 6; ++k) v[k] = cube[k];
    float w = 6.0f / (1.0f + samples);
    w = w / (1.0f + prevSampleCount);
    for (int i = 0; i < samples; ++i)
    {
        const F4 d = directions[i];
        const F4 rad = f4(colors3[3 * (size_t)i], colors3[3 * (size_t)i + 1], colors3[3 * (size_t)i + 2], 0.0f);
        const float dx = (d.x * d.x) * w, dy = (d.y * d.y) * w, dz = (d.z * d.z) * w;
        const int ix = d.x > 0.0f ? 0 : 1, iy = d.y > 0.0f ? 2 : 3, iz = d.z > 0.0f ? 4 : 5;
        v[ix] = f4_lerp(v[ix], rad, dx);
        v[iy] = f4_lerp(v[iy], rad, dy);
        v[iz] = f4_lerp(v[iz], rad, dz);
    }
    for (int k = 0; k < 6; ++k) cube[k] = v[k];
}

// Atmosphere (src/math/atmosphere.h:79-181): single-scattering raymarch with a fixed step of
// the median free path, marching towards the sun at every view step.
// The view ray starts exactly on the crust (height 0): the length uses the IEEE square root so that
// sqrt(r * r) == r there, as on the CPU; the approximate one can land one ulp below and end the march.
__device__ __forceinline__ float lengthExact(V3 a) { return sqrtr(dot(a, a)); }

// `lanes` > 1: the texel belongs to a warp. Every lane marches the view ray (cheap: one step per
// view sample), `lanes` samples at a time, and lane l then runs the sun march of sample l of that
// group, side by side with the other lanes; the partial in-scattering sums are added across the warp
// afterwards. That turns ~190 x 190 dependent steps per
// texel into ~190 + 6 x 190, which is what a 64^2 sky (24 576 texels, a sixth of the GPU's threads) needs.
__device__ V3 atmosphere(const SkyParams& A, V3 ro, V3 rd, V3 lightDir, V3 luminance, int steps, int lane, int lanes)
{
    const float rCrust = A.rCrust;
    const float rhoR = A.rhoR, rhoM = A.rhoM;
    const V3 muR = A.muR;
    const float muM = A.muM;
    const float majorant = f1_max(muM, hmax(muR)) * steps;
    const float mfp = SampleFreePath(0.5f, 1.0f / majorant);
    const float kMinDensity = 1e-5f;

    V3 tr_r = v3s(0.0f), tr_m = v3s(0.0f);
    float odR_v = 0.0f, odM_v = 0.0f, t_v = 0.0f;
    bool viewDone = false;
    while (!viewDone)
    {
        // `lanes` view samples in lockstep (every lane computes all of them); lane j keeps sample j
        bool mine = false;
        V3 myPos = v3s(0.0f);
        float myOdR_v = 0.0f, myOdM_v = 0.0f, myOdR_vi = 0.0f, myOdM_vi = 0.0f;
        for (int j = 0; j < lanes; ++j)
        {
            const V3 pos_v = ro + rd * t_v;
            const float h_v = lengthExact(pos_v) - rCrust;
            if (h_v < 0.0f) { viewDone = true; break; }
            const float densityR_v = expf(-h_v * rhoR);
            const float densityM_v = expf(-h_v * rhoM);
            if ((densityR_v + densityM_v) < kMinDensity) { viewDone = true; break; }
            const float dt_v = mfp;
            t_v += dt_v;
            const float odR_vi = densityR_v * dt_v, odM_vi = densityM_v * dt_v;
            odR_v += odR_vi;
            odM_v += odM_vi;
            if (j == lane)
            {
                mine = true; myPos = pos_v;
                myOdR_v = odR_v; myOdM_v = odM_v; myOdR_vi = odR_vi; myOdM_vi = odM_vi;
            }
        }
        if (!mine) continue;
        // the sun march of this lane's sample; the lanes of a warp run theirs side by side
        float odR_l = 0.0f, odM_l = 0.0f, t_l = 0.0f;
        bool hitCrust = false;
        while (true)
        {
            const V3 pos_l = myPos + lightDir * t_l;
            const float h_l = lengthExact(pos_l) - rCrust;
            if (h_l < 0.0f) { hitCrust = true; break; }
            const float densityR_l = expf(-h_l * rhoR);
            const float densityM_l = expf(-h_l * rhoM);
            if ((densityR_l + densityM_l) < kMinDensity) break;
            const float dt_l = mfp;
            t_l += dt_l;
            odR_l += densityR_l * dt_l;
            odM_l += densityM_l * dt_l;
        }
        if (!hitCrust)
        {
            const float m = muM * (myOdM_v + odM_l);
            const V3 od_i = muR * (myOdR_v + odR_l) + v3s(m);
            const V3 tr_i = v3(expf(-od_i.x), expf(-od_i.y), expf(-od_i.z));
            tr_r = tr_r + tr_i * myOdR_vi;
            tr_m = tr_m + tr_i * myOdM_vi;
        }
    }
    if (lanes > 1)
    {
        for (int o = lanes >> 1; o > 0; o >>= 1)
        {
            tr_r.x += __shfl_xor_sync(0xffffffffu, tr_r.x, o); tr_r.y += __shfl_xor_sync(0xffffffffu, tr_r.y, o);
            tr_r.z += __shfl_xor_sync(0xffffffffu, tr_r.z, o); tr_m.x += __shfl_xor_sync(0xffffffffu, tr_m.x, o);
            tr_m.y += __shfl_xor_sync(0xffffffffu, tr_m.y, o); tr_m.z += __shfl_xor_sync(0xffffffffu, tr_m.z, o);
        }
    }
    {
        const float cosTheta = dot(rd, lightDir);
        const float phR = (3.0f / (16.0f * kPi)) * (1.0f + cosTheta * cosTheta);      // RayleighPhase :31-34
        const float phM = MiePhase(cosTheta, A.gM);
        tr_r = tr_r * (muR * phR);
        tr_m = tr_m * (muM * phM);
    }
    return (tr_r + tr_m) * luminance;
}

// BakeSkyFn (render_system.c:403-425): the view ray starts on the crust at the north pole.
// kWarp: one warp per texel (small cubemaps), else one thread per texel.
template <bool kWarp>
__global__ void __launch_bounds__(128) k_skyBake(SkyParams A, int size, V3 sunDir, float sunLum, int steps, float* __restrict__ color)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = kWarp ? (tid >> 5) : tid;
    const int len = size * size;
    if (i >= 6ll * len) return;        // uniform per warp in the warp form
    const int face = (int)(i / len), t = (int)(i % len);
    const V3 rd = Cubemap_CalcDir(size, face, t % size, t / size, v2(0.0f, 0.0f));
    const V3 ro = v3(0.0f, A.rCrust, 0.0f);
    const V3 c = atmosphere(A, ro, rd, sunDir, v3s(sunLum), steps, kWarp ? (int)(threadIdx.x & 31) : 0, kWarp ? 32 : 1);
    if (!kWarp || (threadIdx.x & 31) == 0)
    {
        color[3 * (size_t)i] = c.x; color[3 * (size_t)i + 1] = c.y; color[3 * (size_t)i + 2] = c.z;
    }
}

// f1_GTTonemap (src/math/color.h:388-408) with CmdScreenshot's parameters (render_system.c:697-705)
__device__ __forceinline__ float gtTonemap(float x)
{
    const float P = 1.0f, a = 1.0f, m = 0.5f, l = 0.4f, c = 1.33f, b = 0.0f;
    const float l0 = ((P - m) * l) / a;
    const float S0 = m + l0;
    const float S1 = m + a * l0;
    const float C2 = (a * P) / (P - S1);
    const float L = m + a * (x - m);
    const float t = f1_sat((x - 0.0f) / (m - 0.0f));         // f1_smoothstep(0, m, x), scalar.h:188-250
    const float w0 = 1.0f - t * t * ((t * -2.0f) + 3.0f);
    const float w2 = (x < m + l0) ? 0.0f : 1.0f;
    const float w1 = 1.0f - w0 - w2;
    // the toe (a power) only counts below m and the shoulder (an exponential) only from m + l0 on: their
    // weights are exactly 0 elsewhere, so each is evaluated only where it is used (5 MUFU per channel
    // in the straight-line form made this kernel XU bound at 42 % of the HBM rate)
    const float T = (w0 != 0.0f) ? m * powf(x / m, c) + b : 0.0f;
    const float S = (w2 != 0.0f) ? P - (P - S1) * expf(-(C2 * (x - S0)) / P) : 0.0f;
    return T * w0 + L * w1 + S * w2;
}
__device__ __forceinline__ float srgbInverseEotf(float L)     // color.h:165-172
{
    if (L <= 0.0031308f) return L * 12.92f;
    return 1.055f * powf(L, 1.0f / 2.4f) - 0.055f;
}

// CmdScreenshot's pixel loop (render_system.c:707-727). Scene colours are Rec709
// (r_config.h:107), so Color_SceneToSDR is the identity. The 1/255 dither term is seeded per
// pixel (the reference draws it from the calling thread's serial generator).
__global__ void __launch_bounds__(256) k_tonemap(const F4* __restrict__ light, int64_t n, float exposure, uint32_t pass,
                                                 uint32_t seed, uint32_t* __restrict__ rgba8)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F4 v = light[i];
    v.x *= exposure; v.y *= exposure; v.z *= exposure;
    v.x = f1_max(v.x, 0.0f); v.y = f1_max(v.y, 0.0f); v.z = f1_max(v.z, 0.0f);
    v.x = gtTonemap(v.x); v.y = gtTonemap(v.y); v.z = gtTonemap(v.z);
    v.x = srgbInverseEotf(v.x); v.y = srgbInverseEotf(v.y); v.z = srgbInverseEotf(v.z);
    Rng rng = seed_prng((uint32_t)i, pass, seed, Salt_Tonemap);
    const F4 noise = rng_f4(rng);
    v = f4_lerp(v, noise, 1.0f / 255.0f);
    v.w = 1.0f;
    // f4_rgba8 (color.h:102-112)
    const uint32_t r = (uint32_t)(f1_sat(v.x) * 255.0f + 0.5f), g = (uint32_t)(f1_sat(v.y) * 255.0f + 0.5f);
    const uint32_t b = (uint32_t)(f1_sat(v.z) * 255.0f + 0.5f), a = (uint32_t)(f1_sat(v.w) * 255.0f + 0.5f);
    rgba8[i] = r | (g << 8) | (b << 16) | (a << 24);
}

// ---- N1: EmbedTaskFn (src/rendering/lightmap.c:947-1034) --------------------------------------
// The reference asks, for every texel of every page, every triangle of every drawable for its
// signed distance in lightmap space (texels x triangles). Here each triangle visits only the
// texels of its own bounding box grown by the fill padding and bids (distance, triangle index)
// on them with one 64-bit atomicMin; the lowest distance wins and equal distances go to the
// lower index — the order the reference's `dist < lmDist` scan keeps. A second pass recomputes
// the winner's barycentrics per texel. (A triangle with zero area in lightmap space is the one
// case that differs: sdTriangle2D's sign term is 0 for it, so the reference sees a negative
// distance on every texel of the page; here it only bids inside its padded box.)

constexpr float kFillPadding = 2.0f;                       // lightmap.c:40
constexpr unsigned long long kEmbedNone = ~0ull;

// float bits -> unsigned with the same order (negative distances are inside the triangle)
__device__ __forceinline__ uint32_t orderedBits(float f)
{
    const uint32_t u = f2u(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ void embedTriUv(const F4* __restrict__ uvs, int64_t tri, float texelSize, V2& A, V2& B, V2& C)
{
    const F4 a = uvs[3 * tri], b = uvs[3 * tri + 1], c = uvs[3 * tri + 2];
    A = v2(a.z * texelSize, a.w * texelSize);     // f2_mulvs(f2_v(uvs[a].z, uvs[a].w), texelSize)
    B = v2(b.z * texelSize, b.w * texelSize);
    C = v2(c.z * texelSize, c.w * texelSize);
}

// one warp per triangle, lanes stride over the padded bounding box
__global__ void __launch_bounds__(256) k_embedRaster(const EmbedMesh M, int lmSize, unsigned long long* __restrict__ keys)
{
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= M.triCount) return;
    const int page = M.page[warp];
    if (page < 0) return;
    V2 A, B, C;
    embedTriUv(M.uvs, warp, (float)lmSize, A, B, C);
    if (!(isfinite(A.x) && isfinite(A.y) && isfinite(B.x) && isfinite(B.y) && isfinite(C.x) && isfinite(C.y))) return;
    // texel centres within kFillPadding of the triangle lie inside its box grown by the padding
    const float pad = kFillPadding + 0.5f;
    int x0 = (int)floorf(f1_min(A.x, f1_min(B.x, C.x)) - pad), x1 = (int)ceilf(f1_max(A.x, f1_max(B.x, C.x)) + pad);
    int y0 = (int)floorf(f1_min(A.y, f1_min(B.y, C.y)) - pad), y1 = (int)ceilf(f1_max(A.y, f1_max(B.y, C.y)) + pad);
    x0 = x0 < 0 ? 0 : x0; y0 = y0 < 0 ? 0 : y0;
    x1 = x1 > lmSize - 1 ? lmSize - 1 : x1; y1 = y1 > lmSize - 1 ? lmSize - 1 : y1;
    if (x1 < x0 || y1 < y0) return;
    const int w = x1 - x0 + 1;
    const int64_t count = (int64_t)w * (y1 - y0 + 1);
    unsigned long long* dst = keys + (size_t)page * lmSize * lmSize;
    for (int64_t i = lane; i < count; i += 32)
    {
        const int x = x0 + (int)(i % w), y = y0 + (int)(i / w);
        const float dist = sdTriangle2D(A, B, C, v2(x + 0.5f, y + 0.5f));
        if (dist < kFillPadding)
        {
            const unsigned long long key = ((unsigned long long)orderedBits(dist) << 32) | (unsigned long long)(uint32_t)warp;
            atomicMin(dst + (size_t)y * lmSize + x, key);
        }
    }
}

__global__ void __launch_bounds__(256) k_embedResolve(const EmbedMesh M, const unsigned long long* __restrict__ keys,
                                                      int lmSize, int64_t texels, float* __restrict__ position3,
                                                      float* __restrict__ normal3, float* __restrict__ sampleCounts)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= texels) return;
    const unsigned long long key = keys[i];
    V3 pos = v3s(0.0f), nor = v3s(0.0f);
    float count = 0.0f;
    if (key != kEmbedNone)
    {
        const int64_t tri = (int64_t)(uint32_t)key;
        const int iTexel = (int)(i % ((int64_t)lmSize * lmSize));
        const V2 px = v2((iTexel % lmSize) + 0.5f, (iTexel / lmSize) + 0.5f);
        V2 A, B, C;
        embedTriUv(M.uvs, tri, (float)lmSize, A, B, C);
        float area = sdEdge2D(A, B, C);
        area = f1_max(area, 1e-5f);
        const float rcpArea = 1.0f / area;
        float w = sdEdge2D(B, C, px) * rcpArea, u = sdEdge2D(C, A, px) * rcpArea, v = sdEdge2D(A, B, px) * rcpArea;
        const float r = 1.0f / (w + u + v);                 // f4_divvs
        w *= r; u *= r; v *= r;
        pos = (v3(M.positions[3 * tri]) * w + v3(M.positions[3 * tri + 1]) * u) + v3(M.positions[3 * tri + 2]) * v;   // f4_blend
        nor = normalize((v3(M.normals[3 * tri]) * w + v3(M.normals[3 * tri + 1]) * u) + v3(M.normals[3 * tri + 2]) * v);
        nor = normalize(nor);                                // lightmap.c:1026-1029
        count = 1.0f;
    }
    sampleCounts[i] = count;
    position3[3 * i] = pos.x; position3[3 * i + 1] = pos.y; position3[3 * i + 2] = pos.z;
    normal3[3 * i] = nor.x; normal3[3 * i + 1] = nor.y; normal3[3 * i + 2] = nor.z;
}

// CreateRomeTexture's texel loop (src/rendering/gltf_model.c:1118-1139): the glTF metallic-roughness,
// occlusion and emission images, each already float4 and of its own size, resampled onto one
// size x size RGBA8 texture: R roughness (mr.y), O occlusion (occ.x), M metallic (mr.x), E packed emission.
__device__ __forceinline__ F4 bilinearClampF4(const F4* __restrict__ tex, int w, int h, V2 uv)
{
    const Bilinear bi = BilinearClamp(w, h, uv);
    return f4_bilerp(tex[bi.a], tex[bi.b], tex[bi.c], tex[bi.d], bi.fx, bi.fy);
}
__global__ void __launch_bounds__(256) k_repackRome(const F4* __restrict__ mr, int mrW, int mrH, const F4* __restrict__ occ, int occW,
                                                    int occH, const F4* __restrict__ em, int emW, int emH, int size,
                                                    uint32_t* __restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= size * size) return;
    const float rcpSize = 1.0f / (size - 1);
    const V2 uv = v2((i % size) * rcpSize, (i / size) * rcpSize);
    const F4 m = bilinearClampF4(mr, mrW, mrH, uv);
    const float o = bilinearClampF4(occ, occW, occH, uv).x;
    const F4 e = bilinearClampF4(em, emW, emH, uv);
    // PackEmission (src/math/color.h:582-586)
    const float pe = f1_sat(fmaxf(e.x, fmaxf(e.y, e.z)) * (1.0f / kEmissionScale));
    const float w = (pe > kEpsilon) ? sqrtf(pe) : 0.0f;
    const uint32_t r = (uint32_t)(f1_sat(m.y) * 255.0f + 0.5f), g = (uint32_t)(f1_sat(o) * 255.0f + 0.5f);
    const uint32_t b = (uint32_t)(f1_sat(m.x) * 255.0f + 0.5f), a = (uint32_t)(f1_sat(w) * 255.0f + 0.5f);
    dst[i] = r | (g << 8) | (b << 16) | (a << 24);         // f4_rgba8
}

} // namespace

int cubemapMipCount(int size)       // CalcMipCount, sampler.h:23-26
{
    int m = 0;
    while ((1 << (m + 1)) <= size) ++m;
    return 1 + m;
}
// cubemap.c:104-105: CalcMipOffset clamps its argument to the last mip, whose one texel is the + 1
int cubemapElemCount(int size) { return mipOffset(size, cubemapMipCount(size) - 1) + 1; }

bool launchCubeRays(int size, V3 origin, uint32_t pass, uint32_t seed, F4* dRo, F4* dRd, cudaStream_t stream)
{
    const int n = 6 * size * size;
    if (n <= 0) return true;
    k_cubeRays<<<gridFor(n, 256), 256, 0, stream>>>(size, origin, pass, seed, dRo, dRd);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

bool launchCubeConvolve(const float* dColor, float* dConvolved, int size, uint32_t sampleCount, float weight, uint32_t pass,
                        uint32_t seed, cudaStream_t stream)
{
    const int mipCount = cubemapMipCount(size), elemCount = cubemapElemCount(size);
    int64_t n = 0;
    for (int m = 0; m < mipCount; ++m) n += 6ll * (size >> m) * (size >> m);
    if (n <= 0) return true;
    // small cubemaps: 8 lanes per mip texel (every mip's texel count x 8 is a multiple of 8, so lane
    // groups never straddle a mip or the end of the range)
    if (n < 148 * 2048)
        k_cubeConvolve<8><<<gridFor(n * 8, 128), 128, 0, stream>>>(dColor, dConvolved, size, elemCount, mipCount, sampleCount, weight, pass, seed);
    else
        k_cubeConvolve<1><<<gridFor(n, 128), 128, 0, stream>>>(dColor, dConvolved, size, elemCount, mipCount, sampleCount, weight, pass, seed);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

bool launchAmbCubeFit(F4* dCube, const F4* dDirections, const float* dColors3, int samples, int prevSampleCount,
                      cudaStream_t stream)
{
    k_ambcubeFit<<<1, 32, 0, stream>>>(dCube, dDirections, dColors3, samples, prevSampleCount);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

bool launchSkyBake(const SkyParams& sky, int size, V3 sunDir, float sunLum, int steps, float* dColor, cudaStream_t stream)
{
    const int n = 6 * size * size;
    if (n <= 0) return true;
    // a thread per texel fills the GPU from about 300 K texels (148 SMs x 2048 threads); below that a
    // warp per texel shortens each texel's serial march 27 times
    if (n < 148 * 2048) k_skyBake<true><<<gridFor((int64_t)n * 32, 128), 128, 0, stream>>>(sky, size, sunDir, sunLum, steps, dColor);
    else k_skyBake<false><<<gridFor(n, 128), 128, 0, stream>>>(sky, size, sunDir, sunLum, steps, dColor);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

bool launchTonemap(const F4* dLight, int64_t n, float exposure, uint32_t pass, uint32_t seed, uint32_t* dRgba8,
                   cudaStream_t stream)
{
    if (n <= 0) return true;
    k_tonemap<<<gridFor(n, 256), 256, 0, stream>>>(dLight, n, exposure, pass, seed, dRgba8);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

bool launchRepackRome(const F4* dMr, int mrW, int mrH, const F4* dOcc, int occW, int occH, const F4* dEm, int emW, int emH,
                      int size, uint32_t* dDst, cudaStream_t stream)
{
    if (size <= 0) return true;
    k_repackRome<<<gridFor((int64_t)size * size, 256), 256, 0, stream>>>(dMr, mrW, mrH, dOcc, occW, occH, dEm, emW, emH, size, dDst);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// keys: one 64-bit word per texel of all pages, filled with 0xff.. by the caller
bool launchEmbed(const EmbedMesh& mesh, int lmSize, int lmCount, unsigned long long* dKeys,
                 float* dPosition3, float* dNormal3, float* dSampleCounts, cudaStream_t stream)
{
    const int64_t texels = (int64_t)lmCount * lmSize * lmSize;
    if (texels <= 0) return true;
    if (mesh.triCount > 0)
    {
        k_embedRaster<<<gridFor(mesh.triCount * 32, 256), 256, 0, stream>>>(mesh, lmSize, dKeys);
        g_counters.kernelLaunches++;
    }
    k_embedResolve<<<gridFor(texels, 256), 256, 0, stream>>>(mesh, dKeys, lmSize, texels, dPosition3, dNormal3, dSampleCounts);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

} // namespace pt
