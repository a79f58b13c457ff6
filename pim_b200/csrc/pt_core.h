// pt_core.h — scalar math of the path-tracing hot path as device functions.
//
// Every function names the reference function whose arithmetic it follows
// (file:line under /root/reference/src). They are written from the formulas,
// for one ray per thread, on a 3-float vector type that lives in registers.
// The file also compiles as plain C++ (PT_HD collapses to `inline`) so the
// host-side unit tests in tests/hostsim can step the same code on a CPU; the
// shipped library only ever instantiates it inside CUDA kernels.
#pragma once

#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define PT_HD __host__ __device__ __forceinline__
#define PT_D __device__ __forceinline__
#define PT_D_RARE __device__ __forceinline__   // cold paths; a __noinline__ call here measured 8% slower (profiles/README.md)
#else
#define PT_HD inline
#define PT_D inline
#define PT_D_RARE inline
#endif

namespace pt {

// ---------------------------------------------------------------- constants
// src/math/scalar.h:21-28
constexpr float kPi = 3.1415926535897932384626433832795f;
constexpr float kTau = 6.283185307179586476925286766559f;
constexpr float kRcpEpsilon = 8388608.0f;
constexpr float kEpsilon = 1.0f / 8388607.0f;
constexpr float kEpsilonSq = (float)(1.0 / 70368744177663.0);
constexpr float kLog2Epsilon = -23.0f;
constexpr float kMilli = 1e-3f;
constexpr float kEmissionScale = 100.0f;       // src/rendering/r_config.h:113
constexpr float kMinAlpha = 1.0f / 1024.0f;    // src/math/lighting.h:36-39

enum : uint32_t
{
    MatFlag_Emissive = 1u << 0,
    MatFlag_Sky = 1u << 1,
    MatFlag_Refractive = 1u << 5,
};
enum : int { Hit_Nothing = 0, Hit_Backface = 1, Hit_Triangle = 2 };

// ------------------------------------------------------------------ vectors
struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct alignas(16) F4 { float x, y, z, w; };
struct alignas(16) U4 { uint32_t x, y, z, w; };

PT_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
PT_HD V3 v3s(float s) { return v3(s, s, s); }
PT_HD V3 v3(const F4& a) { return v3(a.x, a.y, a.z); }
PT_HD V2 v2(float x, float y) { V2 r; r.x = x; r.y = y; return r; }
PT_HD F4 f4(float x, float y, float z, float w) { F4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
PT_HD F4 f4(const V3& a, float w) { return f4(a.x, a.y, a.z, w); }
PT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PT_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
PT_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
// one reciprocal and three multiplies, as -ffast-math (-freciprocal-math) compiles the reference's f4_divvs
PT_HD V3 operator/(V3 a, float s) { const float r = 1.0f / s; return v3(a.x * r, a.y * r, a.z * r); }
PT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
PT_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PT_HD V3 cross(V3 a, V3 b)
{
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
PT_HD float length(V3 a) { return sqrtf(dot(a, a)); }
PT_HD V3 normalize(V3 a) { return a / length(a); }                 // f4_normalize3
PT_HD float hmax(V3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }      // f4_hmax3
PT_HD float sum3(V3 a) { return a.x + a.y + a.z; }
PT_HD float avglum(V3 a) { return (a.x + a.y + a.z) * (1.0f / 3.0f); } // color.h:261
PT_HD V3 lerp(V3 a, V3 b, float t) { return a + (b - a) * t; }     // f4_lerpvs
PT_HD V3 inv(V3 a) { return v3(1.0f - a.x, 1.0f - a.y, 1.0f - a.z); } // f4_inv
PT_HD V3 vmax(V3 a, V3 b) { return v3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
PT_HD V3 reflect(V3 i, V3 n) { return i - (n * dot(i, n)) * 2.0f; } // f4_reflect3

// src/math/scalar.h
PT_HD float f1_max(float a, float b) { return a > b ? a : b; }
PT_HD float f1_min(float a, float b) { return a < b ? a : b; }
PT_HD float f1_sat(float x)          // NaN -> 0 (scalar.h:66-72)
{
    x = (x > 0.0f) ? x : 0.0f;
    x = (x < 1.0f) ? x : 1.0f;
    return x;
}
PT_HD float f1_clamp(float x, float lo, float hi) { return f1_min(f1_max(x, lo), hi); }
PT_HD float f1_abs(float x) { return f1_max(x, -x); }
PT_HD float f1_lerp(float a, float b, float t) { return a + (b - a) * t; }
PT_HD float f1_sq(float x) { return x * x; }
PT_HD float f1_floor(float x)        // scalar.h:139-147 (truncate then fix up)
{
    float i = (float)(int32_t)x;
    return i - ((i > x) ? 1.0f : 0.0f);
}
PT_HD float f1_frac(float x) { return x - f1_floor(x); }
PT_HD float f1_sign(float x) { return ((x > 0.0f) ? 1.0f : 0.0f) - ((x < 0.0f) ? 1.0f : 0.0f); }
PT_HD int i1_clamp(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }

// f4_refract3 (float4_funcs.h:713-719)
PT_HD V3 refract(V3 i, V3 n, float k)
{
    float cosTheta = f1_min(1.0f, dot(-i, n));
    V3 perp = (n * cosTheta + i) * k;
    V3 par = n * (-sqrtf(f1_abs(1.0f - dot(perp, perp))));
    return normalize(perp + par);
}

// ---------------------------------------------------------------------- rng
// src/math/pcg.h:26-172, src/common/random.h:13-152
struct Rng { uint32_t x, y, z, w; };

PT_HD uint32_t Pcg1(uint32_t v)
{
    v = v * 747796405u + 2891336453u;
    v = ((v >> ((v >> 28) + 4u)) ^ v) * 277803737u;
    v = (v >> 22) ^ v;
    return v;
}
PT_HD void Pcg2(uint32_t& x, uint32_t& y)
{
    x = x * 1664525u + 1013904223u;
    y = y * 1664525u + 1013904223u;
    x += y * 1664525u;
    y += x * 1664525u;
    x ^= x >> 16;
    y ^= y >> 16;
    x += y * 1664525u;
    y += x * 1664525u;
}
PT_HD Rng Pcg4(uint32_t x, uint32_t y, uint32_t z, uint32_t w)
{
    x = x * 1664525u + 1013904223u;
    y = y * 1664525u + 1013904223u;
    z = z * 1664525u + 1013904223u;
    w = w * 1664525u + 1013904223u;
    x += y * w; y += z * x; z += x * y; w += y * z;
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16; w ^= w >> 16;
    x += y * w; y += z * x; z += x * y; w += y * z;
    Rng r; r.x = x; r.y = y; r.z = z; r.w = w;
    return r;
}
PT_HD float ToFloat(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }
PT_HD uint32_t rng_u32(Rng& r) { r.x = Pcg1(r.x); return r.x; }
PT_HD float rng_f32(Rng& r) { r.x = Pcg1(r.x); return ToFloat(r.x); }
PT_HD V2 rng_f2(Rng& r) { Pcg2(r.x, r.y); return v2(ToFloat(r.x), ToFloat(r.y)); }
PT_HD F4 rng_f4(Rng& r)
{
    r = Pcg4(r.x, r.y, r.z, r.w);
    return f4(ToFloat(r.x), ToFloat(r.y), ToFloat(r.z), ToFloat(r.w));
}

// Work-item seeding shared with the oracle harness (oracle/ref_stubs.c seed_item):
// the path-tracer context of work item `item` in pass `pass` starts from
// Pcg4(item, pass, seed, salt ^ 0x80000000), the Prng_Get() stream from
// Pcg4(item, pass, seed, salt).
enum : uint32_t
{
    Salt_Emission = 0x10, Salt_LightGrid = 0x11,
    Salt_Trace = 0x21, Salt_RayGen = 0x31, Salt_Bake = 0x41, Salt_Lut = 0x50, Salt_Rays = 0x60,
    Salt_CubeBake = 0x80, Salt_CubeConvolve = 0x90 /* | mip */, Salt_Tonemap = 0xb0,
};
PT_HD Rng seed_ctx(uint32_t item, uint32_t pass, uint32_t seed, uint32_t salt)
{
    return Pcg4(item, pass, seed, salt ^ 0x80000000u);
}
PT_HD Rng seed_prng(uint32_t item, uint32_t pass, uint32_t seed, uint32_t salt)
{
    return Pcg4(item, pass, seed, salt);
}

// ----------------------------------------------------------------- sampling
// src/math/sampling.h
struct Tbn { V3 c0, c1, c2; };

PT_HD Tbn NormalToTBN(V3 N)              // sampling.h:26-59 (Duff et al.)
{
    float s = (N.z < 0.0f) ? -1.0f : 1.0f;
    const float a = -1.0f / (s + N.z);
    const float b = N.x * N.y * a;
    Tbn t;
    t.c0 = v3(1.0f + s * N.x * N.x * a, s * b, -s * N.x);
    t.c1 = v3(b, s + N.y * N.y * a, -N.y);
    t.c2 = N;
    return t;
}
PT_HD V3 TbnToWorld(const Tbn& t, V3 n)  // f3x3_mul_col
{
    return t.c0 * n.x + (t.c1 * n.y + t.c2 * n.z);
}
PT_HD V3 TanToWorld(V3 N, V3 nTS) { return TbnToWorld(NormalToTBN(N), nTS); }

PT_HD float RadicalInverseBase2(uint32_t bits)   // sampling.h:75-83
{
    bits = (bits << 16u) | (bits >> 16u);
    bits = ((bits & 0x55555555u) << 1u) | ((bits & 0xAAAAAAAAu) >> 1u);
    bits = ((bits & 0x33333333u) << 2u) | ((bits & 0xCCCCCCCCu) >> 2u);
    bits = ((bits & 0x0F0F0F0Fu) << 4u) | ((bits & 0xF0F0F0F0u) >> 4u);
    bits = ((bits & 0x00FF00FFu) << 8u) | ((bits & 0xFF00FF00u) >> 8u);
    return (float)(bits) * 2.3283064365386963e-10f;
}
PT_HD V2 Hammersley2D(uint32_t i, uint32_t N) { return v2((i + 0.5f) / N, RadicalInverseBase2(i)); }

PT_HD float PowerHeuristic(float f, float g)     // sampling.h:93-96
{
    return (f * f) / f1_max(f * f + g * g, kEpsilon);
}

PT_HD V2 MapSquareToDisk(V2 Xi)                  // sampling.h:100-118
{
    Xi.x = kEpsilon + Xi.x * ((1.0f - kEpsilon) - kEpsilon);
    Xi.y = kEpsilon + Xi.y * ((1.0f - kEpsilon) - kEpsilon);
    float phi, r;
    float a = 2.0f * Xi.x - 1.0f;
    float b = 2.0f * Xi.y - 1.0f;
    if ((a * a) > (b * b))
    {
        r = a;
        phi = (kPi / 4.0f) * (b / a);
    }
    else
    {
        r = b;
        phi = (kPi / 2.0f) - (kPi / 4.0f) * (a / b);
    }
    return v2(r * cosf(phi), r * sinf(phi));
}

PT_HD V3 SampleBaryCoord(V2 Xi)                  // sampling.h:120-128, returns (w,u,v)
{
    float r1 = sqrtf(f1_max(Xi.x, kEpsilonSq));
    float r2 = Xi.y;
    float u = r1 * (1.0f - r2);
    float v = r2 * r1;
    float w = 1.0f - (u + v);
    return v3(w, u, v);
}

PT_HD V2 SampleNGon(V2 Xi, uint32_t side, uint32_t N, float rot)   // sampling.h:130-139
{
    side %= N;
    const float R = kTau / N;
    const float a = rot + (1 + side) * R;
    const float b = rot + (2 + side) * R;
    V3 wuv = SampleBaryCoord(Xi);
    return v2(cosf(a) * wuv.y + cosf(b) * wuv.z, sinf(a) * wuv.y + sinf(b) * wuv.z);
}

PT_HD V2 SamplePentagram(V2 Xi, uint32_t side)   // sampling.h:141-156
{
    const float R = kTau / 5.0f;
    const float s = kPi * 0.1f;
    const float q = (1.0f - 0.2360679774997896964091736687313f) * 0.5f;
    side %= 5;
    const float a = s + (1.0f + side) * R;
    const float b = s + (1.5f + side) * R;
    const float c = s + (2.0f + side) * R;
    V2 A = v2(q * cosf(a), q * sinf(a));
    V2 B = v2(cosf(b), sinf(b));
    V2 C = v2(q * cosf(c), q * sinf(c));
    // f2_bilerp(A, B, 0, C, Xi)
    float tx = A.x + (B.x - A.x) * Xi.x, ty = A.y + (B.y - A.y) * Xi.x;
    float bx = 0.0f + (C.x - 0.0f) * Xi.x, by = 0.0f + (C.y - 0.0f) * Xi.x;
    return v2(tx + (bx - tx) * Xi.y, ty + (by - ty) * Xi.y);
}

PT_HD V3 SphericalToCartesian(float cosTheta, float phi)   // sampling.h:158-165
{
    float sinTheta = sqrtf(f1_max(1.0f - cosTheta * cosTheta, 0.0f));
    return v3(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
}
PT_HD V3 SampleUnitSphere(V2 Xi) { return SphericalToCartesian(Xi.y * 2.0f - 1.0f, kTau * Xi.x); }
PT_HD V3 SampleUnitHemisphere(V2 Xi) { return SphericalToCartesian(Xi.y, kTau * Xi.x); }

PT_HD V3 SampleCosineHemisphere(V2 Xi)           // sampling.h:267-272
{
    Xi = MapSquareToDisk(Xi);
    float z = sqrtf(f1_max(1.0f - (Xi.x * Xi.x + Xi.y * Xi.y), kEpsilonSq));
    return v3(Xi.x, Xi.y, z);
}

PT_HD V3 SampleGGXMicrofacet(V2 Xi, float alpha) // sampling.h:276-283
{
    float a2 = alpha * alpha;
    float phi = kTau * Xi.x;
    float b = f1_max(1.0f + (a2 - 1.0f) * Xi.y, kEpsilon);
    float cosTheta = sqrtf(f1_max((1.0f - Xi.y) / b, kEpsilonSq));
    return SphericalToCartesian(cosTheta, phi);
}

PT_HD float LightPdf(float area, float cosTheta, float distSq)  // sampling.h:317-321
{
    return distSq / f1_max(cosTheta * area, kEpsilon);
}

PT_HD V2 SampleGaussPixelFilter(V2 Xi, float stddev)   // sampling.h:323-331, scalar.h:299-302
{
    float angle = Xi.x * kTau;
    float radius = 0.0f + stddev * sqrtf(-logf(f1_max(1.0f - Xi.y, kEpsilon)));
    return v2(cosf(angle) * radius, sinf(angle) * radius);
}

PT_HD float SampleFreePath(float Xi, float mfp)  // sampling.h:336-339
{
    return -logf(f1_max(1.0f - Xi, kEpsilon)) * mfp;
}

// ----------------------------------------------------------------- lighting
// src/math/lighting.h
PT_HD float BrdfAlpha(float roughness) { return f1_max(roughness * roughness, kMinAlpha); }
PT_HD V3 F_0(V3 albedo, float metallic) { return lerp(v3s(0.04f), albedo, metallic); }
PT_HD float F_90(V3 F0) { return f1_sat(50.0f * dot(F0, v3s(0.33f))); }
PT_HD float F_Schlick1(float f0, float f90, float cosTheta)
{
    float t = 1.0f - cosTheta;
    float t5 = t * t * t * t * t;
    return f1_lerp(f0, f90, t5);
}
PT_HD float F_Dielectric(float cosThetaI, float etaI, float etaT)   // lighting.h:138-164
{
    cosThetaI = f1_clamp(cosThetaI, -1.0f, 1.0f);
    if (cosThetaI < 0.0f)
    {
        cosThetaI = -cosThetaI;
        float tmp = etaI; etaI = etaT; etaT = tmp;
    }
    float sinThetaI = sqrtf(f1_max(1.0f - cosThetaI * cosThetaI, kEpsilonSq));
    float sinThetaT = (etaI / etaT) * sinThetaI;
    if (sinThetaT >= 1.0f)
    {
        return 1.0f;
    }
    float cosThetaT = sqrtf(f1_max(1.0f - sinThetaT * sinThetaT, kEpsilonSq));
    float Rparl = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    float Rperp = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return f1_sat((Rparl * Rparl + Rperp * Rperp) * 0.5f);
}
PT_HD float D_GTR(float NoH, float alpha)        // lighting.h:220-226
{
    float a2 = alpha * alpha;
    float f = f1_lerp(1.0f, a2, NoH * NoH);
    f = f * f * kPi;
    return a2 / f1_max(f, kEpsilon);
}
PT_HD float GGXPdf(float NoH, float HoV, float alpha)   // sampling.h:306-310
{
    float d = D_GTR(NoH, alpha);
    return (d * NoH) / f1_max(4.0f * HoV, kEpsilon);
}
PT_HD float V_SmithCorrelated(float NoL, float NoV, float alpha)  // lighting.h:249-256
{
    float a2 = alpha * alpha;
    float v = NoL * sqrtf(f1_max(a2 + (NoV - NoV * a2) * NoV, kEpsilonSq));
    float l = NoV * sqrtf(f1_max(a2 + (NoL - NoL * a2) * NoL, kEpsilonSq));
    return 0.5f / f1_max(v + l, kEpsilon);
}
PT_HD float Fd_Burley(float NoL, float NoV, float HoV, float roughness)  // lighting.h:269-279
{
    float fd90 = 0.5f + 2.0f * HoV * HoV * roughness;
    float lightScatter = F_Schlick1(1.0f, fd90, NoL);
    float viewScatter = F_Schlick1(1.0f, fd90, NoV);
    return (lightScatter * viewScatter) / kPi;
}
PT_HD V3 AlbedoToTransmittance(V3 albedo, float beta_n, float thickness)  // lighting.h:196-216
{
    float r2 = beta_n * beta_n;
    float r3 = r2 * beta_n;
    float r4 = r3 * beta_n;
    float r5 = r4 * beta_n;
    float t = 5.969f - 0.215f * beta_n + 2.532f * r2 - 10.73f * r3 + 5.574f * r4 + 0.245f * r5;
    V3 sigmaA = v3(logf(albedo.x), logf(albedo.y), logf(albedo.z));
    sigmaA = sigmaA / f1_max(t, kEpsilon);
    sigmaA = sigmaA * sigmaA;
    return v3(expf(sigmaA.x * -thickness), expf(sigmaA.y * -thickness), expf(sigmaA.z * -thickness));
}
PT_HD float MiePhase(float cosTheta, float g)    // src/math/atmosphere.h:36-43
{
    float k = (3.0f / (8.0f * kPi)) * (1.0f - g * g) / (2.0f + g * g);
    float l = 1.0f + g * g - 2.0f * g * cosTheta;
    l = l * sqrtf(f1_max(kEpsilonSq, l));
    return k * (1.0f + cosTheta * cosTheta) / f1_max(kEpsilon, l);
}

// ------------------------------------------------------------------ texture
// src/rendering/sampler.h:168-250, 327-392; src/math/color.h:185-196, 241-259
struct Bilinear { int a, b, c, d; float fx, fy; };

PT_HD void LinearClamp(int size, float u, int& a, int& b, float& t)
{
    float x = f1_sat(u) * (size - 1);
    a = (int)x;
    t = x - (float)a;
    b = (a + 1 < size - 1) ? a + 1 : size - 1;
}
PT_HD Bilinear BilinearClamp(int w, int h, V2 uv)
{
    int xa, xb, ya, yb; float tx, ty;
    LinearClamp(w, uv.x, xa, xb, tx);
    LinearClamp(h, uv.y, ya, yb, ty);
    Bilinear r;
    r.fx = tx; r.fy = ty;
    r.a = ya * w + xa; r.b = ya * w + xb; r.c = yb * w + xa; r.d = yb * w + xb;
    return r;
}
PT_HD Bilinear BilinearWrap(int w, int h, V2 uv)
{
    uv.x = (uv.x >= 0.0f) ? uv.x : (1.0f - uv.x);
    uv.y = (uv.y >= 0.0f) ? uv.y : (1.0f - uv.y);
    uv.x = f1_frac(uv.x);
    uv.y = f1_frac(uv.y);
    return BilinearClamp(w, h, uv);
}
PT_HD float sRGB_EOTF_Fit(float V)
{
    return 0.020883f * V + 0.656075f * (V * V) + 0.324285f * (V * V * V);
}
PT_HD float sRGB_InverseEOTF_Fit(float L)
{
    float l1 = sqrtf(L);
    float l2 = sqrtf(l1);
    float l3 = sqrtf(l2);
    return 0.658444f * l1 + 0.643378f * l2 - 0.298148f * l3;
}
PT_HD F4 GammaDecode_rgba8(uint32_t c)
{
    const float s = 1.0f / 255.0f;
    return f4(sRGB_EOTF_Fit((float)(c & 0xff) * s), sRGB_EOTF_Fit((float)((c >> 8) & 0xff) * s),
              sRGB_EOTF_Fit((float)((c >> 16) & 0xff) * s), sRGB_EOTF_Fit((float)(c >> 24) * s));
}
PT_HD F4 f4_lerp(F4 a, F4 b, float t)
{
    return f4(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t, a.z + (b.z - a.z) * t, a.w + (b.w - a.w) * t);
}
PT_HD F4 f4_bilerp(F4 a, F4 b, F4 c, F4 d, float fx, float fy)
{
    return f4_lerp(f4_lerp(a, b, fx), f4_lerp(c, d, fx), fy);
}
PT_HD F4 UvBilinearWrap_c32(const uint32_t* texels, int w, int h, V2 uv)
{
    Bilinear bi = BilinearWrap(w, h, uv);
    // all four taps on one texel (1x1 textures, clamped borders): a + (a-a)*t is a, exactly
    if (bi.a == bi.d) return GammaDecode_rgba8(texels[bi.a]);
    return f4_bilerp(GammaDecode_rgba8(texels[bi.a]), GammaDecode_rgba8(texels[bi.b]),
                     GammaDecode_rgba8(texels[bi.c]), GammaDecode_rgba8(texels[bi.d]), bi.fx, bi.fy);
}
PT_HD V3 Xy16ToNormalTs(uint32_t xy)
{
    float x = (float)(int16_t)(xy & 0xffff) * (1.0f / 32768.0f);
    float y = (float)(int16_t)(xy >> 16) * (1.0f / 32768.0f);
    float z = sqrtf(f1_max(kEpsilonSq, 1.0f - (x * x + y * y)));
    return v3(x, y, z);
}
PT_HD V3 UvBilinearWrap_xy16(const uint32_t* texels, int w, int h, V2 uv)
{
    Bilinear bi = BilinearWrap(w, h, uv);
    V3 a = Xy16ToNormalTs(texels[bi.a]), b = Xy16ToNormalTs(texels[bi.b]);
    V3 c = Xy16ToNormalTs(texels[bi.c]), d = Xy16ToNormalTs(texels[bi.d]);
    return normalize(lerp(lerp(a, b, bi.fx), lerp(c, d, bi.fx), bi.fy));      // BilinearBlend_xy16, sampler.h:250-260
}
PT_HD V2 UvBilinearClamp_f2(const float* texels, int w, int h, V2 uv)
{
    Bilinear bi = BilinearClamp(w, h, uv);
    float ax = texels[2 * bi.a], ay = texels[2 * bi.a + 1];
    float bx = texels[2 * bi.b], by = texels[2 * bi.b + 1];
    float cx = texels[2 * bi.c], cy = texels[2 * bi.c + 1];
    float dx = texels[2 * bi.d], dy = texels[2 * bi.d + 1];
    float tx = ax + (bx - ax) * bi.fx, ty = ay + (by - ay) * bi.fx;
    float ux = cx + (dx - cx) * bi.fx, uy = cy + (dy - cy) * bi.fx;
    return v2(tx + (ux - tx) * bi.fy, ty + (uy - ty) * bi.fy);
}
PT_HD V3 UvBilinearClamp_f3(const float* texels, int w, int h, V2 uv)
{
    Bilinear bi = BilinearClamp(w, h, uv);
    V3 a = v3(texels[3 * bi.a], texels[3 * bi.a + 1], texels[3 * bi.a + 2]);
    V3 b = v3(texels[3 * bi.b], texels[3 * bi.b + 1], texels[3 * bi.b + 2]);
    V3 c = v3(texels[3 * bi.c], texels[3 * bi.c + 1], texels[3 * bi.c + 2]);
    V3 d = v3(texels[3 * bi.d], texels[3 * bi.d + 1], texels[3 * bi.d + 2]);
    return lerp(lerp(a, b, bi.fx), lerp(c, d, bi.fx), bi.fy);
}
// sampler.h:105-139 (point sampling used by EmissionPdf)
PT_HD int PointWrap1D(int size, float u)
{
    u = (u >= 0.0f) ? u : (1.0f - u);
    u = f1_frac(u);
    return (int)(f1_sat(u) * (size - 1) + 0.5f);
}

// ------------------------------------------------------------------ cubemap
// src/rendering/cubemap.h:71-99 + cubemap.c:24-42; returns the face, writes uv
PT_HD int Cubemap_CalcUv(V3 dir, V2& uvOut)
{
    V3 a = v3(f1_abs(dir.x), f1_abs(dir.y), f1_abs(dir.z));
    float v = hmax(a);
    float ma = 0.5f / v;
    int face;
    V3 right, up;
    if (v == a.x)
    {
        face = dir.x < 0.0f ? 1 : 0;
        right = dir.x < 0.0f ? v3(0, 0, 1) : v3(0, 0, -1);
        up = v3(0, 1, 0);
    }
    else if (v == a.y)
    {
        face = dir.y < 0.0f ? 3 : 2;
        right = dir.y < 0.0f ? v3(-1, 0, 0) : v3(1, 0, 0);
        up = v3(0, 0, -1);
    }
    else
    {
        face = dir.z < 0.0f ? 5 : 4;
        right = dir.z < 0.0f ? v3(-1, 0, 0) : v3(1, 0, 0);
        up = v3(0, 1, 0);
    }
    uvOut = v2(dot(right, dir) * ma + 0.5f, dot(up, dir) * ma + 0.5f);
    return face;
}

// f2_tent (src/math/float2_funcs.h:630-636). The y lane reads 2 - Xi.x after Xi.x has already
// been replaced: kept as the reference has it.
PT_HD V2 f2_tent(V2 Xi)
{
    Xi.x *= 2.0f; Xi.y *= 2.0f;
    Xi.x = Xi.x < 1.0f ? sqrtf(Xi.x) - 1.0f : 1.0f - sqrtf(2.0f - Xi.x);
    Xi.y = Xi.y < 1.0f ? sqrtf(Xi.y) - 1.0f : 1.0f - sqrtf(2.0f - Xi.x);
    return Xi;
}

// Cubemap_CalcDir (src/rendering/cubemap.h:190-207): BilinearInvert, jitter, snorm, proj_dir with
// the face frames of src/rendering/cubemap.c:14-42
PT_HD V3 Cubemap_CalcDir(int size, int face, int x, int y, V2 Xi)
{
    V2 uv = v2((float)x / (float)(size - 1), (float)y / (float)(size - 1));
    const float r = 1.0f / size;
    uv.x += Xi.x * r; uv.y += Xi.y * r;
    uv.x = uv.x * 2.0f - 1.0f; uv.y = uv.y * 2.0f - 1.0f;
    V3 fwd, up, right;
    switch (face)
    {
    case 0: fwd = v3(1, 0, 0); up = v3(0, 1, 0); right = v3(0, 0, -1); break;
    case 1: fwd = v3(-1, 0, 0); up = v3(0, 1, 0); right = v3(0, 0, 1); break;
    case 2: fwd = v3(0, 1, 0); up = v3(0, 0, -1); right = v3(1, 0, 0); break;
    case 3: fwd = v3(0, -1, 0); up = v3(0, 0, -1); right = v3(-1, 0, 0); break;
    case 4: fwd = v3(0, 0, 1); up = v3(0, 1, 0); right = v3(1, 0, 0); break;
    default: fwd = v3(0, 0, -1); up = v3(0, 1, 0); right = v3(-1, 0, 0); break;
    }
    return normalize((fwd + right * uv.x) + up * uv.y);     // proj_dir, frustum.h:33-47, slope 1
}

// --------------------------------------------------------------------- grid
// src/math/grid.h:8-49
struct Grid
{
    V3 lo, hi;
    int sx, sy, sz;
    float cellsPerMeter;
};
PT_HD int Grid_Len(const Grid& g) { return g.sx * g.sy * g.sz; }
PT_HD V3 Grid_Position(const Grid& g, int index)
{
    const float metersPerCell = 1.0f / g.cellsPerMeter;
    int ix = index % g.sx;
    int iy = (index / g.sx) % g.sy;
    int iz = index / (g.sx * g.sy);
    return g.lo + v3((ix + 0.5f) * metersPerCell, (iy + 0.5f) * metersPerCell, (iz + 0.5f) * metersPerCell);
}
PT_HD int Grid_Index(const Grid& g, V3 p)
{
    V3 o = (p - g.lo) * g.cellsPerMeter;
    int x = i1_clamp((int)o.x, 0, g.sx - 1);
    int y = i1_clamp((int)o.y, 0, g.sy - 1);
    int z = i1_clamp((int)o.z, 0, g.sz - 1);
    return x + y * g.sx + z * g.sx * g.sy;
}

// ------------------------------------------------------------------- dist1d
// src/math/dist1d.c:75-93 (FindInterval over cdf[size])
PT_HD int FindInterval(const float* cdf, int size, float u)
{
    int first = 0;
    int len = size;
    while (len > 0)
    {
        int half = len >> 1;
        int middle = first + half;
        if (cdf[middle] <= u)
        {
            first = middle + 1;
            len = len - (half + 1);
        }
        else
        {
            len = half;
        }
    }
    return i1_clamp(first - 1, 0, size - 2);
}

// sdEdge2D, sdTriangle2D: src/math/sdf.h:99-156
PT_HD float sdEdge2D(V2 A, V2 B, V2 P)
{
    const float apx = P.x - A.x, apy = P.y - A.y, abx = B.x - A.x, aby = B.y - A.y;
    return 0.5f * (apy * abx - apx * aby);
}
PT_HD float dot2(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }
PT_HD float sdTriangle2D(V2 a, V2 b, V2 c, V2 p)
{
    const V2 ba = v2(b.x - a.x, b.y - a.y), cb = v2(c.x - b.x, c.y - b.y), ac = v2(a.x - c.x, a.y - c.y);
    const V2 pa = v2(p.x - a.x, p.y - a.y), pb = v2(p.x - b.x, p.y - b.y), pc = v2(p.x - c.x, p.y - c.y);
    const float h0 = f1_sat(dot2(pa, ba) / dot2(ba, ba));
    const float h1 = f1_sat(dot2(pb, cb) / dot2(cb, cb));
    const float h2 = f1_sat(dot2(pc, ac) / dot2(ac, ac));
    const V2 pq0 = v2(pa.x - ba.x * h0, pa.y - ba.y * h0);
    const V2 pq1 = v2(pb.x - cb.x * h1, pb.y - cb.y * h1);
    const V2 pq2 = v2(pc.x - ac.x * h2, pc.y - ac.y * h2);
    const float s = f1_sign(ba.x * ac.y - ba.y * ac.x);
    const float dx = f1_min(dot2(pq0, pq0), f1_min(dot2(pq1, pq1), dot2(pq2, pq2)));
    const float dy = f1_min(s * (pa.x * ba.y - pa.y * ba.x), f1_min(s * (pb.x * cb.y - pb.y * cb.x), s * (pc.x * ac.y - pc.y * ac.x)));
    const float dySign = (dy >= 0.0f) ? 1.0f : -1.0f;
    return -sqrtf(f1_max(dx, kEpsilonSq)) * dySign;
}

// src/math/sdf.h:158-189
PT_HD float sdTriangle3D(V3 a, V3 b, V3 c, V3 p)
{
    V3 ba = b - a, pa = p - a;
    V3 cb = c - b, pb = p - b;
    V3 ac = a - c, pc = p - c;
    V3 nor = cross(ba, ac);
    float t0 = f1_sign(dot(cross(ba, nor), pa));
    float t1 = f1_sign(dot(cross(cb, nor), pb));
    float t2 = f1_sign(dot(cross(ac, nor), pc));
    float t;
    if ((t0 + t1 + t2) < 2.0f)
    {
        t0 = f1_sat(dot(ba, pa) / dot(ba, ba));
        t1 = f1_sat(dot(cb, pb) / dot(cb, cb));
        t2 = f1_sat(dot(ac, pc) / dot(ac, ac));
        V3 d0 = ba * t0 - pa, d1 = cb * t1 - pb, d2 = ac * t2 - pc;
        t = f1_min(dot(d0, d0), f1_min(dot(d1, d1), dot(d2, d2)));
    }
    else
    {
        t0 = dot(nor, pa);
        t = t0 * t0 / dot(nor, nor);
    }
    return sqrtf(f1_max(t, kEpsilonSq));
}

// src/math/area.h:28-31
PT_HD float TriArea3D(V3 A, V3 B, V3 C) { return 0.5f * length(cross(B - A, C - A)); }

// src/math/sphgauss.h:22-37, sphgauss.c:19-56 (5 lobes, axis.w = sharpness, amp.w = basis weight)
PT_HD float SG_BasisEval(F4 axis, V3 dir)
{
    float cosTheta = dir.x * axis.x + dir.y * axis.y + dir.z * axis.z;
    return expf(axis.w * (cosTheta - 1.0f));
}
PT_HD void SG_Accumulate(float sampleWeight, V3 dir, V3 rad, const F4* axii, F4* amplitudes, int length)
{
    if (sampleWeight == 1.0f)
    {
        for (int i = 0; i < length; ++i) amplitudes[i] = f4(0, 0, 0, 0);
    }
    F4 estimate = f4(0, 0, 0, 0);
    for (int i = 0; i < length; ++i)
    {
        float b = SG_BasisEval(axii[i], dir);
        estimate.x += amplitudes[i].x * b; estimate.y += amplitudes[i].y * b;
        estimate.z += amplitudes[i].z * b; estimate.w += amplitudes[i].w * b;
    }
    for (int i = 0; i < length; ++i)
    {
        float basis = SG_BasisEval(axii[i], dir);
        if (basis > 0.0f)
        {
            F4 amplitude = amplitudes[i];
            float weight = amplitude.w;
            weight = f1_lerp(weight, basis, sampleWeight);
            float s = basis / weight;
            // rad.w is 0 in the reference's call (lightmap.c:1189-1195)
            F4 other = f4(estimate.x - amplitude.x * basis, estimate.y - amplitude.y * basis,
                          estimate.z - amplitude.z * basis, estimate.w - amplitude.w * basis);
            F4 thisLobe = f4((rad.x - other.x) * s, (rad.y - other.y) * s, (rad.z - other.z) * s, (0.0f - other.w) * s);
            amplitude = f4_lerp(amplitude, thisLobe, sampleWeight);
            amplitude.x = f1_max(amplitude.x, 0.0f); amplitude.y = f1_max(amplitude.y, 0.0f);
            amplitude.z = f1_max(amplitude.z, 0.0f);
            amplitude.w = weight;
            amplitudes[i] = amplitude;
        }
    }
}

} // namespace pt
