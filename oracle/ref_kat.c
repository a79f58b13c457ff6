/* TEST INFRASTRUCTURE — not part of the shipped library.
 *
 * Known-answer vectors for the helper functions of SURVEY.md 8(a), produced by calling the
 * reference's own inline functions (headers under /root/reference/src, used as they are).
 * Each KAT maps n rows of `in` (kin floats per row) to n rows of `out` (kout floats per row);
 * integer results are written as float bit patterns. tests/golden/make_golden.py stores the
 * outputs for fixed pcg32-seeded inputs in tests/golden/kat.npz; tests/test_cpu_kat.py checks
 * the library's device headers (compiled for the host) and this oracle against them.
 */
#include "math/types.h"
#include "math/float2_funcs.h"
#include "math/float4_funcs.h"
#include "math/pcg.h"
#include "math/sampling.h"
#include "math/lighting.h"
#include "math/sdf.h"
#include "math/grid.h"
#include "math/color.h"
#include "math/sphgauss.h"
#include "math/dist1d.h"
#include "math/atmosphere.h"
#include "rendering/cubemap.h"
#include "rendering/sampler.h"
#include <string.h>

static float u2f(u32 u) { float f; memcpy(&f, &u, 4); return f; }
static u32 f2u(float f) { u32 u; memcpy(&u, &f, 4); return u; }

typedef struct Kat { const char* name; int kin, kout; void (*fn)(const float* in, float* out); } Kat;

static void k_pcg1(const float* i, float* o) { o[0] = u2f(Pcg1(f2u(i[0]))); }
static void k_pcg2(const float* i, float* o) { uint2 v = Pcg2(f2u(i[0]), f2u(i[1])); o[0] = u2f(v.x); o[1] = u2f(v.y); }
static void k_pcg4(const float* i, float* o)
{
    uint4 v = Pcg4(f2u(i[0]), f2u(i[1]), f2u(i[2]), f2u(i[3]));
    o[0] = u2f(v.x); o[1] = u2f(v.y); o[2] = u2f(v.z); o[3] = u2f(v.w);
}
static void k_tbn(const float* i, float* o)
{
    float3x3 m = NormalToTBN(f4_normalize3(f4_v(i[0], i[1], i[2], 0.0f)));
    o[0] = m.c0.x; o[1] = m.c0.y; o[2] = m.c0.z; o[3] = m.c1.x; o[4] = m.c1.y; o[5] = m.c1.z; o[6] = m.c2.x; o[7] = m.c2.y; o[8] = m.c2.z;
}
static void k_power(const float* i, float* o) { o[0] = PowerHeuristic(i[0], i[1]); }
static void k_disk(const float* i, float* o) { float2 d = MapSquareToDisk(f2_v(i[0], i[1])); o[0] = d.x; o[1] = d.y; }
static void k_bary(const float* i, float* o) { float4 w = SampleBaryCoord(f2_v(i[0], i[1])); o[0] = w.x; o[1] = w.y; o[2] = w.z; }
static void k_ngon(const float* i, float* o)
{
    float2 p = SampleNGon(f2_v(i[0], i[1]), f2u(i[2]) % 7u, 7, 0.3f); o[0] = p.x; o[1] = p.y;
    p = SamplePentagram(f2_v(i[0], i[1]), f2u(i[2])); o[2] = p.x; o[3] = p.y;
}
static void k_sphere(const float* i, float* o) { float4 d = SampleUnitSphere(f2_v(i[0], i[1])); o[0] = d.x; o[1] = d.y; o[2] = d.z; }
static void k_coshemi(const float* i, float* o) { float4 d = SampleCosineHemisphere(f2_v(i[0], i[1])); o[0] = d.x; o[1] = d.y; o[2] = d.z; }
static void k_ggx(const float* i, float* o)
{
    float alpha = BrdfAlpha(i[2]);
    float4 d = SampleGGXMicrofacet(f2_v(i[0], i[1]), alpha); o[0] = d.x; o[1] = d.y; o[2] = d.z;
    o[3] = GGXPdf(i[0], i[1], alpha);
    o[4] = D_GTR(i[0], alpha);
    o[5] = V_SmithCorrelated(i[0], i[1], alpha);
}
static void k_lightpdf(const float* i, float* o) { o[0] = LightPdf(i[0] + 0.01f, i[1], i[2] * 10.0f); }
static void k_gauss(const float* i, float* o) { float2 d = SampleGaussPixelFilter(f2_v(i[0], i[1]), 1.0f); o[0] = d.x; o[1] = d.y; }
static void k_freepath(const float* i, float* o) { o[0] = SampleFreePath(i[0], 1.0f + 100.0f * i[1]); }
static void k_fresnel(const float* i, float* o)
{
    o[0] = F_Dielectric(i[0], 1.0f, 1.0f + i[1]);
    o[1] = F_Dielectric(i[0], 1.0f + i[1], 1.0f);
    float4 f0 = F_0(f4_v(i[0], i[1], i[2], 0.0f), i[1]);
    o[2] = f0.x; o[3] = f0.y; o[4] = f0.z;
    o[5] = F_90(f0);
    o[6] = F_Schlick1(i[2], 1.0f, i[0]);
}
static void k_burley(const float* i, float* o) { o[0] = Fd_Burley(i[0], i[1], i[2], i[3]); }
static void k_transmit(const float* i, float* o)
{
    float4 t = AlbedoToTransmittance(f4_v(i[0], i[1], i[2], 0.0f), i[3], 0.01f + i[3]); o[0] = t.x; o[1] = t.y; o[2] = t.z;
}
static void k_gamma(const float* i, float* o)
{
    R8G8B8A8_t c; u32 bits = f2u(i[0]); memcpy(&c, &bits, 4);
    float4 v = GammaDecode_rgba8(c); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
}
static void k_cubeuv(const float* i, float* o)
{
    float2 uv; Cubeface f = Cubemap_CalcUv(f4_v(i[0] - 0.5f, i[1] - 0.5f, i[2] - 0.5f, 0.0f), &uv);
    o[0] = (float)f; o[1] = uv.x; o[2] = uv.y;
    float4 d = Cubemap_CalcDir(16, (i32)(f2u(i[3]) % 6u), i2_v((i32)(i[0] * 15.99f), (i32)(i[1] * 15.99f)), f2_tent(f2_v(i[1], i[2])));
    o[3] = d.x; o[4] = d.y; o[5] = d.z;
}
static void k_grid(const float* i, float* o)
{
    Grid g; Box3D b; b.lo = f4_v(-5.0f, -3.0f, -4.0f, 0.0f); b.hi = f4_v(5.0f, 3.0f, 6.0f, 0.0f);
    Grid_New(&g, b, 1.0f / 1.5f);
    float4 p = f4_v(i[0] * 14.0f - 7.0f, i[1] * 8.0f - 4.0f, i[2] * 12.0f - 5.0f, 0.0f);
    i32 idx = Grid_Index(&g, p);
    float4 c = Grid_Position(&g, idx);
    o[0] = (float)idx; o[1] = c.x; o[2] = c.y; o[3] = c.z; o[4] = (float)Grid_Len(&g);
}
static void k_sdtri(const float* i, float* o)
{
    o[0] = sdTriangle3D(f4_v(i[0], i[1], i[2], 0), f4_v(i[3], i[4], i[5], 0), f4_v(i[6], i[7], i[8], 0), f4_v(i[9] * 2.0f, i[10] * 2.0f, i[11] * 2.0f, 0));
    o[1] = sdTriangle2D(f2_v(i[0], i[1]), f2_v(i[3], i[4]), f2_v(i[6], i[7]), f2_v(i[9] * 1.5f, i[10] * 1.5f));
}
static void k_mie(const float* i, float* o) { o[0] = MiePhase(i[0] * 2.0f - 1.0f, 0.758f); o[1] = RayleighPhase(i[0] * 2.0f - 1.0f); }
/* 64 SG_Accumulate steps from one seed: the running fit of 5 lobes (sphgauss.c:19-56).
 * in: seed.xy, then the 5 lobe axes (xyz, sharpness) */
static void k_sgfit(const float* i, float* o)
{
    float4 axii[5], amps[5];
    memcpy(axii, i + 2, sizeof(axii));
    memset(amps, 0, sizeof(amps));
    uint4 s = { f2u(i[0]), f2u(i[1]), 7u, 9u };
    for (int k = 0; k < 64; ++k)
    {
        s = Pcg4(s.x, s.y, s.z, s.w);
        float2 Xi = { (float)(s.x >> 8) * (1.0f / 16777216.0f), (float)(s.y >> 8) * (1.0f / 16777216.0f) };
        float4 dir = SampleUnitHemisphere(Xi);
        float4 rad = f4_v((float)(s.z >> 8) * (1.0f / 16777216.0f), (float)(s.w >> 8) * (1.0f / 16777216.0f), 0.5f, 0.0f);
        SG_Accumulate(1.0f / (1.0f + k), dir, rad, axii, amps, 5);
    }
    for (int k = 0; k < 5; ++k) { o[4 * k] = amps[k].x; o[4 * k + 1] = amps[k].y; o[4 * k + 2] = amps[k].z; o[4 * k + 3] = amps[k].w; }
}
void oracle_sg_axii(float* out20)       /* SG_Generate(SGDist_Hemi), what LmPack_Pack stores (lightmap.c:1087) */
{
    float4 axii[5];
    SG_Generate(axii, 5, SGDist_Hemi);
    memcpy(out20, axii, sizeof(axii));
}
/* Dist1D over 8 bins (dist1d.c:75-118). in: cdf[9], pdf[8], u -> sampled bin, its pdf */
static void k_dist1d(const float* i, float* o)
{
    Dist1D d; Dist1D_New(&d, 8);
    memcpy(d.cdf, i, 9 * sizeof(float));
    memcpy(d.pdf, i + 9, 8 * sizeof(float));
    d.integral = 1.0f;
    i32 s = Dist1D_SampleD(&d, i[17]);
    o[0] = (float)s; o[1] = Dist1D_PdfD(&d, s);
    Dist1D_Del(&d);
}

static const Kat kKats[] =
{
    { "pcg1", 1, 1, k_pcg1 }, { "pcg2", 2, 2, k_pcg2 }, { "pcg4", 4, 4, k_pcg4 },
    { "tbn", 3, 9, k_tbn }, { "power", 2, 1, k_power }, { "disk", 2, 2, k_disk }, { "bary", 2, 3, k_bary },
    { "ngon", 3, 4, k_ngon }, { "sphere", 2, 3, k_sphere }, { "coshemi", 2, 3, k_coshemi }, { "ggx", 3, 6, k_ggx },
    { "lightpdf", 3, 1, k_lightpdf }, { "gauss", 2, 2, k_gauss }, { "freepath", 2, 1, k_freepath },
    { "fresnel", 3, 7, k_fresnel }, { "burley", 4, 1, k_burley }, { "transmit", 4, 3, k_transmit },
    { "gamma", 1, 4, k_gamma }, { "cubeuv", 4, 6, k_cubeuv }, { "grid", 3, 5, k_grid }, { "sdtri", 12, 2, k_sdtri },
    { "mie", 1, 2, k_mie }, { "sgfit", 22, 20, k_sgfit }, { "dist1d", 18, 2, k_dist1d },
};

int oracle_kat_count(void) { return (int)(sizeof(kKats) / sizeof(kKats[0])); }
const char* oracle_kat_name(int k, int* kin, int* kout) { *kin = kKats[k].kin; *kout = kKats[k].kout; return kKats[k].name; }
void oracle_kat_run(int k, const float* in, int n, float* out)
{
    for (int r = 0; r < n; ++r) kKats[k].fn(in + (size_t)r * kKats[k].kin, out + (size_t)r * kKats[k].kout);
}
