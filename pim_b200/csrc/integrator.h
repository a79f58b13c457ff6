// integrator.h — the per-path arithmetic of Pt_TraceRay
// (src/rendering/path_tracer.c:2306-2416) cut into two stages around its two
// BVH queries per bounce, so that a wavefront can run them as separate kernels
// (extend -> stageA -> shadow/MIS ray -> stageB) while a one-thread-per-path
// loop (lightmap bake, Pt_TraceRayN, the wavefront's tail) runs the same code
// back to back. RNG words are drawn in exactly the reference's order.
#pragma once

#include "traverse.h"

namespace pt {

struct DMaterial        // src/rendering/material.h:22-31, texture ids resolved to table indices
{
    int32_t albedo, rome, normal;
    uint32_t flags;
    float mfp[4];
    float ior;
    float bumpiness;
    float pad[2];
};

struct DTexture
{
    const uint32_t* texels;   // RGBA8 or RG16 packed in 32 bits
    int32_t w, h;
};

struct MediaDesc        // the fields of PtMediaDesc the integrator reads (path_tracer.c:91-111)
{
    V3 constantMu, noiseMu;
    float rcpMajorant, absorption;
    int32_t noiseOctaves;
    float noiseGain, noiseLacunarity, noiseFreq, noiseScale, noiseHeight, noiseRange;
    float phaseDirA, phaseDirB, phaseBlend;
};

// media_desc_new + media_desc_update (path_tracer.c:1944-1987)
PT_HD MediaDesc MediaDesc_Default()
{
    MediaDesc d;
    const float mfp = 40.0f * 1e3f;
    // f4_lerpsv(mfp*0.5, mfp*2, color=0.5)
    float m = mfp * 0.5f + 0.5f * (mfp * 2.0f - mfp * 0.5f);
    d.constantMu = v3s(1.0f / m);
    d.noiseMu = v3s(1.0f / m);
    d.absorption = 0.1f;
    d.noiseOctaves = 1;
    d.noiseGain = 0.9f;
    d.noiseLacunarity = 2.0666f;
    d.noiseFreq = 1.0f;
    d.noiseHeight = 20.0f;
    d.noiseScale = 1.0f;
    d.phaseDirA = 0.0f; d.phaseDirB = 0.0f; d.phaseBlend = 0.5f;
    float sum = 0.0f, amplitude = 1.0f;
    for (int i = 0; i < d.noiseOctaves; ++i) { sum += amplitude; amplitude *= d.noiseGain; }
    d.noiseRange = sum * d.noiseScale * 1.5f;
    float a = 1.0f + d.absorption;
    V3 maj = (d.constantMu * a + d.noiseMu * a) * 2.0f;     // CalcMajorant :2190-2196
    d.rcpMajorant = 1.0f / f1_max(maj.x, f1_max(maj.y, maj.z));
    return d;
}

struct DScene           // device view of PtScene (path_tracer.c:122-166)
{
    Bvh bvh;
    const F4* positions;
    const F4* normals;
    const V2* uvs;
    const int32_t* matIds;
    const int32_t* vertToEmit;
    const int32_t* emitToVert;
    const DMaterial* materials;
    const DTexture* textures;
    const float* sky;       // 6 * skySize^2 float3, or null
    int32_t skySize;
    const float* lut;       // lutW*lutH float2
    int32_t lutW, lutH;
    Grid grid;
    // Dist1D per cell, dense: pdf[cell*E+i], cdf[cell*(E+1)+i], live[cell*E+i]
    float* distPdf;
    float* distCdf;
    uint32_t* distLive;
    int32_t* distLength;    // E or 0 per cell
    uint32_t* distSum;      // Dist1D.sum per cell
    int32_t vertCount, matCount, emissiveCount;
    MediaDesc media;
};

struct RayHit           // PtRayHit (path_tracer.h:24-31)
{
    float w, u, v, t;
    V3 normal;
    int32_t type;
    int32_t iVert;
    uint32_t flags;
};

struct SurfHit          // PtSurfHit (path_tracer.c:58-72)
{
    V3 P, M, N;
    V3 albedo, emission;
    float roughness, occlusion, metallic, ior;
    uint32_t flags;
    int32_t type;
};

struct Scatter { V3 pos, dir, attenuation, luminance; float pdf; float evalW; };

struct Path
{
    Rng rng;
    V3 ro, rd;
    V3 atten, lum;
    V3 albedo, normal;
    float weight;
    int32_t bounce;
    uint32_t prevFlags;
};

enum : int { Stage_Done = 0, Stage_Ray2 = 1, Stage_NoRay2 = 2, Stage_Next = 3 };
enum : int { Pend_None = 0, Pend_Light = 1, Pend_Bsdf = 2 };

struct Pending
{
    int32_t kind;
    V3 ro2, rd2;
    float tFar2;
    float a0, a1, a2, a3;   // light: selectPdf, distSq, -, - | bsdf: attenuation.xyz, brdfPdf
    int32_t iVertLight;
};

PT_D void atomicAddU32(uint32_t* p, uint32_t v)
{
#if defined(__CUDA_ARCH__)
    atomicAdd(p, v);
#else
    *p += v;
#endif
}

// ------------------------------------------------------------ ray queries

PT_D RayHit finishHit(const DScene& S, V3 rd, const Hit& h)
{
    // pt_intersect_local (path_tracer.c:1421-1464)
    RayHit hit;
    hit.w = hit.u = hit.v = 0.0f; hit.t = -1.0f;
    hit.iVert = -1; hit.flags = 0; hit.type = Hit_Nothing;
    hit.normal = h.ng;
    if (h.prim < 0 || h.t <= 0.0f) return hit;
    hit.type = Hit_Triangle;
    if (dot(hit.normal, rd) > 0.0f)
    {
        hit.normal = -hit.normal;
        hit.type = Hit_Backface;
    }
    hit.normal = normalize(hit.normal);
    float u = f1_sat(h.u), v = f1_sat(h.v);
    hit.w = f1_sat(1.0f - (u + v));
    hit.u = u; hit.v = v; hit.t = h.t;
    hit.iVert = h.prim * 3;
    hit.flags = S.materials[S.matIds[hit.iVert]].flags;
    return hit;
}

template <class Stack>
PT_D RayHit intersectLocal(const DScene& S, V3 ro, V3 rd, float tNear, float tFar, Stack& st)
{
    Hit h;
    traverse<false, false>(S.bvh, ro, rd, tNear, tFar, h, st, (TravCounters*)nullptr);
    return finishHit(S, rd, h);
}

// Same query for the places the shading stages need one in the middle of a bounce (glass
// thickness, light sample from inside the medium): rare, so not inlined into the stage.
template <class Stack>
PT_D_RARE RayHit intersectLocalRare(const DScene& S, V3 ro, V3 rd, float tNear, float tFar, Stack& st)
{
    return intersectLocal(S, ro, rd, tNear, tFar, st);
}


// ------------------------------------------------------------ attributes

PT_D V3 ld3(const F4* p, int i)
{
    U4 a = ldg128(p + i);
    return v3(u2f(a.x), u2f(a.y), u2f(a.z));
}
PT_D V3 blend(V3 a, V3 b, V3 c, float w, float u, float v) { return (a * w + b * u) + c * v; }

PT_D V3 GetPosition(const DScene& S, const RayHit& h)
{
    return blend(ld3(S.positions, h.iVert), ld3(S.positions, h.iVert + 1), ld3(S.positions, h.iVert + 2), h.w, h.u, h.v);
}
PT_D V3 GetNormal(const DScene& S, const RayHit& h)     // :1192-1204
{
    V3 N = blend(ld3(S.normals, h.iVert), ld3(S.normals, h.iVert + 1), ld3(S.normals, h.iVert + 2), h.w, h.u, h.v);
    N = (dot(h.normal, N) > 0.0f) ? N : -N;
    return normalize(N);
}
PT_D V2 GetUV(const DScene& S, const RayHit& h)
{
    V2 a = S.uvs[h.iVert], b = S.uvs[h.iVert + 1], c = S.uvs[h.iVert + 2];
    return v2((a.x * h.w + b.x * h.u) + c.x * h.v, (a.y * h.w + b.y * h.u) + c.y * h.v);
}
PT_D float GetArea(const DScene& S, int iVert)
{
    return TriArea3D(ld3(S.positions, iVert), ld3(S.positions, iVert + 1), ld3(S.positions, iVert + 2));
}

PT_D V3 GetSky(const DScene& S, V3 rd)                   // :1280-1291, cubemap.h:115-127
{
    if (!S.sky) return v3s(0.0f);
    V2 uv;
    int face = Cubemap_CalcUv(rd, uv);
    return UvBilinearClamp_f3(S.sky + (size_t)face * S.skySize * S.skySize * 3, S.skySize, S.skySize, uv);
}

PT_D F4 sampleTex(const DScene& S, int id, V2 uv, F4 dflt)
{
    if (id < 0) return dflt;
    const DTexture t = S.textures[id];
    return UvBilinearWrap_c32(t.texels, t.w, t.h, uv);
}

PT_D V3 GetEmission(const DScene& S, V3 rd, const RayHit& hit)   // :1293-1326
{
    if (hit.flags & MatFlag_Sky) return GetSky(S, rd);
    const DMaterial& mat = S.materials[S.matIds[hit.iVert]];
    V2 uv = GetUV(S, hit);
    F4 albedo = sampleTex(S, mat.albedo, uv, f4(1, 1, 1, 1));
    float e = (mat.rome >= 0) ? sampleTex(S, mat.rome, uv, f4(0, 0, 0, 0)).w : 0.0f;
    return v3(albedo) * ((e * e) * kEmissionScale);
}

PT_D V3 FixShadingNormal(V3 M, V3 N) { return (dot(M, N) > 0.0f) ? N : reflect(N, M); }   // :1356-1361

PT_D SurfHit GetSurface(const DScene& S, V3 rd, const RayHit& hit)   // :1377-1419
{
    SurfHit surf;
    surf.type = hit.type;
    const DMaterial& mat = S.materials[S.matIds[hit.iVert]];
    surf.flags = mat.flags;
    surf.ior = mat.ior;
    V2 uv = GetUV(S, hit);
    surf.M = GetNormal(S, hit);
    surf.N = surf.M;
    surf.P = GetPosition(S, hit);
    surf.P = surf.P + surf.M * (0.01f * kMilli);
    if (hit.flags & MatFlag_Sky)
    {
        surf.albedo = v3s(0.0f);
        surf.roughness = 1.0f;
        surf.occlusion = 0.0f;
        surf.metallic = 0.0f;
        surf.emission = GetSky(S, rd);
        surf.M = -rd;
        surf.N = surf.M;
        surf.ior = 1.0f;
    }
    else
    {
        if (mat.normal >= 0)
        {
            const DTexture t = S.textures[mat.normal];
            V3 Nts = UvBilinearWrap_xy16(t.texels, t.w, t.h, uv);
            surf.N = FixShadingNormal(surf.N, TanToWorld(surf.N, Nts));
        }
        F4 albedo = sampleTex(S, mat.albedo, uv, f4(1, 1, 1, 1));
        F4 rome = sampleTex(S, mat.rome, uv, f4(0.5f, 1.0f, 0.0f, 0.0f));
        surf.albedo = v3(albedo);
        surf.emission = surf.albedo * ((rome.w * rome.w) * kEmissionScale);
        surf.roughness = rome.x;
        surf.occlusion = rome.y;
        surf.metallic = rome.z;
    }
    return surf;
}

// --------------------------------------------------------------- BSDFs

PT_D V3 GGXEnergyCompensation(const DScene& S, V3 f0, float NoV, float alpha)   // lighting.h:294-304
{
    V2 dv = UvBilinearClamp_f2(S.lut, S.lutW, S.lutH, v2(NoV, alpha));
    float t = (1.0f / f1_max(dv.y, kEpsilon)) - 1.0f;
    return f0 * t + v3s(1.0f);
}

// returns (attenuation.xyz, pdf in w)
PT_D F4 Eval_Diffuse(const SurfHit& surf, V3 I, V3 L)           // :1476-1497
{
    F4 result = f4(0, 0, 0, 0);
    V3 N = surf.N;
    float NoL = dot(N, L);
    float pdf = NoL * (1.0f / kPi);
    if (pdf > kEpsilon)
    {
        V3 V = -I;
        V3 H = normalize(V + L);
        float HoV = f1_sat(dot(H, V));
        float NoV = f1_sat(dot(N, V));
        V3 Fd = surf.albedo * Fd_Burley(NoL, NoV, HoV, surf.roughness);
        result = f4(Fd * NoL, pdf);
    }
    return result;
}

PT_D F4 Eval_Specular(const DScene& S, const SurfHit& surf, V3 I, V3 L)   // :1516-1548
{
    F4 result = f4(0, 0, 0, 0);
    V3 N = surf.N;
    float NoL = dot(N, L);
    if (NoL > kEpsilon)
    {
        float alpha = BrdfAlpha(surf.roughness);
        V3 V = -I;
        V3 H = normalize(V + L);
        float NoH = dot(N, H);
        float HoV = dot(H, V);
        float pdf = GGXPdf(NoH, HoV, alpha);
        if (pdf > kEpsilon)
        {
            float NoV = f1_sat(dot(N, V));
            float f = F_Dielectric(HoV, 1.0f, 1.5f);
            V3 f0 = F_0(surf.albedo, surf.metallic);
            V3 F = lerp(f0, v3s(F_90(f0)), f);
            float D = D_GTR(NoH, alpha);
            float G = V_SmithCorrelated(NoL, NoV, alpha);
            V3 Fr = F * (D * G);
            Fr = Fr * GGXEnergyCompensation(S, f0, NoV, alpha);
            result = f4(Fr * NoL, pdf);
        }
    }
    return result;
}

PT_D F4 Eval_Principled(const DScene& S, const SurfHit& surf, V3 I, V3 L)   // :1641-1668
{
    if (surf.flags & MatFlag_Refractive) return f4(0, 0, 0, 0);
    if (!(dot(surf.N, L) > kEpsilon)) return f4(0, 0, 0, 0);
    float amtA = f1_lerp(0.5f, 1.0f, surf.metallic);
    float amtB = 1.0f - amtA;
    F4 evalA = Eval_Specular(S, surf, I, L);
    if (amtB > kEpsilon)
    {
        F4 evalB = Eval_Diffuse(surf, I, L);
        evalA = f4_lerp(evalA, evalB, amtB);
    }
    return evalA;
}

PT_D Scatter Scatter_Diffuse(Rng& rng, const SurfHit& surf, V3 I)            // :1499-1514
{
    Scatter r;
    r.luminance = v3s(0.0f);
    V3 L = TanToWorld(surf.N, SampleCosineHemisphere(rng_f2(rng)));
    F4 ap = Eval_Diffuse(surf, I, L);
    r.dir = L; r.attenuation = v3(ap); r.pdf = ap.w; r.evalW = ap.w; r.pos = surf.P;
    return r;
}

PT_D Scatter Scatter_Specular(Rng& rng, const DScene& S, const SurfHit& surf, V3 I)   // :1550-1565
{
    Scatter r;
    r.luminance = v3s(0.0f);
    V3 m = SampleGGXMicrofacet(rng_f2(rng), BrdfAlpha(surf.roughness));
    m = TanToWorld(surf.N, m);
    m = FixShadingNormal(surf.M, m);
    r.dir = reflect(I, m);
    F4 ap = Eval_Specular(S, surf, I, r.dir);
    r.attenuation = v3(ap); r.pdf = ap.w; r.evalW = ap.w; r.pos = surf.P;
    return r;
}

template <class Stack>
PT_D Scatter Scatter_Refractive(Rng& rng, const DScene& S, const SurfHit& surf, V3 I, Stack& st)   // :1576-1638
{
    const float etaI = 1.000277f;
    const float etaT = f1_max(1.0f, surf.ior);
    const float alpha = BrdfAlpha(surf.roughness);
    V3 V = -I;
    V3 M = surf.M;
    V3 m = TanToWorld(surf.N, SampleGGXMicrofacet(rng_f2(rng), alpha));
    m = FixShadingNormal(M, m);
    bool entering = surf.type != Hit_Backface;
    float cosThetaI = f1_sat(f1_abs(dot(V, m)));
    float pdf = F_Dielectric(entering ? cosThetaI : -cosThetaI, etaI, etaT);
    V3 L;
    bool refr = false;
    if (rng_f32(rng) < pdf)
    {
        L = reflect(I, m);
    }
    else
    {
        float k = entering ? etaI / etaT : etaT / etaI;
        L = refract(I, m, k);
        pdf = 1.0f - pdf;
        refr = true;
    }
    V3 P = surf.P;
    if (dot(L, M) < 0.0f) P = P - M * (kMilli * 0.1f);
    Scatter r;
    r.luminance = v3s(0.0f);
    r.pos = P; r.dir = L; r.pdf = pdf; r.evalW = 0.0f;
    if (refr && entering)
    {
        RayHit hit = intersectLocalRare(S, P, L, 0.0f, kRcpEpsilon, st);
        float thickness = f1_max(hit.t, kEpsilon);
        if (hit.type == Hit_Nothing) thickness = kRcpEpsilon;
        V3 tr = AlbedoToTransmittance(surf.albedo, surf.roughness, thickness);
        r.attenuation = tr * pdf;
    }
    else
    {
        r.attenuation = v3s(pdf);
    }
    return r;
}

template <class Stack>
PT_D Scatter Scatter_Principled(Rng& rng, const DScene& S, const SurfHit& surf, V3 I, Stack& st)   // :1670-1707
{
    if (surf.flags & MatFlag_Refractive) return Scatter_Refractive(rng, S, surf, I, st);
    // Same arithmetic as the reference's two branches, arranged so that the two expensive lobe
    // evaluations run once for the whole warp: only the direction sampling diverges.
    Scatter sc;
    sc.luminance = v3s(0.0f);
    sc.pos = surf.P;
    const float amtSpec = f1_lerp(0.5f, 1.0f, surf.metallic);
    const float amtDiffuse = 1.0f - amtSpec;
    const bool spec = rng_f32(rng) < amtSpec;
    if (spec)
    {
        V3 m = SampleGGXMicrofacet(rng_f2(rng), BrdfAlpha(surf.roughness));      // Scatter_Specular :1550-1565
        m = TanToWorld(surf.N, m);
        m = FixShadingNormal(surf.M, m);
        sc.dir = reflect(I, m);
    }
    else
    {
        sc.dir = TanToWorld(surf.N, SampleCosineHemisphere(rng_f2(rng)));        // Scatter_Diffuse :1499-1514
    }
    const F4 evS = Eval_Specular(S, surf, I, sc.dir);
    F4 evD = f4(0, 0, 0, 0);
    if (!spec || amtDiffuse > kEpsilon) evD = Eval_Diffuse(surf, I, sc.dir);
    if (spec)
    {
        sc.attenuation = v3(evS); sc.pdf = evS.w;
        if (amtDiffuse > kEpsilon)
        {
            sc.attenuation = lerp(sc.attenuation, v3(evD), amtDiffuse);
            sc.pdf = f1_lerp(sc.pdf, evD.w, amtDiffuse);
        }
    }
    else
    {
        sc.attenuation = lerp(v3(evD), v3(evS), amtSpec);
        sc.pdf = f1_lerp(evD.w, evS.w, amtSpec);
    }
    sc.evalW = sc.pdf;
    return sc;
}

// ---------------------------------------------------------------- lights

PT_D void LightOnHit(const DScene& S, V3 ro, V3 lum3, int iVert)   // :1709-1733
{
    float lum = avglum(lum3);
    if (lum <= kEpsilon) return;
    float loglum = log2f(lum) - kLog2Epsilon;
    loglum = f1_clamp(loglum, 0.0f, 46.0f);
    uint32_t amt = (uint32_t)(loglum * (255.0f / 46.0f) + 0.5f);
    int iGrid = Grid_Index(S.grid, ro);
    int iEmit = S.vertToEmit[iVert];
    if (iEmit >= 0 && S.distLength[iGrid])
    {
        atomicAddU32(&S.distLive[(size_t)iGrid * S.emissiveCount + iEmit], amt);
    }
}

PT_D float Dist_PdfD(const DScene& S, int cell, int i)    // dist1d.c:113-116
{
    return S.distPdf[(size_t)cell * S.emissiveCount + i] / (float)S.emissiveCount;
}

PT_D bool LightSelect(Rng& rng, const DScene& S, V3 position, int& iVertOut, float& pdfOut)   // :1735-1764
{
    if (S.emissiveCount == 0)
    {
        iVertOut = -1; pdfOut = 0.0f;
        return false;
    }
    int iCell = Grid_Index(S.grid, position);
    if (!S.distLength[iCell]) return false;
    int iEmit = FindInterval(S.distCdf + (size_t)iCell * (S.emissiveCount + 1), S.emissiveCount + 1, rng_f32(rng));
    float pdf = Dist_PdfD(S, iCell, iEmit);
    iVertOut = S.emitToVert[iEmit];
    pdfOut = pdf;
    return pdf > kEpsilon;
}

PT_D float LightSelectPdf(const DScene& S, int iVert, V3 ro)   // :1766-1783
{
    float selectPdf = 1.0f;
    int iGrid = Grid_Index(S.grid, ro);
    int iEmit = S.vertToEmit[iVert];
    if (iEmit >= 0 && S.distLength[iGrid]) selectPdf = Dist_PdfD(S, iGrid, iEmit);
    return selectPdf;
}

// ----------------------------------------------------------------- media

PT_D float GradientCell3Dot(int x, int y, int z, uint32_t seed, V3 d)   // noise.h:16-25
{
    uint32_t i = Pcg4((uint32_t)x, (uint32_t)y, (uint32_t)z, seed).x;
    float gx = (i & (1u << 31)) ? 1.0f : -1.0f;
    float gy = (i & (1u << 30)) ? 1.0f : -1.0f;
    float gz = (i & (1u << 29)) ? 1.0f : -1.0f;
    return gx * d.x + gy * d.y + gz * d.z;
}
PT_D float unormstep(float t) { return ((t * -2.0f + 3.0f) * t) * t; }
PT_D float GradientNoise3(V3 p, uint32_t seed)            // noise.h:27-68
{
    V3 fl = v3(f1_floor(p.x), f1_floor(p.y), f1_floor(p.z));
    int ix = (int)fl.x, iy = (int)fl.y, iz = (int)fl.z;
    V3 f = p - fl;
    float c000 = GradientCell3Dot(ix, iy, iz, seed, f - v3(0, 0, 0));
    float c001 = GradientCell3Dot(ix, iy, iz + 1, seed, f - v3(0, 0, 1));
    float c010 = GradientCell3Dot(ix, iy + 1, iz, seed, f - v3(0, 1, 0));
    float c011 = GradientCell3Dot(ix, iy + 1, iz + 1, seed, f - v3(0, 1, 1));
    float c100 = GradientCell3Dot(ix + 1, iy, iz, seed, f - v3(1, 0, 0));
    float c101 = GradientCell3Dot(ix + 1, iy, iz + 1, seed, f - v3(1, 0, 1));
    float c110 = GradientCell3Dot(ix + 1, iy + 1, iz, seed, f - v3(1, 1, 0));
    float c111 = GradientCell3Dot(ix + 1, iy + 1, iz + 1, seed, f - v3(1, 1, 1));
    float ux = unormstep(f.x), uy = unormstep(f.y), uz = unormstep(f.z);
    float a = f1_lerp(c000, c001, uz), b = f1_lerp(c010, c011, uz);
    float c = f1_lerp(c100, c101, uz), d = f1_lerp(c110, c111, uz);
    return f1_lerp(f1_lerp(a, b, uy), f1_lerp(c, d, uy), ux);
}

struct Media { V3 scattering, extinction; float g0, g1, gBlend; };

PT_D Media Media_Sample(const MediaDesc& desc, V3 P)      // :2146-2181
{
    Media hit;
    hit.scattering = desc.constantMu;
    hit.g0 = desc.phaseDirA; hit.g1 = desc.phaseDirB; hit.gBlend = desc.phaseBlend;
    if (f1_abs(desc.noiseHeight - P.y) <= desc.noiseRange)
    {
        float sum = 0.0f, freq = 1.0f, ampl = 1.0f;
        V3 p = P * desc.noiseFreq;
        for (int i = 0; i < desc.noiseOctaves; ++i)     // FbmGradientNoise3, noise.h:70-83
        {
            sum += GradientNoise3(p * freq, 1u + (uint32_t)i + 1u) * ampl;
            freq *= desc.noiseLacunarity;
            ampl *= desc.noiseGain;
        }
        float height = desc.noiseHeight + desc.noiseScale * sum;
        float dist = f1_abs(height - P.y) / desc.noiseScale;
        float heightDensity = f1_sat(1.0f - dist);
        hit.scattering = hit.scattering + desc.noiseMu * heightDensity;
    }
    hit.extinction = hit.scattering * (1.0f + desc.absorption);
    return hit;
}

PT_D float CalcPhase(const Media& m, float cosTheta)      // :2198-2206
{
    return f1_lerp(MiePhase(cosTheta, m.g0), MiePhase(cosTheta, m.g1), m.gBlend);
}

PT_D V3 CalcTransmittance(Rng& rng, const DScene& S, V3 ro, V3 rd, float rayLen)   // :2223-2249
{
    const float rcpMaj = S.media.rcpMajorant;
    float t = 0.0f;
    V3 attenuation = v3s(1.0f);
    for (;;)
    {
        float dt = SampleFreePath(rng_f32(rng), rcpMaj);
        if ((t + dt) >= rayLen) break;
        Media media = Media_Sample(S.media, ro + rd * t);
        attenuation = attenuation * inv(media.extinction * rcpMaj);
        t += dt;
    }
    return attenuation;
}

// SampleLight (:1785-1827) with its closest-hit query done in place; used by the
// volumetric branch, where the wavefront does not split the bounce.
template <class Stack>
PT_D_RARE bool EvaluateLight(Rng& rng, const DScene& S, V3 P, V3& lightOut, V3& dirOut, Stack& st)   // :1921-1942
{
    int iVert; float selectPdf;
    if (!LightSelect(rng, S, P, iVert, selectPdf)) return false;
    V3 wuv = SampleBaryCoord(rng_f2(rng));
    V3 A = ld3(S.positions, iVert), B = ld3(S.positions, iVert + 1), C = ld3(S.positions, iVert + 2);
    V3 pt = blend(A, B, C, wuv.x, wuv.y, wuv.z);
    float area = TriArea3D(A, B, C);
    V3 rd = pt - P;
    float distSq = dot(rd, rd);
    float distance = sqrtf(f1_max(distSq, kEpsilonSq));
    rd = rd / distance;
    float pdf = 0.0f;
    V3 lum = v3s(0.0f);
    RayHit hit = intersectLocalRare(S, P, rd, 0.0f, distance + 0.01f * kMilli, st);
    if (hit.type != Hit_Nothing && hit.iVert == iVert)
    {
        float cosTheta = f1_abs(dot(rd, hit.normal));
        pdf = LightPdf(area, cosTheta, distSq);
        lum = GetEmission(S, rd, hit);
        if (sum3(lum) > kEpsilon) lum = lum * CalcTransmittance(rng, S, P, rd, hit.t);
    }
    if (pdf > kEpsilon)
    {
        lightOut = lum / (pdf * selectPdf);
        dirOut = rd;
        return true;
    }
    return false;
}

template <class Stack>
PT_D Scatter ScatterRay(Rng& rng, const DScene& S, V3 ro, V3 rd, float rayLen, Stack& st)   // :2251-2304
{
    Scatter result;
    result.pdf = 0.0f; result.evalW = 0.0f;
    result.luminance = v3s(0.0f);
    result.pos = ro; result.dir = rd;
    const float rcpMaj = S.media.rcpMajorant;
    float t = 0.0f;
    result.attenuation = v3s(1.0f);
    for (;;)
    {
        float dt = SampleFreePath(rng_f32(rng), rcpMaj);
        t += dt;
        if (t >= rayLen) break;
        V3 P = ro + rd * t;
        Media media = Media_Sample(S.media, P);
        result.attenuation = result.attenuation * inv(media.extinction * rcpMaj);
        float scatterProb = hmax(media.scattering) * rcpMaj;
        if (rng_f32(rng) < scatterProb)
        {
            V3 lum, L;
            if (EvaluateLight(rng, S, P, lum, L, st))
            {
                float ph = CalcPhase(media, dot(rd, L));
                lum = lum * (ph * dt);
                result.luminance = result.attenuation * lum;
            }
            result.pos = P;
            V3 dir; float w;
            do                                              // SamplePhaseDir :2208-2221
            {
                dir = SampleUnitSphere(rng_f2(rng));
                w = CalcPhase(media, dot(rd, dir));
            } while (rng_f32(rng) > w);
            result.dir = dir;
            result.pdf = w;
            float ph = CalcPhase(media, dot(rd, dir));
            result.attenuation = result.attenuation * ph;
            break;
        }
    }
    return result;
}

// ---------------------------------------------------------------- stages

PT_D void pathInit(Path& p, Rng rng, V3 ro, V3 rd)
{
    p.rng = rng; p.ro = ro; p.rd = rd;
    p.atten = v3s(1.0f); p.lum = v3s(0.0f);
    p.albedo = v3s(0.0f); p.normal = v3s(0.0f);
    p.weight = 0.0f; p.bounce = 0; p.prevFlags = 0;
}

// Russian roulette at the head of every bounce (:2321-2331); false = path ends
PT_D bool pathRoulette(Path& p)
{
    float pr = f1_sat(avglum(p.atten));
    if (rng_f32(p.rng) < pr)
    {
        p.atten = p.atten / pr;
        return true;
    }
    return false;
}

// After Stage_Next: advance the bounce counter and roll the roulette.
PT_D bool pathAdvance(Path& p)
{
    p.bounce++;
    if (p.bounce >= 666) return false;
    return pathRoulette(p);
}

PT_D void pathFinish(const Path& p, V3& color, V3& albedo, V3& normal)   // :2409-2414
{
    float s = 1.0f / f1_max(p.weight, kEpsilon);
    color = p.lum;
    albedo = p.albedo * s;
    normal = p.normal * s;
}

// Stage A: everything between the path ray's hit and EstimateDirect's ray. On Stage_Ray2 /
// Stage_NoRay2 `surf` holds the surface interaction stage B continues from.
template <class Stack>
PT_D int pathStageA(const DScene& S, Path& p, const RayHit& hit, Pending& pend, SurfHit& surf, Stack& st)
{
    pend.kind = Pend_None;
    if (hit.type == Hit_Nothing)
    {
        p.lum = p.lum + p.atten * GetSky(S, p.rd);
        return Stage_Done;
    }
    if (hit.type == Hit_Backface && !(hit.flags & MatFlag_Refractive)) return Stage_Done;

    {
        Scatter sc = ScatterRay(p.rng, S, p.ro, p.rd, hit.t, st);
        if (sc.pdf > kEpsilon)
        {
            p.lum = p.lum + p.atten * sc.luminance;
            p.atten = p.atten * (sc.attenuation / sc.pdf);
            V3 a = p.atten * (1.0f / kTau);
            float w = f1_sat(1.0f - avglum(a));
            p.weight += w;
            p.albedo = p.albedo + a * w;
            p.normal = p.normal + (-p.rd) * w;
            p.ro = sc.pos; p.rd = sc.dir;
            p.prevFlags = 0;
            return Stage_Next;
        }
        p.atten = p.atten * sc.attenuation;
    }

    surf = GetSurface(S, p.rd, hit);
    pend.ro2 = surf.P;
    if (p.bounce > 0) LightOnHit(S, p.ro, surf.emission, hit.iVert);
    if (p.bounce == 0 || (p.prevFlags & MatFlag_Refractive)) p.lum = p.lum + surf.emission * p.atten;
    if (hit.flags & MatFlag_Sky) return Stage_Done;

    // EstimateDirect up to its ray (:1849-1919)
    if (!(surf.flags & MatFlag_Refractive))
    {
        const float pRough = f1_lerp(0.05f, 0.95f, surf.roughness);
        const float pSmooth = 1.0f - pRough;
        if (rng_f32(p.rng) < pRough)
        {
            int iVert; float selectPdf;
            if (LightSelect(p.rng, S, surf.P, iVert, selectPdf))
            {
                if (hit.iVert != iVert)
                {
                    V3 wuv = SampleBaryCoord(rng_f2(p.rng));          // SampleLight :1794-1812
                    V3 A = ld3(S.positions, iVert), B = ld3(S.positions, iVert + 1), C = ld3(S.positions, iVert + 2);
                    V3 pt = blend(A, B, C, wuv.x, wuv.y, wuv.z);
                    V3 rd = pt - surf.P;
                    float distSq = dot(rd, rd);
                    float distance = sqrtf(f1_max(distSq, kEpsilonSq));
                    rd = rd / distance;
                    pend.kind = Pend_Light;
                    pend.ro2 = surf.P; pend.rd2 = rd; pend.tFar2 = distance + 0.01f * kMilli;
                    pend.a0 = selectPdf; pend.a1 = distSq; pend.a2 = 0.0f; pend.a3 = 0.0f;
                    pend.iVertLight = iVert;
                }
            }
        }
        else
        {
            Scatter sample = Scatter_Principled(p.rng, S, surf, p.rd, st);
            float brdfPdf = sample.pdf * pSmooth;
            if (brdfPdf > kEpsilon)
            {
                pend.kind = Pend_Bsdf;
                pend.ro2 = surf.P; pend.rd2 = sample.dir; pend.tFar2 = kRcpEpsilon;
                pend.a0 = sample.attenuation.x; pend.a1 = sample.attenuation.y; pend.a2 = sample.attenuation.z;
                pend.a3 = brdfPdf;
                pend.iVertLight = -1;
            }
        }
    }
    return pend.kind ? Stage_Ray2 : Stage_NoRay2;
}

// Stage B: the rest of EstimateDirect, then the BSDF scatter that sets up the next bounce.
template <class Stack>
PT_D int pathStageB(const DScene& S, Path& p, const SurfHit& surf, const Pending& pend, const RayHit& hit2, Stack& st)
{
    const float pRough = f1_lerp(0.05f, 0.95f, surf.roughness);
    const float pSmooth = 1.0f - pRough;
    V3 direct = v3s(0.0f);
    if (pend.kind == Pend_Light)
    {
        float pdf = 0.0f;
        V3 Li = v3s(0.0f);
        if (hit2.type != Hit_Nothing && hit2.iVert == pend.iVertLight)
        {
            float cosTheta = f1_abs(dot(pend.rd2, hit2.normal));
            pdf = LightPdf(GetArea(S, pend.iVertLight), cosTheta, pend.a1);
            Li = GetEmission(S, pend.rd2, hit2);
            if (sum3(Li) > kEpsilon) Li = Li * CalcTransmittance(p.rng, S, pend.ro2, pend.rd2, hit2.t);
        }
        float lightPdf = pdf * pend.a0 * pRough;
        if (lightPdf > kEpsilon && hmax(Li) > kEpsilon)
        {
            F4 brdf = Eval_Principled(S, surf, p.rd, pend.rd2);
            Li = Li * v3(brdf);
            float brdfPdf = brdf.w * pSmooth;
            if (brdfPdf > kEpsilon)
            {
                Li = Li * (PowerHeuristic(lightPdf, brdfPdf) / lightPdf);
                direct = direct + Li;
            }
        }
    }
    else if (pend.kind == Pend_Bsdf)
    {
        float lightPdf = 0.0f;                               // LightEvalPdf :1829-1847
        if (hit2.type != Hit_Nothing)
        {
            float cosTheta = f1_abs(dot(pend.rd2, hit2.normal));
            float area = GetArea(S, hit2.iVert);
            float distSq = f1_max(kEpsilon, hit2.t * hit2.t);
            lightPdf = LightPdf(area, cosTheta, distSq);
        }
        lightPdf *= pRough;
        if (lightPdf > kEpsilon)
        {
            lightPdf *= LightSelectPdf(S, hit2.iVert, pend.ro2);
            V3 Li = GetEmission(S, pend.rd2, hit2) * v3(pend.a0, pend.a1, pend.a2);
            if (hmax(Li) > kEpsilon)
            {
                Li = Li * (PowerHeuristic(pend.a3, lightPdf) / pend.a3);
                Li = Li * CalcTransmittance(p.rng, S, pend.ro2, pend.rd2, hit2.t);
                direct = direct + Li;
            }
        }
    }
    p.lum = p.lum + direct * p.atten;

    Scatter sc = Scatter_Principled(p.rng, S, surf, p.rd, st);
    if (!(sc.pdf > kEpsilon)) return Stage_Done;
    p.ro = sc.pos;
    p.rd = sc.dir;
    p.atten = p.atten * (sc.attenuation / sc.pdf);
    p.prevFlags = surf.flags;
    {
        V3 a = p.atten * (1.0f / kPi);
        float w = f1_sat(1.0f - avglum(a));
        p.weight += w;
        p.albedo = p.albedo + surf.albedo * w;
        p.normal = p.normal + surf.N * w;
    }
    return Stage_Next;
}

// Whole path in one thread: Pt_TraceRay (:2306-2416)
template <class Stack>
PT_D void traceRay(const DScene& S, Rng& rng, V3 ro, V3 rd, V3& color, V3& albedo, V3& normal, Stack& st,
                   uint32_t* rayCount)
{
    Path p;
    pathInit(p, rng, ro, rd);
    bool alive = pathRoulette(p);
    while (alive)
    {
        RayHit hit = intersectLocal(S, p.ro, p.rd, 0.0f, kRcpEpsilon, st);
        if (rayCount) (*rayCount)++;
        Pending pend;
        SurfHit surf;
        int r = pathStageA(S, p, hit, pend, surf, st);
        if (r == Stage_Done) break;
        if (r != Stage_Next)
        {
            RayHit hit2;
            hit2.type = Hit_Nothing; hit2.iVert = -1; hit2.t = -1.0f; hit2.flags = 0; hit2.normal = v3s(0.0f);
            hit2.w = hit2.u = hit2.v = 0.0f;
            if (r == Stage_Ray2)
            {
                hit2 = intersectLocal(S, pend.ro2, pend.rd2, 0.0f, pend.tFar2, st);
                if (rayCount) (*rayCount)++;
            }
            r = pathStageB(S, p, surf, pend, hit2, st);
            if (r == Stage_Done) break;
        }
        alive = pathAdvance(p);
    }
    rng = p.rng;
    pathFinish(p, color, albedo, normal);
}

// ------------------------------------------------------------------ camera

struct CamParams        // TraceFn's per-frame constants (:2527-2536)
{
    V3 eye, right, up, fwd;
    V2 slope;
    int32_t width, height;
    // PtDofInfo (:33-42)
    float aperture, focalLength;
    int32_t bladeCount;
    float bladeRot, focalPlaneCurvature;
};

// TraceFn's ray setup for pixel i (:2541-2547) + CalculateDof (:2418-2452)
PT_D void cameraRay(const CamParams& C, Rng& rng, int i, V3& ro, V3& rd)
{
    const float rcpW = 1.0f / (float)C.width, rcpH = 1.0f / (float)C.height;
    const int cx = i % C.width, cy = i / C.width;
    const V2 baseUv = v2((cx + 0.5f) * rcpW, (cy + 0.5f) * rcpH);
    const V2 aa = SampleGaussPixelFilter(rng_f2(rng), 1.0f);
    const V2 rayUv = v2(baseUv.x + aa.x * rcpW, baseUv.y + aa.y * rcpH);
    V2 coord = v2((2.0f * rayUv.x - 1.0f) * C.slope.x, (2.0f * rayUv.y - 1.0f) * C.slope.y);   // proj_dir, frustum.h:33-47
    V3 dir = normalize((C.fwd + C.right * coord.x) + C.up * coord.y);
    ro = C.eye; rd = dir;

    uint32_t side = rng_u32(rng);
    V2 Xi = rng_f2(rng);
    V2 offset = (C.bladeCount == 666) ? SamplePentagram(Xi, side)
                                      : SampleNGon(Xi, side, (uint32_t)C.bladeCount, C.bladeRot);
    offset.x *= C.aperture; offset.y *= C.aperture;
    float t = f1_lerp(C.focalLength / dot(rd, C.fwd), C.focalLength, C.focalPlaneCurvature);
    V3 focusPos = ro + rd * t;
    V3 aperturePos = ro + (C.right * offset.x + C.up * offset.y);
    ro = aperturePos;
    rd = normalize(focusPos - aperturePos);
}

} // namespace pt
