// TEST INFRASTRUCTURE — steps the library's device headers (pt_core.h, bvh.h,
// traverse.h, integrator.h) on the CPU, one "thread" at a time, so the CPU test
// suite can check the kernels' arithmetic against the oracle without a GPU.
// Nothing here is linked into libpimpt.so; the shipped path has no CPU route.
#include "../../pim_b200/csrc/integrator.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace pt;

struct HsTexDesc { const uint32_t* texels; int32_t w, h; };
struct HsMaterial { int32_t albedo, rome, normal; uint32_t flags; float mfp[4]; float ior, bumpiness; };

struct HostScene
{
    DScene S;
    std::vector<WideNode> nodes;
    std::vector<PackedTri> tris;
    std::vector<F4> pos, nrm;
    std::vector<V2> uvs;
    std::vector<int32_t> matIds, vertToEmit, emitToVert, distLength;
    std::vector<DMaterial> mats;
    std::vector<DTexture> tex;
    std::vector<std::vector<uint32_t>> texData;
    std::vector<float> pdf, cdf, lut, sky;
    std::vector<uint32_t> live, distSum;
    uint64_t nodeFetches = 0, triFetches = 0, rays = 0;
};

struct LocalStack : TravStack<64>
{
    unsigned long long mem[64];
    LocalStack() { init(mem, 1); }
};

// Host rendition of the GPU builder: Morton sort -> Karras radix tree -> bottom-up
// boxes -> greedy 8-wide collapse. Same PT_HD helpers as the kernels use.
static void buildBvh(HostScene& H)
{
    const int n = (int)H.pos.size() / 3;
    H.nodes.clear(); H.tris.clear();
    if (n == 0) { H.S.bvh = Bvh{ nullptr, nullptr, 0, 0 }; return; }
    V3 lo = v3s(3e38f), hi = v3s(-3e38f);
    std::vector<V3> cent(n);
    for (int t = 0; t < n; ++t)
    {
        V3 a = v3(H.pos[3 * t]), b = v3(H.pos[3 * t + 1]), c = v3(H.pos[3 * t + 2]);
        V3 l = v3(f1_min(a.x, f1_min(b.x, c.x)), f1_min(a.y, f1_min(b.y, c.y)), f1_min(a.z, f1_min(b.z, c.z)));
        V3 h = v3(f1_max(a.x, f1_max(b.x, c.x)), f1_max(a.y, f1_max(b.y, c.y)), f1_max(a.z, f1_max(b.z, c.z)));
        cent[t] = (l + h) * 0.5f;
        lo = v3(f1_min(lo.x, cent[t].x), f1_min(lo.y, cent[t].y), f1_min(lo.z, cent[t].z));
        hi = v3(f1_max(hi.x, cent[t].x), f1_max(hi.y, cent[t].y), f1_max(hi.z, cent[t].z));
    }
    V3 ext = hi - lo;
    V3 invExt = v3(ext.x > 0 ? 1.0f / ext.x : 0.0f, ext.y > 0 ? 1.0f / ext.y : 0.0f, ext.z > 0 ? 1.0f / ext.z : 0.0f);
    std::vector<std::pair<uint64_t, int>> kv(n);
    for (int t = 0; t < n; ++t) kv[t] = { morton63(cent[t], lo, invExt), t };
    std::stable_sort(kv.begin(), kv.end(), [](auto& a, auto& b) { return a.first < b.first; });
    std::vector<uint64_t> keys(n);
    for (int i = 0; i < n; ++i) keys[i] = kv[i].first;

    const int total = 2 * n - 1;
    std::vector<F4> blo(total), bhi(total);
    std::vector<int32_t> first(total), last(total), parent(total, -1);
    for (int k = 0; k < n; ++k)
    {
        int t = kv[k].second, id = n - 1 + k;
        V3 a = v3(H.pos[3 * t]), b = v3(H.pos[3 * t + 1]), c = v3(H.pos[3 * t + 2]);
        blo[id] = f4(f1_min(a.x, f1_min(b.x, c.x)), f1_min(a.y, f1_min(b.y, c.y)), f1_min(a.z, f1_min(b.z, c.z)), 0);
        bhi[id] = f4(f1_max(a.x, f1_max(b.x, c.x)), f1_max(a.y, f1_max(b.y, c.y)), f1_max(a.z, f1_max(b.z, c.z)), 0);
        first[id] = last[id] = k;
    }
    for (int i = 0; i < n - 1; ++i)
    {
        int f, l, s;
        lbvhRange(keys.data(), n, i, f, l, s);
        int left = (s == f) ? (n - 1 + s) : s;
        int right = (s + 1 == l) ? (n - 1 + s + 1) : (s + 1);
        first[i] = f; last[i] = l;
        blo[i].w = u2f((uint32_t)left); bhi[i].w = u2f((uint32_t)right);
        parent[left] = i; parent[right] = i;
    }
    // bottom-up boxes (the kernel does this with one atomic flag per internal node)
    std::vector<int> visits(n > 1 ? n - 1 : 0, 0);
    for (int k = 0; k < n; ++k)
    {
        int node = parent[n - 1 + k];
        while (node >= 0)
        {
            if (++visits[node] < 2) break;
            int l = (int)f2u(blo[node].w), r = (int)f2u(bhi[node].w);
            blo[node].x = f1_min(blo[l].x, blo[r].x); blo[node].y = f1_min(blo[l].y, blo[r].y); blo[node].z = f1_min(blo[l].z, blo[r].z);
            bhi[node].x = f1_max(bhi[l].x, bhi[r].x); bhi[node].y = f1_max(bhi[l].y, bhi[r].y); bhi[node].z = f1_max(bhi[l].z, bhi[r].z);
            node = parent[node];
        }
    }
    BinTree T{ blo.data(), bhi.data(), first.data(), last.data(), n };
    int root = (n == 1) ? 0 + (n - 1) : 0;
    // breadth-first collapse
    struct Work { int bin; uint32_t wide; };
    std::vector<Work> queue{ { root, 0 } };
    H.nodes.resize(1);
    size_t head = 0;
    while (head < queue.size())
    {
        Work w = queue[head++];
        int kids[8];
        int nk = collapseChildren(T, w.bin, kids);
        uint32_t nInt = 0, nTri = 0;
        for (int i = 0; i < nk; ++i)
        {
            if (binIsLeafCluster(T, kids[i])) nTri += (uint32_t)(last[kids[i]] - first[kids[i]] + 1);
            else nInt++;
        }
        uint32_t childBase = (uint32_t)H.nodes.size();
        H.nodes.resize(H.nodes.size() + nInt);
        uint32_t triBase = (uint32_t)H.tris.size();
        H.tris.resize(H.tris.size() + nTri);
        encodeWideNode(H.nodes[w.wide], T, kids, nk, childBase, triBase);
        uint32_t ci = 0, ti = 0;
        for (int i = 0; i < nk; ++i)
        {
            if (binIsLeafCluster(T, kids[i]))
            {
                for (int k = first[kids[i]]; k <= last[kids[i]]; ++k)
                {
                    int t = kv[k].second;
                    V3 a = v3(H.pos[3 * t]), b = v3(H.pos[3 * t + 1]), c = v3(H.pos[3 * t + 2]);
                    PackedTri pt;
                    pt.v0x = a.x; pt.v0y = a.y; pt.v0z = a.z; pt.prim = (uint32_t)t;
                    pt.e1x = b.x - a.x; pt.e1y = b.y - a.y; pt.e1z = b.z - a.z; pt.pad0 = 0;
                    pt.e2x = c.x - a.x; pt.e2y = c.y - a.y; pt.e2z = c.z - a.z; pt.pad1 = 0;
                    H.tris[triBase + ti++] = pt;
                }
            }
            else
            {
                queue.push_back({ kids[i], childBase + ci++ });
            }
        }
    }
    H.S.bvh = Bvh{ H.nodes.data(), H.tris.data(), (int32_t)H.nodes.size(), (int32_t)H.tris.size() };
}

extern "C" {

void* hs_new(const float* pos4, const float* nrm4, const float* uv2, const int32_t* matIds, int vertCount,
             const HsMaterial* mats, int matCount, const HsTexDesc* tex, int texCount,
             const int32_t* vertToEmit, const int32_t* emitToVert, int emissiveCount,
             const float* gridLo, const float* gridHi, const int32_t* gridSize, float cellsPerMeter,
             const float* pdf, const float* cdf, const int32_t* distLength,
             const float* lut, int lutW, int lutH, const float* sky, int skySize)
{
    HostScene* H = new HostScene();
    H->pos.resize(vertCount); H->nrm.resize(vertCount); H->uvs.resize(vertCount);
    memcpy(H->pos.data(), pos4, sizeof(F4) * vertCount);
    memcpy(H->nrm.data(), nrm4, sizeof(F4) * vertCount);
    memcpy(H->uvs.data(), uv2, sizeof(V2) * vertCount);
    H->matIds.assign(matIds, matIds + vertCount);
    H->vertToEmit.assign(vertToEmit, vertToEmit + vertCount);
    H->emitToVert.assign(emitToVert, emitToVert + emissiveCount);
    H->texData.resize(texCount); H->tex.resize(texCount);
    for (int i = 0; i < texCount; ++i)
    {
        H->texData[i].assign(tex[i].texels, tex[i].texels + (size_t)tex[i].w * tex[i].h);
        H->tex[i] = DTexture{ H->texData[i].data(), tex[i].w, tex[i].h };
    }
    H->mats.resize(matCount);
    for (int i = 0; i < matCount; ++i)
    {
        DMaterial m; memset(&m, 0, sizeof(m));
        m.albedo = mats[i].albedo; m.rome = mats[i].rome; m.normal = mats[i].normal;
        m.flags = mats[i].flags; memcpy(m.mfp, mats[i].mfp, 16); m.ior = mats[i].ior; m.bumpiness = mats[i].bumpiness;
        H->mats[i] = m;
    }
    DScene& S = H->S;
    memset(&S, 0, sizeof(S));
    S.grid.lo = v3(gridLo[0], gridLo[1], gridLo[2]); S.grid.hi = v3(gridHi[0], gridHi[1], gridHi[2]);
    S.grid.sx = gridSize[0]; S.grid.sy = gridSize[1]; S.grid.sz = gridSize[2]; S.grid.cellsPerMeter = cellsPerMeter;
    int cells = Grid_Len(S.grid);
    H->pdf.assign(pdf, pdf + (size_t)cells * emissiveCount);
    H->cdf.assign(cdf, cdf + (size_t)cells * (emissiveCount + 1));
    H->distLength.assign(distLength, distLength + cells);
    H->live.assign((size_t)cells * emissiveCount, 0u);
    H->distSum.assign(cells, 0u);
    H->lut.assign(lut, lut + (size_t)lutW * lutH * 2);
    if (sky) H->sky.assign(sky, sky + (size_t)6 * skySize * skySize * 3);
    S.positions = H->pos.data(); S.normals = H->nrm.data(); S.uvs = H->uvs.data();
    S.matIds = H->matIds.data(); S.vertToEmit = H->vertToEmit.data(); S.emitToVert = H->emitToVert.data();
    S.materials = H->mats.data(); S.textures = H->tex.data();
    S.sky = sky ? H->sky.data() : nullptr; S.skySize = skySize;
    S.lut = H->lut.data(); S.lutW = lutW; S.lutH = lutH;
    S.distPdf = H->pdf.data(); S.distCdf = H->cdf.data(); S.distLive = H->live.data();
    S.distLength = H->distLength.data(); S.distSum = H->distSum.data();
    S.vertCount = vertCount; S.matCount = matCount; S.emissiveCount = emissiveCount;
    S.media = MediaDesc_Default();
    buildBvh(*H);
    return H;
}

void hs_del(void* h) { delete (HostScene*)h; }

int hs_node_count(void* h) { return (int)((HostScene*)h)->nodes.size(); }

void hs_set_dists(void* h, const float* pdf, const float* cdf, const uint32_t* live)
{
    HostScene* H = (HostScene*)h;
    memcpy(H->pdf.data(), pdf, sizeof(float) * H->pdf.size());
    memcpy(H->cdf.data(), cdf, sizeof(float) * H->cdf.size());
    if (live) memcpy(H->live.data(), live, sizeof(uint32_t) * H->live.size());
}
void hs_get_live(void* h, uint32_t* live)
{
    HostScene* H = (HostScene*)h;
    memcpy(live, H->live.data(), sizeof(uint32_t) * H->live.size());
}

// out: 12 floats per ray = PtRayHit (wuvt, normal) + 3 ints in a parallel array
void hs_intersect_n(void* h, const float* ro4, const float* rd4, float tNear, int64_t n,
                    float* wuvt, float* normal, int32_t* typeVertFlags)
{
    HostScene* H = (HostScene*)h;
    LocalStack st;
    TravCounters c{ 0, 0 };
    for (int64_t i = 0; i < n; ++i)
    {
        V3 ro = v3(ro4[4 * i], ro4[4 * i + 1], ro4[4 * i + 2]);
        V3 rd = v3(rd4[4 * i], rd4[4 * i + 1], rd4[4 * i + 2]);
        Hit hit;
        traverse<false, true>(H->S.bvh, ro, rd, tNear, rd4[4 * i + 3], hit, st, &c);
        RayHit r = finishHit(H->S, rd, hit);
        wuvt[4 * i] = r.w; wuvt[4 * i + 1] = r.u; wuvt[4 * i + 2] = r.v; wuvt[4 * i + 3] = r.t;
        normal[4 * i] = r.normal.x; normal[4 * i + 1] = r.normal.y; normal[4 * i + 2] = r.normal.z; normal[4 * i + 3] = 0;
        typeVertFlags[3 * i] = r.type; typeVertFlags[3 * i + 1] = r.iVert; typeVertFlags[3 * i + 2] = (int32_t)r.flags;
    }
    H->nodeFetches += c.nodes; H->triFetches += c.tris; H->rays += (uint64_t)n;
}

void hs_occluded_n(void* h, const float* ro4, const float* rd4, float tNear, int64_t n, uint8_t* out)
{
    HostScene* H = (HostScene*)h;
    LocalStack st;
    for (int64_t i = 0; i < n; ++i)
    {
        V3 ro = v3(ro4[4 * i], ro4[4 * i + 1], ro4[4 * i + 2]);
        V3 rd = v3(rd4[4 * i], rd4[4 * i + 1], rd4[4 * i + 2]);
        Hit hit;
        out[i] = traverse<true, false>(H->S.bvh, ro, rd, tNear, rd4[4 * i + 3], hit, st, (TravCounters*)nullptr) ? 1 : 0;
    }
}

void hs_counters(void* h, uint64_t* nodes, uint64_t* tris, uint64_t* rays)
{
    HostScene* H = (HostScene*)h;
    *nodes = H->nodeFetches; *tris = H->triFetches; *rays = H->rays;
}

struct HsCam
{
    float eye[3], right[3], up[3], fwd[3]; float slope[2]; int32_t width, height;
    float aperture, focalLength; int32_t bladeCount; float bladeRot, focalPlaneCurvature;
};

// one 1-spp frame in sum form: out += sample
void hs_trace_frame(void* h, const HsCam* cam, uint32_t pass, uint32_t seed, float* color3, float* albedo3, float* normal3)
{
    HostScene* H = (HostScene*)h;
    CamParams C;
    C.eye = v3(cam->eye[0], cam->eye[1], cam->eye[2]); C.right = v3(cam->right[0], cam->right[1], cam->right[2]);
    C.up = v3(cam->up[0], cam->up[1], cam->up[2]); C.fwd = v3(cam->fwd[0], cam->fwd[1], cam->fwd[2]);
    C.slope = v2(cam->slope[0], cam->slope[1]); C.width = cam->width; C.height = cam->height;
    C.aperture = cam->aperture; C.focalLength = cam->focalLength; C.bladeCount = cam->bladeCount;
    C.bladeRot = cam->bladeRot; C.focalPlaneCurvature = cam->focalPlaneCurvature;
    LocalStack st;
    uint32_t rays = 0;
    for (int i = 0; i < C.width * C.height; ++i)
    {
        Rng rng = seed_ctx((uint32_t)i, pass, seed, Salt_Trace);
        V3 ro, rd;
        cameraRay(C, rng, i, ro, rd);
        V3 c, a, nrm;
        traceRay(H->S, rng, ro, rd, c, a, nrm, st, &rays);
        color3[3 * i] = c.x; color3[3 * i + 1] = c.y; color3[3 * i + 2] = c.z;
        albedo3[3 * i] = a.x; albedo3[3 * i + 1] = a.y; albedo3[3 * i + 2] = a.z;
        normal3[3 * i] = nrm.x; normal3[3 * i + 1] = nrm.y; normal3[3 * i + 2] = nrm.z;
    }
    H->rays += rays;
}

void hs_trace_rays(void* h, const float* ro4, const float* rd4, int64_t n, uint32_t pass, uint32_t seed, float* out9)
{
    HostScene* H = (HostScene*)h;
    LocalStack st;
    for (int64_t i = 0; i < n; ++i)
    {
        Rng rng = seed_ctx((uint32_t)i, pass, seed, Salt_Rays);
        V3 c, a, nrm;
        traceRay(H->S, rng, v3(ro4[4 * i], ro4[4 * i + 1], ro4[4 * i + 2]), v3(rd4[4 * i], rd4[4 * i + 1], rd4[4 * i + 2]),
                 c, a, nrm, st, nullptr);
        float* o = out9 + 9 * i;
        o[0] = c.x; o[1] = c.y; o[2] = c.z; o[3] = a.x; o[4] = a.y; o[5] = a.z; o[6] = nrm.x; o[7] = nrm.y; o[8] = nrm.z;
    }
}

} // extern "C"

// ---- known-answer vectors of the helpers (the counterpart of oracle/ref_kat.c): the same rows
// through the library's PT_HD functions, compiled for the host
namespace {
struct HsKat { const char* name; int kin, kout; void (*fn)(const float* in, float* out); };

void hk_pcg1(const float* i, float* o) { o[0] = u2f(Pcg1(f2u(i[0]))); }
void hk_pcg2(const float* i, float* o) { uint32_t x = f2u(i[0]), y = f2u(i[1]); Pcg2(x, y); o[0] = u2f(x); o[1] = u2f(y); }
void hk_pcg4(const float* i, float* o)
{
    Rng r = Pcg4(f2u(i[0]), f2u(i[1]), f2u(i[2]), f2u(i[3]));
    o[0] = u2f(r.x); o[1] = u2f(r.y); o[2] = u2f(r.z); o[3] = u2f(r.w);
}
void hk_tbn(const float* i, float* o)
{
    Tbn m = NormalToTBN(normalize(v3(i[0], i[1], i[2])));
    o[0] = m.c0.x; o[1] = m.c0.y; o[2] = m.c0.z; o[3] = m.c1.x; o[4] = m.c1.y; o[5] = m.c1.z; o[6] = m.c2.x; o[7] = m.c2.y; o[8] = m.c2.z;
}
void hk_power(const float* i, float* o) { o[0] = PowerHeuristic(i[0], i[1]); }
void hk_disk(const float* i, float* o) { V2 d = MapSquareToDisk(v2(i[0], i[1])); o[0] = d.x; o[1] = d.y; }
void hk_bary(const float* i, float* o) { V3 w = SampleBaryCoord(v2(i[0], i[1])); o[0] = w.x; o[1] = w.y; o[2] = w.z; }
void hk_ngon(const float* i, float* o)
{
    V2 p = SampleNGon(v2(i[0], i[1]), f2u(i[2]) % 7u, 7, 0.3f); o[0] = p.x; o[1] = p.y;
    p = SamplePentagram(v2(i[0], i[1]), f2u(i[2])); o[2] = p.x; o[3] = p.y;
}
void hk_sphere(const float* i, float* o) { V3 d = SampleUnitSphere(v2(i[0], i[1])); o[0] = d.x; o[1] = d.y; o[2] = d.z; }
void hk_coshemi(const float* i, float* o) { V3 d = SampleCosineHemisphere(v2(i[0], i[1])); o[0] = d.x; o[1] = d.y; o[2] = d.z; }
void hk_ggx(const float* i, float* o)
{
    float alpha = BrdfAlpha(i[2]);
    V3 d = SampleGGXMicrofacet(v2(i[0], i[1]), alpha); o[0] = d.x; o[1] = d.y; o[2] = d.z;
    o[3] = GGXPdf(i[0], i[1], alpha);
    o[4] = D_GTR(i[0], alpha);
    o[5] = V_SmithCorrelated(i[0], i[1], alpha);
}
void hk_lightpdf(const float* i, float* o) { o[0] = LightPdf(i[0] + 0.01f, i[1], i[2] * 10.0f); }
void hk_gauss(const float* i, float* o) { V2 d = SampleGaussPixelFilter(v2(i[0], i[1]), 1.0f); o[0] = d.x; o[1] = d.y; }
void hk_freepath(const float* i, float* o) { o[0] = SampleFreePath(i[0], 1.0f + 100.0f * i[1]); }
void hk_fresnel(const float* i, float* o)
{
    o[0] = F_Dielectric(i[0], 1.0f, 1.0f + i[1]);
    o[1] = F_Dielectric(i[0], 1.0f + i[1], 1.0f);
    V3 f0 = F_0(v3(i[0], i[1], i[2]), i[1]);
    o[2] = f0.x; o[3] = f0.y; o[4] = f0.z;
    o[5] = F_90(f0);
    o[6] = F_Schlick1(i[2], 1.0f, i[0]);
}
void hk_burley(const float* i, float* o) { o[0] = Fd_Burley(i[0], i[1], i[2], i[3]); }
void hk_transmit(const float* i, float* o)
{
    V3 t = AlbedoToTransmittance(v3(i[0], i[1], i[2]), i[3], 0.01f + i[3]); o[0] = t.x; o[1] = t.y; o[2] = t.z;
}
void hk_gamma(const float* i, float* o) { F4 v = GammaDecode_rgba8(f2u(i[0])); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
void hk_cubeuv(const float* i, float* o)
{
    V2 uv; int f = Cubemap_CalcUv(v3(i[0] - 0.5f, i[1] - 0.5f, i[2] - 0.5f), uv);
    o[0] = (float)f; o[1] = uv.x; o[2] = uv.y;
    V3 d = Cubemap_CalcDir(16, (int)(f2u(i[3]) % 6u), (int)(i[0] * 15.99f), (int)(i[1] * 15.99f), f2_tent(v2(i[1], i[2])));
    o[3] = d.x; o[4] = d.y; o[5] = d.z;
}
void hk_grid(const float* i, float* o)
{
    // Grid_New (src/math/grid.h:8-16) on the host side of the library: api.cu sceneBuild
    Grid g; g.lo = v3(-5.0f, -3.0f, -4.0f); g.hi = v3(5.0f, 3.0f, 6.0f); g.cellsPerMeter = 1.0f / 1.5f;
    g.sx = (int)ceilf((g.hi.x - g.lo.x) * g.cellsPerMeter); g.sy = (int)ceilf((g.hi.y - g.lo.y) * g.cellsPerMeter);
    g.sz = (int)ceilf((g.hi.z - g.lo.z) * g.cellsPerMeter);
    V3 p = v3(i[0] * 14.0f - 7.0f, i[1] * 8.0f - 4.0f, i[2] * 12.0f - 5.0f);
    int idx = Grid_Index(g, p);
    V3 c = Grid_Position(g, idx);
    o[0] = (float)idx; o[1] = c.x; o[2] = c.y; o[3] = c.z; o[4] = (float)Grid_Len(g);
}
void hk_sdtri(const float* i, float* o)
{
    o[0] = sdTriangle3D(v3(i[0], i[1], i[2]), v3(i[3], i[4], i[5]), v3(i[6], i[7], i[8]), v3(i[9] * 2.0f, i[10] * 2.0f, i[11] * 2.0f));
    o[1] = sdTriangle2D(v2(i[0], i[1]), v2(i[3], i[4]), v2(i[6], i[7]), v2(i[9] * 1.5f, i[10] * 1.5f));
}
void hk_mie(const float* i, float* o)
{
    float c = i[0] * 2.0f - 1.0f;
    o[0] = MiePhase(c, 0.758f);
    o[1] = (3.0f / (16.0f * kPi)) * (1.0f + c * c);        // RayleighPhase as k_skyBake writes it
}
void hk_sgfit(const float* i, float* o)
{
    F4 axii[5], amps[5];
    memcpy(axii, i + 2, sizeof(axii));
    memset(amps, 0, sizeof(amps));
    Rng s; s.x = f2u(i[0]); s.y = f2u(i[1]); s.z = 7u; s.w = 9u;
    for (int k = 0; k < 64; ++k)
    {
        s = Pcg4(s.x, s.y, s.z, s.w);
        V2 Xi = v2(ToFloat(s.x), ToFloat(s.y));
        V3 dir = SampleUnitHemisphere(Xi);
        V3 rad = v3(ToFloat(s.z), ToFloat(s.w), 0.5f);
        SG_Accumulate(1.0f / (1.0f + k), dir, rad, axii, amps, 5);
    }
    for (int k = 0; k < 5; ++k) { o[4 * k] = amps[k].x; o[4 * k + 1] = amps[k].y; o[4 * k + 2] = amps[k].z; o[4 * k + 3] = amps[k].w; }
}
void hk_dist1d(const float* i, float* o)
{
    int s = FindInterval(i, 9, i[17]);           // Dist1D_SampleD
    o[0] = (float)s; o[1] = i[9 + s] / 8.0f;     // Dist1D_PdfD (integrator.h Dist_PdfD)
}

const HsKat kHsKats[] =
{
    { "pcg1", 1, 1, hk_pcg1 }, { "pcg2", 2, 2, hk_pcg2 }, { "pcg4", 4, 4, hk_pcg4 },
    { "tbn", 3, 9, hk_tbn }, { "power", 2, 1, hk_power }, { "disk", 2, 2, hk_disk }, { "bary", 2, 3, hk_bary },
    { "ngon", 3, 4, hk_ngon }, { "sphere", 2, 3, hk_sphere }, { "coshemi", 2, 3, hk_coshemi }, { "ggx", 3, 6, hk_ggx },
    { "lightpdf", 3, 1, hk_lightpdf }, { "gauss", 2, 2, hk_gauss }, { "freepath", 2, 1, hk_freepath },
    { "fresnel", 3, 7, hk_fresnel }, { "burley", 4, 1, hk_burley }, { "transmit", 4, 3, hk_transmit },
    { "gamma", 1, 4, hk_gamma }, { "cubeuv", 4, 6, hk_cubeuv }, { "grid", 3, 5, hk_grid }, { "sdtri", 12, 2, hk_sdtri },
    { "mie", 1, 2, hk_mie }, { "sgfit", 22, 20, hk_sgfit }, { "dist1d", 18, 2, hk_dist1d },
};
} // namespace

extern "C" int hs_kat_run(const char* name, const float* in, int n, float* out)
{
    for (const HsKat& k : kHsKats)
    {
        if (strcmp(k.name, name) == 0)
        {
            for (int r = 0; r < n; ++r) k.fn(in + (size_t)r * k.kin, out + (size_t)r * k.kout);
            return k.kout;
        }
    }
    return -1;
}
