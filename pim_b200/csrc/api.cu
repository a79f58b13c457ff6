// api.cu — the C ABI of libpimpt.so (include/pim_pt.h): the reference's entry
// points of src/rendering/path_tracer.h:67-103 and LmPack_Bake, implemented on
// the CUDA kernels of this directory. No CPU implementation exists behind any
// of them: without a CUDA device every call fails with Pt_LastError set.
#include <cuda_runtime.h>
#define PIM_PT_NO_VECTOR_TYPES 1
#include "../../include/pim_pt.h"
#include "pt_device.h"
#include "arena.h"
#include "hostpool.h"
#include "hostcopy.h"

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <thread>
#include <string>

namespace pt {

static thread_local char t_error[512];
LibCounters g_counters;
static float g_lastKernelMs = 0.0f;

void setError(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_error, sizeof(t_error), fmt, ap);
    va_end(ap);
}

bool cudaOk(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return true;
    setError("CUDA error %s (%d) at %s", cudaGetErrorString(e), (int)e, what);
    return false;
}

Arena& stagingArena(int device)
{
    static Arena arenas[kMaxDevices];
    return arenas[(device >= 0 && device < kMaxDevices) ? device : 0];
}
int currentDevice()
{
    int d = 0;
    cudaGetDevice(&d);
    return d;
}

} // namespace pt

using namespace pt;

struct DeviceImages
{
    float* image[3] = { nullptr, nullptr, nullptr };   // running means, w*h*3 floats each
    float* sum[3] = { nullptr, nullptr, nullptr };     // per-call sums
    int64_t texels = 0;
};

struct PtScene_s
{
    cudaStream_t stream = nullptr;
    DScene view;
    BvhBuffers bvh;
    std::vector<void*> allocs;
    PathPool pool;
    PtConfig config;
    float* dLut = nullptr;
    int32_t gridCells = 0;
    uint32_t nextPass = 0;
    float buildMs = 0.0f;
    float phaseMs[6] = { 0, 0, 0, 0, 0, 0 };
    // Pt_RayGen results: library-owned, valid until the second next call (the reference
    // hands out per-frame linear-allocator memory, path_tracer.c:2630-2631)
    pim_float4* rayGenColors[2] = { nullptr, nullptr };
    pim_float4* rayGenDirs[2] = { nullptr, nullptr };
    int32_t rayGenCap[2] = { 0, 0 };
    int rayGenFlip = 0;
    int device = 0;                 // the CUDA device the scene lives on (current when it was built)
    uint64_t descVersion = 0;       // PtSys_SetSceneDesc generation the scene was built from (0: an explicit description)
    std::vector<int32_t> entityLengths;     // vertices per entity, in order: the shape a refit must find again
    bool refitted = false;          // the tree was refitted from the scene this one replaced, not built
    // ray cursors of the persistent batch kernels: one 128-byte line per call in flight, handed out
    // round robin, so two batches on different streams never share a cursor
    static constexpr uint32_t kCursorRing = 64;
    unsigned long long* cursorRing = nullptr;
    unsigned long long* nextCursor() const;
    // in-process multi-GPU (PtSys_SetDevices): full replicas of this scene on the other devices, and the
    // buffers their partial results pass through. Growable, owned by the scene, freed with it.
    std::vector<PtScene_s*> replicas;
    float* partial[3] = { nullptr, nullptr, nullptr };  // replica: its partial sum images; primary: where a peer's arrive
    size_t partialFloats = 0;
    uint32_t* liveBase = nullptr;                       // light-learning counters before the call (replica: then its delta)
    size_t liveWords = 0;
};
static std::atomic<uint32_t> g_cursorNext{ 0 };
unsigned long long* PtScene_s::nextCursor() const
{
    return cursorRing + 16 * (size_t)(g_cursorNext.fetch_add(1u) % kCursorRing);
}

static bool g_initialised = false;
static std::map<PtTrace*, DeviceImages> g_traces;
static std::mutex g_mutex;

// PtSys_SetSceneDesc keeps its own copy of everything the description points to (the adapter of
// INTEGRATION.md fills it from per-frame memory): mesh arrays once per distinct pointer, so
// instanced meshes stay shared, texel blocks, sky faces and the BRDF table.
struct SceneDescCopy
{
    PtSceneDesc desc;
    std::vector<PtEntityDesc> entities;
    std::vector<PtTextureDesc> textures;
    std::vector<std::vector<pim_float4>> meshArrays;
    std::vector<std::vector<uint32_t>> texelBlocks;
    std::vector<float> sky, lut;
    bool valid = false;
    uint64_t version = 0;

    void clear()
    {
        entities.clear(); textures.clear(); meshArrays.clear(); texelBlocks.clear(); sky.clear(); lut.clear();
        memset(&desc, 0, sizeof(desc));
        valid = false;
    }
    void assign(const PtSceneDesc* src)
    {
        clear();
        ++version;
        if (!src) return;
        desc = *src;
        std::map<const void*, const pim_float4*> seen;
        auto dup = [&](const pim_float4* p, int32_t n) -> const pim_float4*
        {
            if (!p || n <= 0) return nullptr;
            auto it = seen.find(p);
            if (it != seen.end()) return it->second;
            meshArrays.emplace_back(p, p + n);
            seen[p] = meshArrays.back().data();
            return meshArrays.back().data();
        };
        entities.assign(src->entities, src->entities + std::max(src->entityCount, 0));
        for (PtEntityDesc& e : entities)
        {
            e.positions = dup(e.positions, e.length);
            e.normals = dup(e.normals, e.length);
            e.uvs = dup(e.uvs, e.length);
        }
        textures.assign(src->textures, src->textures + std::max(src->textureCount, 0));
        for (PtTextureDesc& t : textures)
        {
            const size_t n = (size_t)std::max(t.width, 0) * (size_t)std::max(t.height, 0);
            const uint32_t* q = (const uint32_t*)t.texels;
            texelBlocks.emplace_back(q ? q : nullptr, q ? q + n : nullptr);     // RGBA8 and RG16 texels are both 4 bytes
            t.texels = texelBlocks.back().data();
        }
        if (src->skyFaces && src->skySize > 0)
        {
            sky.assign(src->skyFaces, src->skyFaces + (size_t)6 * src->skySize * src->skySize * 3);
            desc.skyFaces = sky.data();
        }
        if (src->brdfLut && src->lutWidth > 0 && src->lutHeight > 0)
        {
            lut.assign(src->brdfLut, src->brdfLut + (size_t)2 * src->lutWidth * src->lutHeight);
            desc.brdfLut = lut.data();
        }
        desc.entities = entities.data();
        desc.textures = textures.data();
        valid = true;
    }
};
static SceneDescCopy g_sceneDesc;

// the lightmap pack LmPack_Bake works on (the reference's LmPack_Get(), lightmap.c:136): page
// descriptions copied, the arrays stay the engine's (they are where the result has to appear)
static std::vector<PtLightmapDesc> g_lmPack;
static float g_lmTexelsPerMeter = 0.0f;

// In-process multi-GPU (PtSys_SetDevices): pim is one process, so the library itself can spread a call over
// the first g_deviceCount devices: every scene built afterwards carries a full replica per extra device, one
// host thread per device drives it, partial results come to the scene's own device by peer copy.
static int g_deviceCount = 1;
static bool g_refitOnUpdate = false;    // Pt_SetOption("refit"): PtScene_Update refits the tree when the geometry kept its shape

struct DeviceGuard
{
    int prev = 0;
    explicit DeviceGuard(int d) { cudaGetDevice(&prev); if (d != prev) cudaSetDevice(d); }
    ~DeviceGuard() { int cur = 0; cudaGetDevice(&cur); if (cur != prev) cudaSetDevice(prev); }
};

// runs fn(k) for k in [0, count): k = 0 on the calling thread, the others on a thread each; false (and the
// first failing thread's message in Pt_LastError) when any of them failed
template <class Fn>
static bool forEachDevice(int count, Fn fn)
{
    if (count <= 1) return fn(0);
    std::vector<std::thread> threads;
    std::vector<int> okFlags((size_t)count, 1);
    std::vector<std::string> errors((size_t)count);
    for (int k = 1; k < count; ++k)
    {
        threads.emplace_back([&, k]()
        {
            t_workerThread = true;
            t_error[0] = 0;
            okFlags[(size_t)k] = fn(k) ? 1 : 0;
            if (!okFlags[(size_t)k]) errors[(size_t)k] = t_error;
        });
    }
    okFlags[0] = fn(0) ? 1 : 0;
    for (std::thread& t : threads) t.join();
    for (int k = 1; k < count; ++k)
    {
        if (!okFlags[(size_t)k]) { if (okFlags[0]) setError("device thread %d: %s", k, errors[(size_t)k].c_str()); okFlags[0] = 0; }
    }
    return okFlags[0] != 0;
}

template <class T>
static bool sceneAlloc(PtScene* s, T*& p, size_t count)
{
    p = nullptr;
    if (count == 0) count = 1;
    if (!PT_CUDA(cudaMalloc((void**)&p, sizeof(T) * count))) return false;
    s->allocs.push_back(p);
    return true;
}

static bool ensureInit()
{
    if (g_initialised) return true;
    int n = 0;
    if (!PT_CUDA(cudaGetDeviceCount(&n))) return false;
    if (n <= 0) { setError("no CUDA device: libpimpt has no CPU path"); return false; }
    if (const char* e = getenv("PIMPT_PERSISTENT")) g_persistentTraversal = atoi(e);   // tuning switches
    if (const char* e = getenv("PIMPT_FINISH")) g_finishThreshold = atoi(e);
    g_initialised = true;
    return true;
}

// ---- host math for the camera and entity matrices (same formulas as the reference) ----

static V3 quatMulDir(const float q[4], V3 dir)       // quat_funcs.h:86-97
{
    V3 qv = v3(q[0], q[1], q[2]);
    V3 uv = cross(qv, dir);
    V3 uuv = cross(qv, uv);
    V3 y = uv * q[3];
    y = y + uuv;
    y = y * 2.0f;
    return y + dir;
}

static void inverseTranspose3x3(const float* M, float* out)   // f3x3_IM, float3x3_funcs.h:54-110
{
    // m = transpose(upper 3x3 of M); [column][row]
    float m00 = M[0], m01 = M[4], m02 = M[8];
    float m10 = M[1], m11 = M[5], m12 = M[9];
    float m20 = M[2], m21 = M[6], m22 = M[10];
    float m_00_11_10_01 = m00 * m11 - m10 * m01;
    float m_00_12_10_02 = m00 * m12 - m10 * m02;
    float m_00_21_20_01 = m00 * m21 - m20 * m01;
    float m_00_22_20_02 = m00 * m22 - m20 * m02;
    float m_01_12_11_02 = m01 * m12 - m11 * m02;
    float m_01_22_21_02 = m01 * m22 - m21 * m02;
    float m_10_21_20_11 = m10 * m21 - m20 * m11;
    float m_10_22_20_12 = m10 * m22 - m20 * m12;
    float m_11_22_21_12 = m11 * m22 - m21 * m12;
    float d0 = m00 * m_11_22_21_12;
    float d1 = -m10 * m_01_22_21_02;
    float d2 = m20 * m_01_12_11_02;
    float rcpDet = 1.0f / (d0 + d1 + d2);
    out[0] = rcpDet * m_11_22_21_12;   // 00
    out[3] = -rcpDet * m_10_22_20_12;  // 10
    out[6] = rcpDet * m_10_21_20_11;   // 20
    out[1] = -rcpDet * m_01_22_21_02;  // 01
    out[4] = rcpDet * m_00_22_20_02;   // 11
    out[7] = -rcpDet * m_00_21_20_01;  // 21
    out[2] = rcpDet * m_01_12_11_02;   // 02
    out[5] = -rcpDet * m_00_12_10_02;  // 12
    out[8] = rcpDet * m_00_11_10_01;   // 22
}

static float f1_ceil_ref(float x)     // scalar.h:149-157
{
    float i = (float)(int32_t)x;
    return i + ((x > i) ? 1.0f : 0.0f);
}

static CamParams makeCam(const Camera* camera, const PtDofInfo* dof, int w, int h)
{
    CamParams C;
    const float q[4] = { camera->rotation.x, camera->rotation.y, camera->rotation.z, camera->rotation.w };
    C.eye = v3(camera->position.x, camera->position.y, camera->position.z);
    C.right = quatMulDir(q, v3(1.0f, 0.0f, 0.0f));
    C.up = quatMulDir(q, v3(0.0f, 1.0f, 0.0f));
    C.fwd = quatMulDir(q, v3(0.0f, 0.0f, -1.0f));
    const float fovy = camera->fovy * (kTau / 360.0f);
    const float tanHalfFov = tanf(fovy * 0.5f);
    C.slope = v2(((float)w / (float)h) * tanHalfFov, tanHalfFov);
    C.width = w; C.height = h;
    C.aperture = dof->aperture; C.focalLength = dof->focalLength; C.bladeCount = dof->bladeCount;
    C.bladeRot = dof->bladeRot; C.focalPlaneCurvature = dof->focalPlaneCurvature;
    return C;
}

// ---- scene construction -------------------------------------------------------

static void sceneFree(PtScene* s)
{
    if (!s) return;
    for (PtScene_s* r : s->replicas) sceneFree(r);
    s->replicas.clear();
    DeviceGuard guard(s->device);
    for (int k = 0; k < 3; ++k) if (s->partial[k]) cudaFree(s->partial[k]);
    if (s->liveBase) cudaFree(s->liveBase);
    for (void* p : s->allocs) cudaFree(p);
    freeBvh(s->bvh);
    poolFree(s->pool);
    for (int i = 0; i < 2; ++i)
    {
        if (s->rayGenColors[i]) cudaFreeHost(s->rayGenColors[i]);
        if (s->rayGenDirs[i]) cudaFreeHost(s->rayGenDirs[i]);
    }
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

// reuse: the tree of the scene this one replaces, refitted to the new vertices when the geometry still has its
// shape (same triangle count, same entity lengths), else dropped
static PtScene* sceneBuildOne(const PtSceneDesc* desc, PtScene* reuse = nullptr)
{
    if (!ensureInit()) return nullptr;
    if (!desc) { setError("PtScene_NewFromDesc: null description"); return nullptr; }
    auto t0 = std::chrono::steady_clock::now();
    auto tPhase = t0;
    PtScene* s = new PtScene_s();
    auto phase = [&](int k)
    {
        cudaStreamSynchronize(s->stream);
        auto now = std::chrono::steady_clock::now();
        s->phaseMs[k] += std::chrono::duration<float, std::milli>(now - tPhase).count();
        tPhase = now;
    };
    memset(&s->view, 0, sizeof(s->view));
    cudaGetDevice(&s->device);
    s->config = desc->config;
    if (s->config.ptDistMeters <= 0.0f) s->config.ptDistMeters = 1.5f;
    bool ok = PT_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    cudaStream_t st = s->stream;
    DScene& V = s->view;
    V.media = MediaDesc_Default();
    ok = ok && sceneAlloc(s, s->cursorRing, (size_t)16 * PtScene_s::kCursorRing);

    // textures
    std::vector<DTexture> texTable((size_t)std::max(desc->textureCount, 0));
    for (int i = 0; ok && i < desc->textureCount; ++i)
    {
        const PtTextureDesc& t = desc->textures[i];
        uint32_t* d = nullptr;
        size_t n = (size_t)t.width * (size_t)t.height;
        ok = sceneAlloc(s, d, n);
        ok = ok && copyIn(d, t.texels, n * 4, st);
        texTable[i].texels = d; texTable[i].w = t.width; texTable[i].h = t.height;
    }
    DTexture* dTex = nullptr;
    ok = ok && sceneAlloc(s, dTex, texTable.size());
    if (ok && !texTable.empty())
        ok = PT_CUDA(cudaMemcpyAsync(dTex, texTable.data(), sizeof(DTexture) * texTable.size(), cudaMemcpyHostToDevice, st));
    V.textures = dTex;

    // entities: concatenate distinct source meshes, one upload each
    int32_t vertCount = 0, srcCount = 0;
    std::map<const void*, int32_t> srcOffsets;
    std::vector<EntityUpload> ents((size_t)std::max(desc->entityCount, 0));
    std::vector<DMaterial> mats((size_t)std::max(desc->entityCount, 0));
    for (int i = 0; i < desc->entityCount; ++i)
    {
        const PtEntityDesc& e = desc->entities[i];
        EntityUpload& u = ents[i];
        memcpy(u.M, e.matrix, sizeof(u.M));
        inverseTranspose3x3(e.matrix, u.IM);
        u.vertOffset = vertCount; u.length = e.length; u.pad = 0;
        auto it = srcOffsets.find(e.positions);
        if (it == srcOffsets.end())
        {
            srcOffsets[e.positions] = srcCount;
            u.srcOffset = srcCount;
            srcCount += e.length;
        }
        else u.srcOffset = it->second;
        vertCount += e.length;
        DMaterial m; memset(&m, 0, sizeof(m));
        m.albedo = e.material.albedo; m.rome = e.material.rome; m.normal = e.material.normal;
        m.flags = e.material.flags; memcpy(m.mfp, e.material.meanFreePath, sizeof(m.mfp));
        m.ior = e.material.ior; m.bumpiness = e.material.bumpiness;
        mats[i] = m;
    }
    F4 *dSrcPos = nullptr, *dSrcNrm = nullptr, *dSrcUv = nullptr;
    ok = ok && PT_CUDA(cudaMalloc((void**)&dSrcPos, sizeof(F4) * (size_t)std::max(srcCount, 1)));
    ok = ok && PT_CUDA(cudaMalloc((void**)&dSrcNrm, sizeof(F4) * (size_t)std::max(srcCount, 1)));
    ok = ok && PT_CUDA(cudaMalloc((void**)&dSrcUv, sizeof(F4) * (size_t)std::max(srcCount, 1)));
    {
        std::map<const void*, bool> done;
        for (int i = 0; ok && i < desc->entityCount; ++i)
        {
            const PtEntityDesc& e = desc->entities[i];
            if (done[e.positions]) continue;
            done[e.positions] = true;
            const size_t bytes = sizeof(F4) * (size_t)e.length;
            ok = ok && copyIn(dSrcPos + ents[i].srcOffset, e.positions, bytes, st);
            ok = ok && copyIn(dSrcNrm + ents[i].srcOffset, e.normals, bytes, st);
            ok = ok && copyIn(dSrcUv + ents[i].srcOffset, e.uvs, bytes, st);
        }
    }
    EntityUpload* dEnts = nullptr;
    DMaterial* dMats = nullptr;
    F4 *dPos = nullptr, *dNrm = nullptr;
    V2* dUv = nullptr;
    int32_t *dMatIds = nullptr, *dVertToEmit = nullptr, *dEmitToVert = nullptr;
    ok = ok && PT_CUDA(cudaMalloc((void**)&dEnts, sizeof(EntityUpload) * std::max<size_t>(ents.size(), 1)));
    ok = ok && sceneAlloc(s, dMats, mats.size()) && sceneAlloc(s, dPos, (size_t)vertCount) &&
         sceneAlloc(s, dNrm, (size_t)vertCount) && sceneAlloc(s, dUv, (size_t)vertCount) &&
         sceneAlloc(s, dMatIds, (size_t)vertCount) && sceneAlloc(s, dVertToEmit, (size_t)vertCount);
    if (ok && !ents.empty())
    {
        ok = PT_CUDA(cudaMemcpyAsync(dEnts, ents.data(), sizeof(EntityUpload) * ents.size(), cudaMemcpyHostToDevice, st));
        ok = ok && PT_CUDA(cudaMemcpyAsync(dMats, mats.data(), sizeof(DMaterial) * mats.size(), cudaMemcpyHostToDevice, st));
    }
    ok = ok && launchFlatten(dEnts, desc->entityCount, dSrcPos, dSrcNrm, dSrcUv, dPos, dNrm, dUv, dMatIds, vertCount, st);
    V.positions = dPos; V.normals = dNrm; V.uvs = dUv; V.matIds = dMatIds; V.materials = dMats;
    V.vertCount = vertCount; V.matCount = desc->entityCount;

    phase(0);
    // sky + BRDF LUT
    if (ok && desc->skyFaces && desc->skySize > 0)
    {
        float* dSky = nullptr;
        size_t n = (size_t)6 * desc->skySize * desc->skySize * 3;
        ok = sceneAlloc(s, dSky, n) && copyIn(dSky, desc->skyFaces, n * 4, st);
        V.sky = dSky; V.skySize = desc->skySize;
    }
    {
        int lw = desc->brdfLut ? desc->lutWidth : 128, lh = desc->brdfLut ? desc->lutHeight : 128;
        ok = ok && sceneAlloc(s, s->dLut, (size_t)lw * lh * 2);
        if (ok && desc->brdfLut)
            ok = PT_CUDA(cudaMemcpyAsync(s->dLut, desc->brdfLut, sizeof(float) * 2 * lw * lh, cudaMemcpyHostToDevice, st));
        else if (ok)
            ok = launchBrdfLutBake(s->dLut, lw, lh, 4096, s->config.seed, st);
        V.lut = s->dLut; V.lutW = lw; V.lutH = lh;
    }

    phase(4);
    // BVH (RtcNewScene)
    const int32_t triCount = vertCount / 3;
    for (int i = 0; i < desc->entityCount; ++i) s->entityLengths.push_back(desc->entities[i].length);
    if (ok && reuse && reuse->bvh.triCount == triCount && reuse->bvh.nodes && reuse->entityLengths == s->entityLengths)
    {
        s->bvh = std::move(reuse->bvh);
        reuse->bvh = BvhBuffers();
        ok = refitBvh(dPos, triCount, s->bvh, st);
        s->refitted = true;
    }
    else ok = ok && buildBvh(dPos, triCount, s->bvh, st);
    V.bvh.nodes = s->bvh.nodes; V.bvh.tris = s->bvh.tris; V.bvh.nodesF = s->bvh.nodesF;
    V.bvh.nodeCount = s->bvh.nodeCount; V.bvh.triCount = s->bvh.triCount;

    phase(1);
    // SetupEmissives (path_tracer.c:845-883): device pdf, ordered compaction
    int32_t emissiveCount = 0;
    if (ok && triCount > 0)
    {
        float* dPdf = nullptr;
        ok = PT_CUDA(cudaMalloc((void**)&dPdf, sizeof(float) * (size_t)triCount));
        ok = ok && launchEmissionPdf(V, triCount, s->config.seed, dPdf, st);
        std::vector<float> pdf((size_t)triCount);
        ok = ok && PT_CUDA(cudaMemcpyAsync(pdf.data(), dPdf, sizeof(float) * (size_t)triCount, cudaMemcpyDeviceToHost, st));
        ok = ok && PT_CUDA(cudaStreamSynchronize(st));
        cudaFree(dPdf);
        if (ok)
        {
            std::vector<int32_t> vertToEmit((size_t)vertCount, -1), emitToVert;
            for (int32_t t = 0; t < triCount; ++t)
            {
                if (pdf[t] > 0.01f)
                {
                    vertToEmit[3 * t] = vertToEmit[3 * t + 1] = vertToEmit[3 * t + 2] = emissiveCount;
                    emitToVert.push_back(3 * t);
                    ++emissiveCount;
                }
            }
            ok = sceneAlloc(s, dEmitToVert, emitToVert.size());
            ok = ok && PT_CUDA(cudaMemcpyAsync(dVertToEmit, vertToEmit.data(), sizeof(int32_t) * vertToEmit.size(), cudaMemcpyHostToDevice, st));
            if (ok && !emitToVert.empty())
                ok = PT_CUDA(cudaMemcpyAsync(dEmitToVert, emitToVert.data(), sizeof(int32_t) * emitToVert.size(), cudaMemcpyHostToDevice, st));
            ok = ok && PT_CUDA(cudaStreamSynchronize(st));
        }
    }
    V.vertToEmit = dVertToEmit; V.emitToVert = dEmitToVert; V.emissiveCount = emissiveCount;

    phase(2);
    // SetupLightGrid (path_tracer.c:992-1009): Grid_New over box_from_pts
    if (ok && vertCount > 0)
    {
        Grid g;
        g.lo = v3(s->bvh.sceneLo[0], s->bvh.sceneLo[1], s->bvh.sceneLo[2]);
        g.hi = v3(s->bvh.sceneHi[0], s->bvh.sceneHi[1], s->bvh.sceneHi[2]);
        g.cellsPerMeter = 1.0f / s->config.ptDistMeters;
        g.sx = (int)f1_ceil_ref((g.hi.x - g.lo.x) * g.cellsPerMeter);
        g.sy = (int)f1_ceil_ref((g.hi.y - g.lo.y) * g.cellsPerMeter);
        g.sz = (int)f1_ceil_ref((g.hi.z - g.lo.z) * g.cellsPerMeter);
        V.grid = g;
        const size_t cells = (size_t)std::max(Grid_Len(g), 0);
        s->gridCells = (int32_t)cells;
        const size_t E = (size_t)emissiveCount;
        ok = sceneAlloc(s, V.distPdf, cells * E) && sceneAlloc(s, V.distCdf, cells * (E + 1)) &&
             sceneAlloc(s, V.distLive, cells * E) && sceneAlloc(s, V.distLength, cells) && sceneAlloc(s, V.distSum, cells);
        ok = ok && PT_CUDA(cudaMemsetAsync(V.distLength, 0, sizeof(int32_t) * std::max<size_t>(cells, 1), st));
        ok = ok && PT_CUDA(cudaMemsetAsync(V.distLive, 0, sizeof(uint32_t) * std::max<size_t>(cells * E, 1), st));
        ok = ok && launchLightGrid(V, s->config.ptDistMeters, s->config.seed, st);
    }

    phase(3);
    // wavefront pool: 8 M slots x 324 B = 2.7 GB. Measured on cornell 1080p, 64 spp per call (Msamples/s):
    // 2 M 2161, 4 M 2336, 6 M 2392, 8 M 2424, 12 M 2427, 16 M 2418, 24 M 2407 - the state streams through
    // HBM fully coalesced either way, and fewer, larger iterations amortise launches and the drain at the
    // end of a call. Calls with less work than slots run a correspondingly smaller range.
    if (ok)
    {
        int32_t cap = s->config.pathPoolSize > 0 ? s->config.pathPoolSize : (1 << 23);
        cap = (cap + 255) & ~255;
        ok = poolAlloc(s->pool, cap);
    }
    ok = ok && PT_CUDA(cudaStreamSynchronize(st));
    phase(5);
    if (dSrcPos) cudaFree(dSrcPos);
    if (dSrcNrm) cudaFree(dSrcNrm);
    if (dSrcUv) cudaFree(dSrcUv);
    if (dEnts) cudaFree(dEnts);
    if (!ok)
    {
        if (s->refitted && reuse) { reuse->bvh = std::move(s->bvh); s->bvh = BvhBuffers(); }     // the tree goes back with its owner
        sceneFree(s);
        return nullptr;
    }
    s->buildMs = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return s;
}

// the scene on the current device plus, in multi-GPU mode, a replica on each of the other devices (the
// build is deterministic, so the replicas hold the same tree and the same light grid)
static PtScene* sceneBuild(const PtSceneDesc* desc, PtScene* reuse = nullptr)
{
    if (!ensureInit()) return nullptr;
    int primary = 0;
    if (!PT_CUDA(cudaGetDevice(&primary))) return nullptr;
    std::vector<int> devices(1, primary);
    for (int d = 0; d < g_deviceCount; ++d) if (d != primary) devices.push_back(d);
    if (g_deviceCount <= 1) devices.resize(1);
    std::vector<PtScene*> built(devices.size(), nullptr);
    const bool ok = forEachDevice((int)devices.size(), [&](int k) -> bool
    {
        DeviceGuard guard(devices[(size_t)k]);
        PtScene* old = nullptr;         // the replaced scene's tree on the same device
        if (reuse && k == 0 && reuse->device == devices[0]) old = reuse;
        else if (reuse && k > 0 && (size_t)k <= reuse->replicas.size() && reuse->replicas[(size_t)k - 1]->device == devices[(size_t)k])
            old = reuse->replicas[(size_t)k - 1];
        built[(size_t)k] = sceneBuildOne(desc, old);
        return built[(size_t)k] != nullptr;
    });
    if (!ok)
    {
        for (PtScene* b : built) sceneFree(b);
        return nullptr;
    }
    for (size_t k = 1; k < built.size(); ++k) built[0]->replicas.push_back(built[k]);
    return built[0];
}

// ---- helpers for host <-> device batches ---------------------------------------

// CUDA events around a call's kernels on their stream; stop() leaves the time in Pt_LastKernelMs().
// The two events belong to the staging arena of the call's lease (created once per device).
struct KernelTimer
{
    cudaEvent_t a, b;
    cudaStream_t st;
    KernelTimer(Lease& lease, cudaStream_t s) : a(lease.arena().evA), b(lease.arena().evB), st(s)
    {
        if (a && b) cudaEventRecord(a, st);
    }
    void stop()
    {
        if (!a || !b) return;
        cudaEventRecord(b, st);
        float ms = 0.0f;
        if (cudaEventSynchronize(b) == cudaSuccess && cudaEventElapsedTime(&ms, a, b) == cudaSuccess) g_lastKernelMs = ms;
    }
};

// PtScene_Update (path_tracer.c:1024-1037): a scene built from the registered description is rebuilt in
// place when the description has changed since (the reference compares Entities_Get()->modtime), then
// the light-grid distributions relearn (UpdateDists)
static bool sceneRefresh(PtScene* scene, cudaStream_t st)
{
    if (scene->descVersion != 0 && g_sceneDesc.valid && scene->descVersion != g_sceneDesc.version)
    {
        // Pt_SetOption("refit", 1): keep the tree's topology and refit it when the geometry has the same shape
        for (PtScene* r : scene->replicas) { DeviceGuard guard(r->device); cudaStreamSynchronize(r->stream); }
        cudaStreamSynchronize(scene->stream);
        PtScene* fresh = sceneBuild(&g_sceneDesc.desc, g_refitOnUpdate ? scene : nullptr);
        if (!fresh) return false;
        fresh->descVersion = g_sceneDesc.version;
        fresh->nextPass = scene->nextPass;
        const bool ownStream = st == scene->stream;
        std::swap(*scene, *fresh);          // the caller's handle now owns the new scene
        sceneFree(fresh);
        if (ownStream) st = scene->stream;
    }
    return launchDistUpdate(scene->view, st);
}

template <class T>
static bool growBuffer(T*& p, size_t& have, size_t need)
{
    if (need <= have) return true;
    if (p) cudaFree(p);
    p = nullptr; have = 0;
    if (!PT_CUDA(cudaMalloc((void**)&p, sizeof(T) * need))) return false;
    have = need;
    return true;
}

// one device's share of a Pt_Trace / Pt_TraceSpp call: samples [sppBegin, +sppCount) of pixels [pixelBegin, +pixelCount)
struct TraceShare { int sppBegin, sppCount, pixelBegin, pixelCount; };

static bool traceShare(PtScene* s, const Camera* camera, const PtDofInfo* dof, int w, int h, const TraceShare& sh,
                       float* dColor, float* dAlbedo, float* dNormal, cudaStream_t st)
{
    if (sh.sppCount <= 0 || sh.pixelCount <= 0) return true;
    WorkDesc W; memset(&W, 0, sizeof(W));
    W.kind = Work_Camera;
    W.cam = makeCam(camera, dof, w, h);
    W.pixelBegin = sh.pixelBegin; W.pixelCount = sh.pixelCount;
    W.count = (int64_t)sh.pixelCount * sh.sppCount;
    W.passBase = (uint32_t)sh.sppBegin;
    W.seed = s->config.seed;
    W.salt = Salt_Trace;
    W.sumColor = dColor; W.sumAlbedo = dAlbedo; W.sumNormal = dNormal;
    W.accumulate = 1;
    float ms = 0.0f;
    const bool ok = runWavefront(s->view, s->pool, W, st, &ms);
    if (!t_workerThread) g_lastKernelMs = ms;
    return ok;
}

// The light-learning counters (Dist1D.live) of the replicas are kept equal to what one device tracing everything
// would hold: every device notes its counters before the call, the replicas turn that into "what this call added",
// the primary adds those and hands the totals back (all-reduce of the u32 counters per frame).
static bool liveSnapshot(PtScene* s, cudaStream_t st)
{
    const size_t words = (size_t)s->gridCells * (size_t)s->view.emissiveCount;
    if (words == 0) return true;
    if (!growBuffer(s->liveBase, s->liveWords, words)) return false;
    return PT_CUDA(cudaMemcpyAsync(s->liveBase, s->view.distLive, 4 * words, cudaMemcpyDeviceToDevice, st));
}

// Before a call is shared out, every replica takes the primary's light grid as it is now (pdf, cdf, counters, sums,
// lengths): the primary has just relearned (PtScene_Update), and calls that run on the primary alone (Pt_RayGen,
// cubemap bakes, ray batches that feed LightOnHit) have moved its counters since the last shared call.
static bool lightGridToReplicas(PtScene* scene, cudaStream_t st)
{
    const size_t cells = (size_t)scene->gridCells, E = (size_t)scene->view.emissiveCount;
    if (cells == 0 || scene->replicas.empty()) return true;
    bool ok = true;
    const DScene& P = scene->view;
    for (PtScene_s* r : scene->replicas)
    {
        const DScene& R = r->view;
        if (E > 0)
        {
            ok = ok && PT_CUDA(cudaMemcpyPeerAsync(R.distPdf, r->device, P.distPdf, scene->device, 4 * cells * E, st)) &&
                 PT_CUDA(cudaMemcpyPeerAsync(R.distCdf, r->device, P.distCdf, scene->device, 4 * cells * (E + 1), st)) &&
                 PT_CUDA(cudaMemcpyPeerAsync(R.distLive, r->device, P.distLive, scene->device, 4 * cells * E, st));
        }
        ok = ok && PT_CUDA(cudaMemcpyPeerAsync(R.distLength, r->device, P.distLength, scene->device, 4 * cells, st)) &&
             PT_CUDA(cudaMemcpyPeerAsync(R.distSum, r->device, P.distSum, scene->device, 4 * cells, st));
    }
    return PT_CUDA(cudaStreamSynchronize(st)) && ok;
}

static bool liveMerge(PtScene* scene, cudaStream_t st)
{
    const size_t words = (size_t)scene->gridCells * (size_t)scene->view.emissiveCount;
    if (words == 0 || scene->replicas.empty()) return true;
    bool ok = true;
    for (PtScene_s* r : scene->replicas)
    {
        // the replica's delta (left in its liveBase by its thread) arrives in the primary's liveBase, which the
        // primary no longer needs once its own kernels are done
        ok = ok && PT_CUDA(cudaMemcpyPeerAsync(scene->liveBase, scene->device, r->liveBase, r->device, 4 * words, st));
        ok = ok && launchU32Add(scene->view.distLive, scene->liveBase, (int64_t)words, st);
    }
    for (PtScene_s* r : scene->replicas)
        ok = ok && PT_CUDA(cudaMemcpyPeerAsync(r->view.distLive, r->device, scene->view.distLive, scene->device, 4 * words, st));
    return ok;
}

static bool traceSppDevice(PtScene* scene, const Camera* camera, PtDofInfo* dof, int w, int h, int sppBegin, int sppCount,
                           float* dColor, float* dAlbedo, float* dNormal, cudaStream_t st, bool refresh = true)
{
    const bool own = st == scene->stream;
    if (refresh && !sceneRefresh(scene, st)) return false;       // PtScene_Update -> UpdateDists
    if (own) st = scene->stream;
    const int D = 1 + (int)scene->replicas.size();
    const int pixels = w * h;
    if (D == 1)
    {
        const TraceShare all = { sppBegin, sppCount, 0, pixels };
        return traceShare(scene, camera, dof, w, h, all, dColor, dAlbedo, dNormal, st);
    }
    // Shares: by sample index when every device gets at least one sample of the frame (each renders the whole
    // frame; the partial images are summed), else by pixel band (disjoint rows; a one-sample Pt_Trace frame)
    std::vector<TraceShare> share((size_t)D);
    const bool bySample = sppCount >= D;
    for (int k = 0; k < D; ++k)
    {
        if (bySample)
        {
            const int base = sppCount / D, rem = sppCount % D;
            share[(size_t)k] = { sppBegin + k * base + std::min(k, rem), base + (k < rem ? 1 : 0), 0, pixels };
        }
        else
        {
            const int rows = h / D, rem = h % D;
            const int y0 = k * rows + std::min(k, rem), ny = rows + (k < rem ? 1 : 0);
            share[(size_t)k] = { sppBegin, sppCount, y0 * w, ny * w };
        }
    }
    const size_t n3 = (size_t)pixels * 3;
    const int images = 1 + (dAlbedo ? 1 : 0) + (dNormal ? 1 : 0);
    bool ok = lightGridToReplicas(scene, st);
    ok = ok && forEachDevice(D, [&](int k) -> bool
    {
        if (k == 0) return liveSnapshot(scene, st) && traceShare(scene, camera, dof, w, h, share[0], dColor, dAlbedo, dNormal, st);
        PtScene_s* r = scene->replicas[(size_t)k - 1];
        DeviceGuard guard(r->device);
        cudaStream_t rs = r->stream;
        if (r->partialFloats < n3)
        {
            for (int i = 0; i < 3; ++i)
            {
                size_t have = 0;
                if (r->partial[i]) { cudaFree(r->partial[i]); r->partial[i] = nullptr; }
                if (!growBuffer(r->partial[i], have, n3)) { r->partialFloats = 0; return false; }
            }
            r->partialFloats = n3;
        }
        const TraceShare& sh = share[(size_t)k];
        const size_t off = (size_t)sh.pixelBegin * 3, cnt = (size_t)sh.pixelCount * 3;
        bool okr = liveSnapshot(r, rs);
        for (int i = 0; okr && i < images; ++i) okr = PT_CUDA(cudaMemsetAsync(r->partial[i] + off, 0, 4 * cnt, rs));
        okr = okr && traceShare(r, camera, dof, w, h, sh, r->partial[0], dAlbedo ? r->partial[1] : nullptr, dNormal ? r->partial[2] : nullptr, rs);
        const size_t words = (size_t)r->gridCells * (size_t)r->view.emissiveCount;
        if (okr && words) okr = launchU32Sub(r->liveBase, r->view.distLive, (int64_t)words, rs);
        return PT_CUDA(cudaStreamSynchronize(rs)) && okr;
    });
    // the partial images come to the primary by peer copy and are added (bands: only their rows)
    float* dst[3] = { dColor, dAlbedo, dNormal };
    for (int k = 1; ok && k < D; ++k)
    {
        PtScene_s* r = scene->replicas[(size_t)k - 1];
        const TraceShare& sh = share[(size_t)k];
        const size_t off = (size_t)sh.pixelBegin * 3, cnt = (size_t)sh.pixelCount * 3;
        if (cnt == 0 || sh.sppCount <= 0) continue;
        ok = growBuffer(scene->partial[0], scene->partialFloats, n3);
        for (int i = 0; ok && i < 3; ++i)
        {
            if (!dst[i]) continue;
            ok = PT_CUDA(cudaMemcpyPeerAsync(scene->partial[0], scene->device, r->partial[i] + off, r->device, 4 * cnt, st)) &&
                 launchAccumulate(dst[i] + off, scene->partial[0], (int64_t)cnt, st);
        }
    }
    ok = ok && liveMerge(scene, st);
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;      // the replicas' counters are current before the next call starts
    return ok;
}

// DofUpdate (path_tracer.c:2485-2503); the frame time is fixed at 1/60 s so results
// do not depend on wall-clock speed
static bool dofUpdate(PtScene* scene, PtDofInfo* dof, const Camera* camera)
{
    if (!dof->autoFocus) return true;
    const float q[4] = { camera->rotation.x, camera->rotation.y, camera->rotation.z, camera->rotation.w };
    V3 fwd = quatMulDir(q, v3(0.0f, 0.0f, -1.0f));
    F4 ro = f4(camera->position.x, camera->position.y, camera->position.z, 0.0f);
    F4 rd = f4(fwd, kRcpEpsilon);
    PtRayHit hit;
    if (Pt_IntersectN(scene, (const pim_float4*)&ro, (const pim_float4*)&rd, 0.0f, &hit, 1) != 0) return false;
    if (hit.type != PtHit_Nothing)
    {
        const float dt = 1.0f / 60.0f;
        float t = 1.0f - expf(-dt * dof->autoFocusSpeed);
        dof->focalLength = f1_lerp(dof->focalLength, hit.wuvt.w, t);
    }
    return true;
}

// the six images are one allocation: image[0..2] then sum[0..2], each w*h*3 floats, so that a frame zeroes the
// sums with one memset, folds them with one launch and, when the host images are adjacent too (PtTrace_New
// makes them so), copies them back with one copy
static void freeImages(DeviceImages& im)
{
    if (im.image[0]) cudaFree(im.image[0]);
    for (int k = 0; k < 3; ++k) im.image[k] = im.sum[k] = nullptr;
}

// The running-mean images of a PtTrace live on the device (the host arrays are refreshed from
// them at the end of every Pt_Trace / Pt_TraceSpp; edits the host makes to them are not read back).
static DeviceImages* traceImages(PtTrace* trace)
{
    std::lock_guard<std::mutex> lock(g_mutex);
    auto it = g_traces.find(trace);
    const int64_t texels = (int64_t)trace->imageSize.x * trace->imageSize.y;
    if (it != g_traces.end() && it->second.texels == texels) return &it->second;
    if (it != g_traces.end()) { freeImages(it->second); g_traces.erase(it); }      // the image size changed
    DeviceImages im;
    im.texels = texels;
    float* block = nullptr;
    const size_t n3 = 3 * (size_t)texels;
    if (!PT_CUDA(cudaMalloc((void**)&block, sizeof(float) * 6 * n3))) return nullptr;
    if (!PT_CUDA(cudaMemset(block, 0, sizeof(float) * 3 * n3))) { cudaFree(block); return nullptr; }
    for (int k = 0; k < 3; ++k) { im.image[k] = block + (size_t)k * n3; im.sum[k] = block + (size_t)(3 + k) * n3; }
    g_traces[trace] = im;
    return &g_traces[trace];
}

static void releaseImages(PtTrace* trace)
{
    std::lock_guard<std::mutex> lock(g_mutex);
    auto it = g_traces.find(trace);
    if (it == g_traces.end()) return;
    freeImages(it->second);
    g_traces.erase(it);
}

static bool onSceneDevice(const PtScene* scene, const char* who)
{
    int dev = 0;
    if (!PT_CUDA(cudaGetDevice(&dev))) return false;
    if (dev != scene->device)
    {
        setError("%s: the scene lives on device %d but device %d is current", who, scene->device, dev);
        return false;
    }
    return true;
}

// ================================================================== C ABI =====

extern "C" {

const char* Pt_LastError(void) { return t_error; }

void PtSys_Init(void) { t_error[0] = 0; ensureInit(); }
void PtSys_Update(void) {}
void PtSys_Shutdown(void)
{
    std::vector<PtTrace*> keys;
    { std::lock_guard<std::mutex> lock(g_mutex); for (auto& kv : g_traces) keys.push_back(kv.first); }
    for (PtTrace* t : keys) releaseImages(t);
    // staging arenas of every device that was used
    int count = 0, cur = 0;
    if (cudaGetDeviceCount(&count) == cudaSuccess && cudaGetDevice(&cur) == cudaSuccess)
    {
        for (int d = 0; d < count && d < kMaxDevices; ++d)
        {
            Arena& A = stagingArena(d);
            CopyRing& R = copyRing(d);
            std::lock_guard<std::mutex> lock(A.mutex);
            std::lock_guard<std::mutex> ringLock(R.mutex);
            if (!A.dev && !A.host && !A.evA && !R.slots[0]) continue;
            cudaSetDevice(d);
            A.release();
            R.release();
        }
        cudaSetDevice(cur);
    }
    HostPool::get().stop();
    g_sceneDesc.clear();
    g_lmPack.clear();
    g_initialised = false;
}

void PtSys_SetSceneDesc(const PtSceneDesc* desc) { g_sceneDesc.assign(desc); }

int32_t PtSys_SetDevices(int32_t count)
{
    if (!ensureInit()) return -1;
    int have = 0;
    if (!PT_CUDA(cudaGetDeviceCount(&have))) return -1;
    if (count < 1) count = 1;
    if (count > have || count > kMaxDevices) { setError("PtSys_SetDevices: %d devices asked for, %d present", count, have); return -1; }
    // peer access both ways between the devices in use, where the hardware has it (NVLink / NVSwitch here);
    // without it the peer copies still work, staged through the host by the driver
    for (int a = 0; a < count; ++a)
    {
        DeviceGuard guard(a);
        for (int b = 0; b < count; ++b)
        {
            int can = 0;
            if (a == b || cudaDeviceCanAccessPeer(&can, a, b) != cudaSuccess || !can) continue;
            const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); }
            else if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        }
    }
    g_deviceCount = count;
    return count;
}
int32_t PtSys_GetDevices(void) { return g_deviceCount; }

PtScene* PtScene_New(void)
{
    if (!g_sceneDesc.valid) { setError("PtScene_New: no scene description registered (PtSys_SetSceneDesc)"); return nullptr; }
    PtScene* s = sceneBuild(&g_sceneDesc.desc);
    if (s) s->descVersion = g_sceneDesc.version;
    return s;
}
PtScene* PtScene_NewFromDesc(const PtSceneDesc* desc) { return sceneBuild(desc); }

void PtScene_Update(PtScene* scene)
{
    if (!scene) return;
    if (sceneRefresh(scene, scene->stream)) PT_CUDA(cudaStreamSynchronize(scene->stream));
}
void PtScene_Del(PtScene* scene) { sceneFree(scene); }
void PtScene_Gui(PtScene* scene) { (void)scene; }

void PtTrace_New(PtTrace* trace, pim_int2 imageSize)      // path_tracer.c:1117-1130
{
    releaseImages(trace);       // a PtTrace re-initialised at the same address starts from empty device images
    memset(trace, 0, sizeof(*trace));
    const size_t texels = (size_t)imageSize.x * (size_t)imageSize.y;
    trace->imageSize = imageSize;
    trace->sampleWeight = 1.0f;
    if (!ensureInit() || texels == 0) return;
    // one page-locked block, carved into the four images (color is its base: PtTrace_Del frees that)
    const size_t bytes = sizeof(float) * 3 * texels;
    char* block = nullptr;
    if (!PT_CUDA(cudaMallocHost((void**)&block, 4 * bytes))) return;
    memset(block, 0, 4 * bytes);
    trace->color = (pim_float3*)block; trace->albedo = (pim_float3*)(block + bytes);
    trace->normal = (pim_float3*)(block + 2 * bytes); trace->denoised = (pim_float3*)(block + 3 * bytes);
}

void PtTrace_Del(PtTrace* trace)
{
    if (!trace) return;
    releaseImages(trace);
    if (trace->color) cudaFreeHost(trace->color);       // the base of the block PtTrace_New carved the four images from
    memset(trace, 0, sizeof(*trace));
}

void DofInfo_New(PtDofInfo* dof)      // path_tracer.c:1141-1153
{
    if (!dof) return;
    dof->aperture = 5.0f * kMilli;
    dof->focalLength = 6.0f;
    dof->bladeCount = 5;
    dof->bladeRot = kPi / 10.0f;
    dof->focalPlaneCurvature = 0.05f;
    dof->autoFocus = true;
    dof->autoFocusSpeed = 3.0f;
}
void DofInfo_Gui(PtDofInfo* dof) { (void)dof; }

int32_t Pt_IntersectN_Device(const PtScene* scene, const void* dRo, const void* dRd, float tNear, void* dHitsOut,
                             int64_t n, void* stream)
{
    if (!scene) { setError("Pt_IntersectN_Device: null scene"); return -1; }
    cudaStream_t st = stream ? (cudaStream_t)stream : scene->stream;
    return launchIntersect(scene->view, (const F4*)dRo, (const F4*)dRd, tNear, (RayHitOut*)dHitsOut, n, scene->nextCursor(), st) ? 0 : -1;
}

int32_t Pt_OccludedN_Device(const PtScene* scene, const void* dRo, const void* dRd, float tNear, void* dOccludedOut,
                            int64_t n, void* stream)
{
    if (!scene) { setError("Pt_OccludedN_Device: null scene"); return -1; }
    cudaStream_t st = stream ? (cudaStream_t)stream : scene->stream;
    return launchOccluded(scene->view, (const F4*)dRo, (const F4*)dRd, tNear, (uint8_t*)dOccludedOut, n, scene->nextCursor(), st) ? 0 : -1;
}

int32_t Pt_IntersectN(const PtScene* scene, const pim_float4* ro, const pim_float4* rd, float tNear, PtRayHit* hitsOut,
                      int64_t n)
{
    if (!scene) { setError("Pt_IntersectN: null scene"); return -1; }
    if (n <= 0) return 0;
    const int64_t chunk = 1 << 24;
    const int64_t m = std::min(n, chunk);
    F4 *dRo = nullptr, *dRd = nullptr; RayHitOut* dOut = nullptr;
    Lease lease;
    if (!lease.dev(dRo, (size_t)m).dev(dRd, (size_t)m).dev(dOut, (size_t)m).commit()) return -1;
    cudaStream_t st = scene->stream;
    for (int64_t base = 0; base < n; base += chunk)
    {
        const int64_t c = std::min(chunk, n - base);
        bool ok = copyIn(dRo, ro + base, sizeof(F4) * c, st) &&
                  copyIn(dRd, rd + base, sizeof(F4) * c, st) &&
                  launchIntersect(scene->view, dRo, dRd, tNear, dOut, c, scene->nextCursor(), st) &&
                  copyOut(hitsOut + base, dOut, sizeof(RayHitOut) * c, st) &&
                  PT_CUDA(cudaStreamSynchronize(st));
        if (!ok) { cudaStreamSynchronize(st); return -1; }
    }
    return 0;
}

int32_t Pt_OccludedN(const PtScene* scene, const pim_float4* ro, const pim_float4* rd, float tNear, uint8_t* occludedOut,
                     int64_t n)
{
    if (!scene) { setError("Pt_OccludedN: null scene"); return -1; }
    if (n <= 0) return 0;
    const int64_t chunk = 1 << 24;
    const int64_t m = std::min(n, chunk);
    F4 *dRo = nullptr, *dRd = nullptr; uint8_t* dOut = nullptr;
    Lease lease;
    if (!lease.dev(dRo, (size_t)m).dev(dRd, (size_t)m).dev(dOut, (size_t)m).commit()) return -1;
    cudaStream_t st = scene->stream;
    for (int64_t base = 0; base < n; base += chunk)
    {
        const int64_t c = std::min(chunk, n - base);
        bool ok = copyIn(dRo, ro + base, sizeof(F4) * c, st) &&
                  copyIn(dRd, rd + base, sizeof(F4) * c, st) &&
                  launchOccluded(scene->view, dRo, dRd, tNear, dOut, c, scene->nextCursor(), st) &&
                  copyOut(occludedOut + base, dOut, c, st) &&
                  PT_CUDA(cudaStreamSynchronize(st));
        if (!ok) { cudaStreamSynchronize(st); return -1; }
    }
    return 0;
}

int32_t Pt_PointQueryN(const PtScene* scene, const pim_float4* points, float* distanceOut, int64_t n)
{
    if (!scene) { setError("Pt_PointQueryN: null scene"); return -1; }
    if (n <= 0) return 0;
    F4* dPts = nullptr; float* dDist = nullptr;
    Lease lease;
    if (!lease.dev(dPts, (size_t)n).dev(dDist, (size_t)n).commit()) return -1;
    cudaStream_t st = scene->stream;
    bool ok = copyIn(dPts, points, sizeof(F4) * n, st) &&
              launchPointQuery(scene->view, dPts, dDist, n, st) &&
              copyOut(distanceOut, dDist, sizeof(float) * n, st);
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

PtRayHit Pt_Intersect(const PtScene* scene, pim_float4 ro, pim_float4 rd, float tNear, float tFar)
{
    PtRayHit hit;
    memset(&hit, 0, sizeof(hit));
    hit.wuvt.w = -1.0f; hit.iVert = -1;
    rd.w = tFar;
    Pt_IntersectN(scene, &ro, &rd, tNear, &hit, 1);
    return hit;
}

int32_t Pt_TraceRayN(PtScene* scene, const pim_float4* ro, const pim_float4* rd, uint32_t pass, PtResult* resultsOut,
                     int64_t n)
{
    if (!scene) { setError("Pt_TraceRayN: null scene"); return -1; }
    if (n <= 0) return 0;
    const bool small = n < 4096;
    F4 *dRo = nullptr, *dRd = nullptr; float *dOut = nullptr, *c = nullptr, *a = nullptr, *nr = nullptr;
    Lease lease;
    lease.dev(dRo, (size_t)n).dev(dRd, (size_t)n).dev(dOut, 9 * (size_t)n);
    if (!small) lease.dev(c, 3 * (size_t)n).dev(a, 3 * (size_t)n).dev(nr, 3 * (size_t)n);
    if (!lease.commit()) return -1;
    cudaStream_t st = scene->stream;
    bool ok = copyIn(dRo, ro, sizeof(F4) * n, st) &&
              copyIn(dRd, rd, sizeof(F4) * n, st);
    if (ok && small)
    {
        ok = launchTraceRays(scene->view, dRo, dRd, n, pass, scene->config.seed, dOut, st);
    }
    else if (ok)
    {
        // results are per item: colour, albedo, normal triples interleaved as PtResult
        WorkDesc W; memset(&W, 0, sizeof(W));
        W.kind = Work_Rays; W.ro = dRo; W.rd = dRd; W.count = n;
        W.passBase = pass; W.seed = scene->config.seed; W.salt = Salt_Rays;
        W.sumColor = c; W.sumAlbedo = a; W.sumNormal = nr; W.accumulate = 0;
        float ms = 0.0f;
        ok = runWavefront(scene->view, scene->pool, W, st, &ms);
        g_lastKernelMs = ms;
        // interleave on the device side of the copy: three strided 2D copies into PtResult
        ok = ok && PT_CUDA(cudaMemcpy2DAsync(dOut + 0, 36, c, 12, 12, (size_t)n, cudaMemcpyDeviceToDevice, st));
        ok = ok && PT_CUDA(cudaMemcpy2DAsync(dOut + 3, 36, a, 12, 12, (size_t)n, cudaMemcpyDeviceToDevice, st));
        ok = ok && PT_CUDA(cudaMemcpy2DAsync(dOut + 6, 36, nr, 12, 12, (size_t)n, cudaMemcpyDeviceToDevice, st));
    }
    ok = ok && copyOut(resultsOut, dOut, sizeof(float) * 9 * n, st);
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

int32_t Pt_TraceRayN_NoAlloc(PtScene* scene, const void* dRo, const void* dRd, uint32_t pass, void* dColor3,
                             void* dAlbedo3, void* dNormal3, int64_t n, void* stream)
{
    if (!scene || (n > 0 && (!dRo || !dRd || !dColor3))) { setError("Pt_TraceRayN_NoAlloc: null argument"); return -1; }
    if (n > PT_NOALLOC_MAX_PATHS)
    {
        setError("Pt_TraceRayN_NoAlloc: n = %lld is beyond PT_NOALLOC_MAX_PATHS = %d; split the batch or use Pt_TraceRayN",
                 (long long)n, PT_NOALLOC_MAX_PATHS);
        return -1;
    }
    // the scratch array is per device: the one resolved is the current device's, which must be the scene's
    if (!onSceneDevice(scene, "Pt_TraceRayN_NoAlloc")) return -1;
    cudaStream_t st = stream ? (cudaStream_t)stream : scene->stream;
    return runWavefrontNoAlloc(scene->view, (const F4*)dRo, (const F4*)dRd, n, pass, scene->config.seed, (float*)dColor3,
                               (float*)dAlbedo3, (float*)dNormal3, st) ? 0 : -1;
}

PtResult Pt_TraceRay(PtScene* scene, pim_float4 ro, pim_float4 rd)
{
    PtResult r;
    memset(&r, 0, sizeof(r));
    if (scene) Pt_TraceRayN(scene, &ro, &rd, scene->nextPass++, &r, 1);
    return r;
}

int32_t Pt_TraceSpp_Device(PtScene* scene, const Camera* camera, PtDofInfo* dof, int32_t width, int32_t height,
                           int32_t sppBegin, int32_t sppCount, void* dSumColor, void* dSumAlbedo, void* dSumNormal,
                           void* stream)
{
    if (!scene || !camera || !dof) { setError("Pt_TraceSpp_Device: null argument"); return -1; }
    cudaStream_t st = stream ? (cudaStream_t)stream : scene->stream;
    return traceSppDevice(scene, camera, dof, width, height, sppBegin, sppCount, (float*)dSumColor, (float*)dSumAlbedo,
                          (float*)dSumNormal, st) ? 0 : -1;
}

int32_t Pt_TraceSpp(PtTrace* trace, PtDofInfo* dof, PtScene* scene, const Camera* camera, int32_t sppBegin, int32_t sppCount)
{
    if (!trace || !dof || !scene || !camera || !trace->color) { setError("Pt_TraceSpp: null argument"); return -1; }
    if (sppCount <= 0) return 0;
    DeviceImages* im = traceImages(trace);
    if (!im) return -1;
    const int w = trace->imageSize.x, h = trace->imageSize.y;
    const int64_t n3 = im->texels * 3;
    bool ok = true;
    ok = ok && PT_CUDA(cudaMemsetAsync(im->sum[0], 0, sizeof(float) * 3 * n3, scene->stream));      // the three sums are adjacent
    ok = ok && traceSppDevice(scene, camera, dof, w, h, sppBegin, sppCount, im->sum[0], im->sum[1], im->sum[2], scene->stream);
    cudaStream_t st = scene->stream;
    float* host[3] = { (float*)trace->color, (float*)trace->albedo, (float*)trace->normal };
    for (int k = 0; ok && k < 3; ++k)
    {
        if (trace->sampleWeight > 0.0f)
        {
            ok = launchResolve(im->image[k], im->sum[k], n3, (float)sppBegin, (float)sppCount, st);
            ok = ok && PT_CUDA(cudaMemcpyAsync(host[k], im->image[k], sizeof(float) * n3, cudaMemcpyDeviceToHost, st));
        }
        else
        {
            ok = PT_CUDA(cudaMemcpyAsync(host[k], im->sum[k], sizeof(float) * n3, cudaMemcpyDeviceToHost, st));
        }
    }
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

void Pt_Trace(PtTrace* trace, PtDofInfo* dof, PtScene* scene, const Camera* camera)      // path_tracer.c:2557-2585
{
    t_error[0] = 0;
    if (!trace || !dof || !scene || !camera || !trace->color) { setError("Pt_Trace: null argument"); return; }
    DeviceImages* im = traceImages(trace);
    if (!im) return;
    const int w = trace->imageSize.x, h = trace->imageSize.y;
    const int64_t n3 = im->texels * 3;
    if (!sceneRefresh(scene, scene->stream)) return;    // PtScene_Update first (:2573), so that DofUpdate sees the current scene
    if (!dofUpdate(scene, dof, camera)) return;
    cudaStream_t st = scene->stream;
    bool ok = true;
    ok = ok && PT_CUDA(cudaMemsetAsync(im->sum[0], 0, sizeof(float) * 3 * n3, st));      // the three sums are adjacent
    // the CDFs were relearned just above: trace without a second PtScene_Update
    ok = ok && traceSppDevice(scene, camera, dof, w, h, (int)scene->nextPass, 1, im->sum[0], im->sum[1], im->sum[2], st, false);
    scene->nextPass++;
    float* host[3] = { (float*)trace->color, (float*)trace->albedo, (float*)trace->normal };
    ok = ok && launchLerp(im->image[0], im->sum[0], 3 * n3, trace->sampleWeight, st);           // and so are the three means
    if (ok && host[1] == host[0] + n3 && host[2] == host[1] + n3)
        ok = PT_CUDA(cudaMemcpyAsync(host[0], im->image[0], sizeof(float) * 3 * n3, cudaMemcpyDeviceToHost, st));
    else
        for (int k = 0; ok && k < 3; ++k) ok = PT_CUDA(cudaMemcpyAsync(host[k], im->image[k], sizeof(float) * n3, cudaMemcpyDeviceToHost, st));
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
}

PtResults Pt_RayGen(PtScene* scene, pim_float4 origin, int32_t count)      // path_tracer.c:2615-2644
{
    PtResults res = { nullptr, nullptr };
    if (!scene || count <= 0) return res;
    const int slot = scene->rayGenFlip;
    scene->rayGenFlip ^= 1;
    if (scene->rayGenCap[slot] < count)
    {
        if (scene->rayGenColors[slot]) cudaFreeHost(scene->rayGenColors[slot]);
        if (scene->rayGenDirs[slot]) cudaFreeHost(scene->rayGenDirs[slot]);
        scene->rayGenColors[slot] = nullptr; scene->rayGenDirs[slot] = nullptr; scene->rayGenCap[slot] = 0;
        if (!PT_CUDA(cudaMallocHost((void**)&scene->rayGenColors[slot], sizeof(F4) * (size_t)count))) return res;
        if (!PT_CUDA(cudaMallocHost((void**)&scene->rayGenDirs[slot], sizeof(F4) * (size_t)count))) return res;
        scene->rayGenCap[slot] = count;
    }
    float *c = nullptr, *hc = nullptr; F4* d = nullptr;
    Lease lease;
    if (!lease.dev(c, 3 * (size_t)count).dev(d, (size_t)count).host(hc, 3 * (size_t)count).commit()) return res;
    bool ok = sceneRefresh(scene, scene->stream);
    cudaStream_t st = scene->stream;
    WorkDesc W; memset(&W, 0, sizeof(W));
    W.kind = Work_Sphere; W.origin = v3(origin.x, origin.y, origin.z); W.count = count;
    W.passBase = scene->nextPass++; W.seed = scene->config.seed; W.salt = Salt_RayGen;
    W.sumColor = c; W.directions = d; W.accumulate = 0;
    float ms = 0.0f;
    ok = ok && runWavefront(scene->view, scene->pool, W, st, &ms);
    g_lastKernelMs = ms;
    // colours come back as float3; widen to the reference's float4 (w = 0)
    ok = ok && PT_CUDA(cudaMemcpyAsync(scene->rayGenDirs[slot], d, sizeof(F4) * (size_t)count, cudaMemcpyDeviceToHost, st));
    ok = ok && PT_CUDA(cudaMemcpyAsync(hc, c, sizeof(float) * 3 * (size_t)count, cudaMemcpyDeviceToHost, st));
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    if (!ok) return res;
    for (int32_t i = 0; i < count; ++i)
    {
        scene->rayGenColors[slot][i].x = hc[3 * (size_t)i]; scene->rayGenColors[slot][i].y = hc[3 * (size_t)i + 1];
        scene->rayGenColors[slot][i].z = hc[3 * (size_t)i + 2]; scene->rayGenColors[slot][i].w = 0.0f;
    }
    res.colors = scene->rayGenColors[slot];
    res.directions = scene->rayGenDirs[slot];
    return res;
}

// one page whose arrays are already on the device
static bool bakePageDevice(PtScene* scene, const LightmapView& view, float texelsPerMeter, float timeSlice, int32_t spp,
                           uint32_t pass, cudaStream_t st)
{
    if (scene->config.bakeMode == PtBake_Sequential)
        return launchLightmapBake(scene->view, view, 1.0f / texelsPerMeter, timeSlice, spp, pass, scene->config.seed, st);
    return launchLightmapBakeWavefront(scene->view, scene->pool, view, 1.0f / texelsPerMeter, timeSlice, spp, pass,
                                       scene->config.seed, st);
}

// device side of one device's share of a host page (LmPack_BakeDesc): the full-size page and the covered texels' records
struct LmDeviceBuffers
{
    F4* probes = nullptr; float *pos = nullptr, *nrm = nullptr, *cnt = nullptr;                 // the full-size device page
    LmTexelRecord* records = nullptr;
};
static void lmLeaseDevice(Lease& lease, LmDeviceBuffers& b, size_t texels)
{
    lease.dev(b.probes, 5 * texels).dev(b.pos, 3 * texels).dev(b.nrm, 3 * texels).dev(b.cnt, texels).dev(b.records, texels);
}
// covered texels [j0, j0 + count) of the page: up in one copy, scatter, bake, gather, down in one copy (into the
// same page-locked records)
static bool lmBakeShare(PtScene_s* s, const LmDeviceBuffers& b, size_t n, int32_t texelBase, int32_t j0, int32_t count,
                        LmTexelRecord* hRecords, float texelsPerMeter, float timeSlice, int32_t spp, uint32_t pass)
{
    if (count <= 0) return true;
    cudaStream_t st = s->stream;
    LightmapView view;
    for (int k = 0; k < 5; ++k) view.probes[k] = b.probes + (size_t)k * n;
    view.position = b.pos; view.normal = b.nrm; view.sampleCounts = b.cnt;
    view.texelCount = (int32_t)n; view.texelBase = texelBase;
    const size_t bytes = sizeof(LmTexelRecord) * (size_t)count;
    bool ok = PT_CUDA(cudaMemsetAsync(b.cnt, 0, sizeof(float) * n, st)) &&
              PT_CUDA(cudaMemcpyAsync(b.records, hRecords + j0, bytes, cudaMemcpyHostToDevice, st)) &&
              launchLightmapScatter(view, b.records, count, st);
    if (ok)
    {
        ok = s->config.bakeMode == PtBake_Sequential
                 ? launchLightmapBake(s->view, view, 1.0f / texelsPerMeter, timeSlice, spp, pass, s->config.seed, st)
                 : launchLightmapBakeWavefront(s->view, s->pool, view, 1.0f / texelsPerMeter, timeSlice, spp, pass, s->config.seed, st);
    }
    ok = ok && launchLightmapGather(view, b.records, count, st) &&
         PT_CUDA(cudaMemcpyAsync(hRecords + j0, b.records, bytes, cudaMemcpyDeviceToHost, st));
    return PT_CUDA(cudaStreamSynchronize(st)) && ok;
}

// Host pages move sparsely. BakeFn skips every texel whose sampleCounts entry is 0 (lightmap.c:1145-1149)
// — the texels no chart covers, usually most of an atlas page (cornell at 16 texels/m: 54 437 of 1 048 576)
// — so only the covered texels' 108 bytes travel: the host compacts them into page-locked staging, one
// copy takes them to the device, a kernel scatters them into a full-size device page whose sampleCounts
// were zeroed, the bake runs on that page exactly as it would on a resident one, and probes + counts of
// the same texels come back the same way. A dense page costs the same bytes as a plain copy plus 4 B of
// index per texel; the cornell page costs 6.1 MB each way instead of 113 MB.
int32_t LmPack_BakeDesc(PtScene* scene, PtLightmapDesc* lightmaps, int32_t lmCount, float texelsPerMeter, float timeSlice,
                        int32_t spp, uint32_t pass)
{
    if (!scene || (!lightmaps && lmCount > 0)) { setError("LmPack_BakeDesc: null argument"); return -1; }
    if (lmCount <= 0) return 0;
    size_t maxTexels = 0;
    for (int32_t i = 0; i < lmCount; ++i)
    {
        const PtLightmapDesc& lm = lightmaps[i];
        if (lm.size <= 0 || !lm.position || !lm.normal || !lm.sampleCounts || !lm.probes[0] || !lm.probes[1] || !lm.probes[2] ||
            !lm.probes[3] || !lm.probes[4]) { setError("LmPack_BakeDesc: page %d has no storage", i); return -1; }
        maxTexels = std::max(maxTexels, (size_t)lm.size * (size_t)lm.size);
    }
    // worst case: every texel of the largest page is covered
    LmTexelRecord* hRecords = nullptr; int32_t* hIdx = nullptr;
    LmDeviceBuffers primary;
    Lease lease;
    lmLeaseDevice(lease, primary, maxTexels);
    lease.host(hRecords, maxTexels).host(hIdx, maxTexels);
    if (!lease.commit()) return -1;
    bool ok = sceneRefresh(scene, scene->stream);       // LmPack_Bake -> PtScene_Update (lightmap.c:1209)
    const int D = 1 + (int)scene->replicas.size();
    // the replicas take the primary's light grid (it relearned in sceneRefresh), then every device notes its
    // light-learning counters so that what the bake adds on each can be merged afterwards (liveMerge)
    if (D > 1)
    {
        ok = ok && lightGridToReplicas(scene, scene->stream);
        ok = ok && forEachDevice(D, [&](int k) -> bool
        {
            PtScene_s* s = k == 0 ? scene : scene->replicas[(size_t)k - 1];
            DeviceGuard guard(s->device);
            const bool okd = liveSnapshot(s, s->stream);
            return PT_CUDA(cudaStreamSynchronize(s->stream)) && okd;
        });
    }
    int32_t texelBase = 0;
    for (int32_t i = 0; ok && i < lmCount; ++i)
    {
        PtLightmapDesc& lm = lightmaps[i];
        const size_t n = (size_t)lm.size * (size_t)lm.size;
        // covered texels, compacted on the host by the resident host threads in three steps: every thread counts
        // the covered texels of its slice of the page; writes their indices (a dense int array) behind those of
        // the slices before it; then packs an equal share of the covered texels' records (charts cluster in the
        // atlas: packing by slice left nearly all of the records, 6 MB of cache lines, to one or two threads)
        HostPool& hostPool = HostPool::get();
        hostPool.start();
        const float* pos = (const float*)lm.position; const float* nrm = (const float*)lm.normal;
        std::vector<int32_t> sliceCount((size_t)hostPool.workers() + 1, 0);
        hostPool.run([&](int w, int workers)
        {
            size_t t0, t1; hostShare(n, w, workers, t0, t1);
            int32_t c = 0;
            for (size_t t = t0; t < t1; ++t) c += lm.sampleCounts[t] != 0.0f ? 1 : 0;
            sliceCount[(size_t)w + 1] = c;
        });
        for (size_t w = 1; w < sliceCount.size(); ++w) sliceCount[w] += sliceCount[w - 1];
        const int32_t m = sliceCount.back();
        hostPool.run([&](int w, int workers)
        {
            size_t t0, t1; hostShare(n, w, workers, t0, t1);
            int32_t j = sliceCount[(size_t)w];
            for (size_t t = t0; t < t1; ++t) if (lm.sampleCounts[t] != 0.0f) hIdx[j++] = (int32_t)t;
        });
        hostPool.run([&](int w, int workers)
        {
            size_t j0, j1; hostShare((size_t)m, w, workers, j0, j1);
            for (size_t j = j0; j < j1; ++j)
            {
                alignas(16) LmTexelRecord r;
                const size_t t = (size_t)hIdx[j];
                r.idx = hIdx[j];
                for (int k = 0; k < 5; ++k) r.probes[k] = *(const F4*)&lm.probes[k][t];
                r.pos[0] = pos[3 * t]; r.pos[1] = pos[3 * t + 1]; r.pos[2] = pos[3 * t + 2];
                r.nrm[0] = nrm[3 * t]; r.nrm[1] = nrm[3 * t + 1]; r.nrm[2] = nrm[3 * t + 2];
                r.cnt = lm.sampleCounts[t];
                // non-temporal stores: the DMA engine reads the records next, and lines left dirty in eight cores'
                // caches made that copy 6.6x slower (6.1 MB: 0.81 ms against 0.12 ms)
                streamStore16(&hRecords[j], &r, sizeof(LmTexelRecord));
            }
        });
        if (m > 0)
        {
            // one contiguous range of the covered texels per device (the fit is sequential per texel, so texels,
            // never samples, are what devices share out); a texel's random streams follow its index in the pack,
            // so the result does not depend on how many devices baked it
            ok = forEachDevice(D, [&](int k) -> bool
            {
                const int32_t base = m / D, rem = m % D;
                const int32_t j0 = k * base + std::min(k, rem), cnt = base + (k < rem ? 1 : 0);
                PtScene_s* s = k == 0 ? scene : scene->replicas[(size_t)k - 1];
                if (k == 0) return lmBakeShare(s, primary, n, texelBase, j0, cnt, hRecords, texelsPerMeter, timeSlice, spp, pass);
                DeviceGuard guard(s->device);
                LmDeviceBuffers mine;
                Lease own;                  // this device's arena
                lmLeaseDevice(own, mine, std::max<size_t>(n, 1));
                if (!own.commit()) return false;
                return lmBakeShare(s, mine, n, texelBase, j0, cnt, hRecords, texelsPerMeter, timeSlice, spp, pass);
            });
            if (ok)
            {
                hostPool.run([&](int w, int workers)
                {
                    size_t j0, j1; hostShare((size_t)m, w, workers, j0, j1);
                    for (size_t j = j0; j < j1; ++j)
                    {
                        const LmTexelRecord& r = hRecords[j];
                        const size_t t = (size_t)r.idx;
                        for (int k = 0; k < 5; ++k) *(F4*)&lm.probes[k][t] = r.probes[k];
                        lm.sampleCounts[t] = r.cnt;
                    }
                });
            }
        }
        texelBase += (int32_t)n;
    }
    if (ok && D > 1)
    {
        ok = forEachDevice(D, [&](int k) -> bool
        {
            if (k == 0) return true;
            PtScene_s* s = scene->replicas[(size_t)k - 1];
            DeviceGuard guard(s->device);
            const size_t words = (size_t)s->gridCells * (size_t)s->view.emissiveCount;
            bool okd = words == 0 || launchU32Sub(s->liveBase, s->view.distLive, (int64_t)words, s->stream);
            return PT_CUDA(cudaStreamSynchronize(s->stream)) && okd;
        });
        ok = ok && liveMerge(scene, scene->stream);
    }
    ok = PT_CUDA(cudaStreamSynchronize(scene->stream)) && ok;
    return ok ? 0 : -1;
}

int32_t LmPack_BakeDesc_Device(PtScene* scene, const PtLightmapDesc* lightmaps, int32_t lmCount, float texelsPerMeter,
                               float timeSlice, int32_t spp, uint32_t pass)
{
    if (!scene || (!lightmaps && lmCount > 0)) { setError("LmPack_BakeDesc_Device: null argument"); return -1; }
    Lease lease;                                // only for the timer's events and the lock
    if (!lease.commit()) return -1;
    bool ok = sceneRefresh(scene, scene->stream);
    cudaStream_t st = scene->stream;
    int32_t texelBase = 0;
    KernelTimer timer(lease, st);
    for (int32_t i = 0; ok && i < lmCount; ++i)
    {
        const PtLightmapDesc& lm = lightmaps[i];
        const size_t n = (size_t)lm.size * (size_t)lm.size;
        LightmapView view;
        for (int k = 0; k < 5; ++k) view.probes[k] = (F4*)lm.probes[k];
        view.position = (const float*)lm.position; view.normal = (const float*)lm.normal; view.sampleCounts = lm.sampleCounts;
        view.texelCount = (int32_t)n; view.texelBase = texelBase;
        if (!view.probes[0] || !view.position || !view.normal || !view.sampleCounts) { setError("LmPack_BakeDesc_Device: page %d has no storage", i); return -1; }
        ok = bakePageDevice(scene, view, texelsPerMeter, timeSlice, spp, pass, st);
        texelBase += (int32_t)n;
    }
    timer.stop();
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

// LmPack_Bake under the reference's name and signature (src/rendering/lightmap.h:69, body lightmap.c:1204-1223).
// The reference bakes the engine's global pack (LmPack_Get()); PtSys_SetLmPack is the adapter's stand-in for
// that global. The pass index is the scene's running counter (the reference draws from its thread generators).
void PtSys_SetLmPack(const PtLightmapDesc* lightmaps, int32_t lmCount, float texelsPerMeter)
{
    g_lmPack.clear();
    g_lmTexelsPerMeter = texelsPerMeter;
    if (lightmaps && lmCount > 0) g_lmPack.assign(lightmaps, lightmaps + lmCount);
}

void LmPack_Bake(PtScene* scene, float timeSlice, int32_t spp)
{
    t_error[0] = 0;
    if (!scene) { setError("LmPack_Bake: null scene"); return; }
    if (g_lmPack.empty())
    {
        PtScene_Update(scene);      // an empty pack still relearns the light grid (lightmap.c:1209)
        return;
    }
    if (!(g_lmTexelsPerMeter > 0.0f)) { setError("LmPack_Bake: the registered pack has no texel density"); return; }
    LmPack_BakeDesc(scene, g_lmPack.data(), (int32_t)g_lmPack.size(), g_lmTexelsPerMeter, timeSlice, spp < 1 ? 1 : spp,
                    scene->nextPass++);
}

// ---- cubemap / ambient-cube bakes, sky bake, screenshot tonemap ----

void Cubemap_New(Cubemap* cm, int32_t size)       // cubemap.c:95-114
{
    if (!cm) return;
    memset(cm, 0, sizeof(*cm));
    if (size <= 0 || !ensureInit()) return;
    const size_t len = (size_t)size * size, elemCount = (size_t)cubemapElemCount(size);
    float* color = nullptr; float* conv = nullptr;
    if (!PT_CUDA(cudaMallocHost((void**)&color, sizeof(float) * 3 * len * 6))) return;
    if (!PT_CUDA(cudaMallocHost((void**)&conv, sizeof(float) * 4 * elemCount * 6))) { cudaFreeHost(color); return; }
    memset(color, 0, sizeof(float) * 3 * len * 6);
    memset(conv, 0, sizeof(float) * 4 * elemCount * 6);
    cm->size = size;
    cm->mipCount = cubemapMipCount(size);
    for (int f = 0; f < 6; ++f)
    {
        cm->color[f] = reinterpret_cast<pim_float3*>(color + 3 * len * f);
        cm->convolved[f] = reinterpret_cast<pim_float4*>(conv + 4 * elemCount * f);
    }
}

void Cubemap_Del(Cubemap* cm)                     // cubemap.c:116-127
{
    if (!cm) return;
    // the six faces of each kind are slices of one page-locked block
    if (cm->color[0]) cudaFreeHost(cm->color[0]);
    if (cm->convolved[0]) cudaFreeHost(cm->convolved[0]);
    memset(cm, 0, sizeof(*cm));
}

// faces may be separate host arrays (a caller-built Cubemap): copy face by face
static bool cubeUpload(const Cubemap* cm, float* dColor, float* dConv, cudaStream_t st)
{
    const size_t len = (size_t)cm->size * cm->size, elemCount = (size_t)cubemapElemCount(cm->size);
    bool ok = true;
    for (int f = 0; ok && f < 6; ++f)
    {
        if (dColor) ok = copyIn(dColor + 3 * len * f, cm->color[f], 12 * len, st);
        if (ok && dConv) ok = copyIn(dConv + 4 * elemCount * f, cm->convolved[f], 16 * elemCount, st);
    }
    return ok;
}
static bool cubeDownload(Cubemap* cm, const float* dColor, const float* dConv, cudaStream_t st)
{
    const size_t len = (size_t)cm->size * cm->size, elemCount = (size_t)cubemapElemCount(cm->size);
    bool ok = true;
    for (int f = 0; ok && f < 6; ++f)
    {
        if (dColor) ok = copyOut(cm->color[f], dColor + 3 * len * f, 12 * len, st);
        if (ok && dConv) ok = copyOut(cm->convolved[f], dConv + 4 * elemCount * f, 16 * elemCount, st);
    }
    return ok && PT_CUDA(cudaStreamSynchronize(st));
}
static bool cubeValid(const Cubemap* cm, const char* who, bool needConvolved)
{
    if (!cm || cm->size <= 0) { setError("%s: null or empty cubemap", who); return false; }
    for (int f = 0; f < 6; ++f)
    {
        if (!cm->color[f] || (needConvolved && !cm->convolved[f])) { setError("%s: cubemap face %d has no storage", who, f); return false; }
    }
    return true;
}

int32_t Cubemap_BakePass(Cubemap* cm, PtScene* scene, pim_float4 origin, float weight, uint32_t pass)
{
    if (!scene) { setError("Cubemap_Bake: null scene"); return -1; }
    if (!cubeValid(cm, "Cubemap_Bake", false)) return -1;
    const int64_t n = 6ll * cm->size * cm->size;
    cudaStream_t st = scene->stream;
    float *dColor = nullptr, *dRes = nullptr; F4 *dRo = nullptr, *dRd = nullptr;
    Lease lease;
    if (!lease.dev(dColor, 3 * (size_t)n).dev(dRo, (size_t)n).dev(dRd, (size_t)n).dev(dRes, 3 * (size_t)n).commit()) return -1;
    bool ok = cubeUpload(cm, dColor, nullptr, st);
    KernelTimer timer(lease, st);
    ok = ok && launchCubeRays(cm->size, v3(origin.x, origin.y, origin.z), pass, scene->config.seed, dRo, dRd, st);
    WorkDesc W; memset(&W, 0, sizeof(W));
    W.kind = Work_Rays; W.ro = dRo; W.rd = dRd; W.count = n;
    W.passBase = pass; W.seed = scene->config.seed; W.salt = Salt_CubeBake;
    W.sumColor = dRes; W.accumulate = 0;
    float ms = 0.0f;
    ok = ok && runWavefront(scene->view, scene->pool, W, st, &ms);
    ok = ok && launchLerp(dColor, dRes, 3 * n, weight, st);      // f3_lerpvs, cubemap.c:158
    timer.stop();
    ok = ok && cubeDownload(cm, dColor, nullptr, st);
    if (!ok) cudaStreamSynchronize(st);
    return ok ? 0 : -1;
}

void Cubemap_Bake(Cubemap* cm, PtScene* scene, pim_float4 origin, float weight)
{
    if (scene) Cubemap_BakePass(cm, scene, origin, weight, scene->nextPass++);
}

static uint32_t g_convolvePass = 0;

int32_t Cubemap_ConvolvePass(Cubemap* cm, uint32_t sampleCount, float weight, uint32_t seed, uint32_t pass)
{
    if (!ensureInit()) return -1;
    if (!cubeValid(cm, "Cubemap_Convolve", true)) return -1;
    const size_t len = (size_t)cm->size * cm->size, elemCount = (size_t)cubemapElemCount(cm->size);
    float *dColor = nullptr, *dConv = nullptr;
    Lease lease;
    if (!lease.dev(dColor, 3 * len * 6).dev(dConv, 4 * elemCount * 6).commit()) return -1;
    cudaStream_t st = nullptr;
    bool ok = cubeUpload(cm, dColor, dConv, st);
    KernelTimer timer(lease, st);
    ok = ok && launchCubeConvolve(dColor, dConv, cm->size, sampleCount, weight, pass, seed, st);
    timer.stop();
    ok = ok && cubeDownload(cm, nullptr, dConv, st);
    if (!ok) cudaStreamSynchronize(st);
    return ok ? 0 : -1;
}

void Cubemap_Convolve(Cubemap* cm, uint32_t sampleCount, float weight)
{
    Cubemap_ConvolvePass(cm, sampleCount, weight, 1u, g_convolvePass++);
}

int32_t AmbCube_Bake(PtScene* scene, AmbCube_t* pCube, pim_float4 origin, int32_t samples, int32_t prevSampleCount)   // ambcube.c:5-33
{
    if (!scene || !pCube || samples < 0 || prevSampleCount < 0) { setError("AmbCube_Bake: bad argument"); return prevSampleCount; }
    if (samples > 0)
    {
        float* c = nullptr; F4 *d = nullptr, *cube = nullptr;
        Lease lease;
        if (!lease.dev(c, 3 * (size_t)samples).dev(d, (size_t)samples).dev(cube, 6).commit()) return prevSampleCount;
        // Pt_RayGen (path_tracer.c:2615-2644), results left on the device
        bool ok = sceneRefresh(scene, scene->stream);
        cudaStream_t st = scene->stream;
        WorkDesc W; memset(&W, 0, sizeof(W));
        W.kind = Work_Sphere; W.origin = v3(origin.x, origin.y, origin.z); W.count = samples;
        W.passBase = scene->nextPass++; W.seed = scene->config.seed; W.salt = Salt_RayGen;
        W.sumColor = c; W.directions = d; W.accumulate = 0;
        float ms = 0.0f;
        ok = ok && runWavefront(scene->view, scene->pool, W, st, &ms);
        g_lastKernelMs = ms;
        ok = ok && PT_CUDA(cudaMemcpyAsync(cube, pCube, sizeof(F4) * 6, cudaMemcpyHostToDevice, st));
        ok = ok && launchAmbCubeFit(cube, d, c, samples, prevSampleCount, st);
        ok = ok && PT_CUDA(cudaMemcpyAsync(pCube, cube, sizeof(F4) * 6, cudaMemcpyDeviceToHost, st));
        ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
        if (!ok) return prevSampleCount;
    }
    return prevSampleCount + 1;
}

void PtMediaParams_Default(PtMediaParams* m)       // media_desc_new, path_tracer.c:1944-1961
{
    if (!m) return;
    memset(m, 0, sizeof(*m));
    m->constantColor = pim_float4{ 0.5f, 0.5f, 0.5f, 2.0f };
    m->noiseColor = pim_float4{ 0.5f, 0.5f, 0.5f, 2.0f };
    m->constantMfp = 40.0f * 1e3f; m->noiseMfp = 40.0f * 1e3f;
    m->absorption = 0.1f; m->noiseOctaves = 1; m->noiseGain = 0.9f; m->noiseLacunarity = 2.0666f;
    m->noiseFreq = 1.0f; m->noiseHeight = 20.0f; m->noiseScale = 1.0f;
    m->phaseDirA = 0.0f; m->phaseDirB = 0.0f; m->phaseBlend = 0.5f;
}

int32_t PtScene_SetMedia(PtScene* scene, const PtMediaParams* m)      // media_desc_update, path_tracer.c:1963-1987
{
    if (!scene || !m) { setError("PtScene_SetMedia: null argument"); return -1; }
    MediaDesc d;
    auto mu = [](float mfp, float color) { return 1.0f / (mfp * 0.5f + color * (mfp * 2.0f - mfp * 0.5f)); };   // f4_lerpsv, MeanFreePathToMu
    d.constantMu = v3(mu(m->constantMfp, m->constantColor.x), mu(m->constantMfp, m->constantColor.y), mu(m->constantMfp, m->constantColor.z));
    d.noiseMu = v3(mu(m->noiseMfp, m->noiseColor.x), mu(m->noiseMfp, m->noiseColor.y), mu(m->noiseMfp, m->noiseColor.z));
    d.absorption = m->absorption;
    d.noiseOctaves = m->noiseOctaves; d.noiseGain = m->noiseGain; d.noiseLacunarity = m->noiseLacunarity;
    d.noiseFreq = m->noiseFreq; d.noiseScale = m->noiseScale; d.noiseHeight = m->noiseHeight;
    d.phaseDirA = f1_clamp(m->phaseDirA, -0.99f, 0.99f);
    d.phaseDirB = f1_clamp(m->phaseDirB, -0.99f, 0.99f);
    d.phaseBlend = f1_sat(m->phaseBlend);
    float sum = 0.0f, amplitude = 1.0f;
    for (int i = 0; i < d.noiseOctaves; ++i) { sum += amplitude; amplitude *= d.noiseGain; }
    d.noiseRange = sum * d.noiseScale * 1.5f;
    const float a = 1.0f + d.absorption;
    const V3 maj = (d.constantMu * a + d.noiseMu * a) * 2.0f;           // CalcMajorant :2190-2196
    d.rcpMajorant = 1.0f / f1_max(maj.x, f1_max(maj.y, maj.z));
    cudaStreamSynchronize(scene->stream);
    scene->view.media = d;
    for (PtScene_s* r : scene->replicas) r->view.media = d;
    return 0;
}

int32_t LmPack_EmbedDesc(const PtEmbedMesh* meshes, int32_t meshCount, PtLightmapDesc* lightmaps, const int32_t* slotIndices,
                         int32_t lmCount)
{
    if (!ensureInit()) return -1;
    if (lmCount <= 0) return 0;
    if (!lightmaps || !slotIndices || (!meshes && meshCount > 0)) { setError("LmPack_EmbedDesc: null argument"); return -1; }
    const int32_t lmSize = lightmaps[0].size;
    for (int32_t k = 0; k < lmCount; ++k)
    {
        if (lightmaps[k].size != lmSize || lmSize <= 0) { setError("LmPack_EmbedDesc: pages differ in size"); return -1; }
        if (!lightmaps[k].position || !lightmaps[k].normal || !lightmaps[k].sampleCounts) { setError("LmPack_EmbedDesc: page %d has no storage", k); return -1; }
    }
    // the reference's scan order: drawables in order, whole triangles of each mesh in order
    const uint32_t kUnmapped = (1u << 1) | (1u << 4);       // kUnmappedMaterials, lightmap.c:38
    int64_t triCount = 0;
    for (int32_t i = 0; i < meshCount; ++i)
    {
        const PtEmbedMesh& m = meshes[i];
        if ((m.materialFlags & kUnmapped) || !m.positions || !m.normals || !m.uvs || !m.texIndices) continue;
        triCount += m.length / 3;
    }
    if (triCount > 0xffffffffll) { setError("LmPack_EmbedDesc: more than 2^32 triangles"); return -1; }
    const size_t texels = (size_t)lmCount * lmSize * lmSize;
    cudaStream_t st = nullptr;
    F4 *dPos = nullptr, *dNor = nullptr, *dUv = nullptr; int32_t* dPage = nullptr; unsigned long long* dKeys = nullptr;
    float *dOutPos = nullptr, *dOutNor = nullptr, *dOutCnt = nullptr;
    Lease lease;
    if (!lease.dev(dPos, 3 * (size_t)triCount).dev(dNor, 3 * (size_t)triCount).dev(dUv, 3 * (size_t)triCount).dev(dPage, (size_t)triCount)
              .dev(dKeys, texels).dev(dOutPos, 3 * texels).dev(dOutNor, 3 * texels).dev(dOutCnt, texels).commit()) return -1;
    std::vector<int32_t> page((size_t)triCount);
    bool ok = true;
    int64_t base = 0;
    for (int32_t i = 0; ok && i < meshCount; ++i)
    {
        const PtEmbedMesh& m = meshes[i];
        if ((m.materialFlags & kUnmapped) || !m.positions || !m.normals || !m.uvs || !m.texIndices) continue;
        const int64_t nt = m.length / 3;
        if (nt <= 0) continue;
        for (int64_t t = 0; t < nt; ++t)
        {
            const int32_t slot = m.texIndices[4 * (3 * t) + 3];     // texIndices[iVert].w, lightmap.c:991
            int32_t p = -1;
            for (int32_t k = 0; k < lmCount; ++k) { if (slotIndices[k] == slot) { p = k; break; } }
            page[(size_t)(base + t)] = p;
        }
        const size_t bytes = 16 * 3 * (size_t)nt, off = 3 * (size_t)base;
        ok = copyIn(dPos + off, m.positions, bytes, st) &&
             copyIn(dNor + off, m.normals, bytes, st) &&
             copyIn(dUv + off, m.uvs, bytes, st);
        base += nt;
    }
    ok = ok && PT_CUDA(cudaMemcpyAsync(dPage, page.data(), 4 * (size_t)triCount, cudaMemcpyHostToDevice, st));
    ok = ok && PT_CUDA(cudaMemsetAsync(dKeys, 0xff, 8 * texels, st));
    EmbedMesh M;
    M.positions = dPos; M.normals = dNor; M.uvs = dUv;
    M.page = dPage; M.triCount = triCount;
    KernelTimer timer(lease, st);
    ok = ok && launchEmbed(M, lmSize, lmCount, dKeys, dOutPos, dOutNor, dOutCnt, st);
    timer.stop();
    const size_t n = (size_t)lmSize * lmSize;
    for (int32_t k = 0; ok && k < lmCount; ++k)
    {
        ok = copyOut((void*)lightmaps[k].position, dOutPos + 3 * n * k, 12 * n, st) &&
             copyOut((void*)lightmaps[k].normal, dOutNor + 3 * n * k, 12 * n, st) &&
             copyOut(lightmaps[k].sampleCounts, dOutCnt + n * k, 4 * n, st);
    }
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

int32_t LmPack_ChartMasksDesc(pim_float2* triCoords, const int32_t* chartFirstNode, int32_t chartCount, float* areaOut,
                              int32_t* maskSizeOut, uint8_t* const* masksOut)
{
    if (!ensureInit()) return -1;
    if (chartCount <= 0) return 0;
    if (!triCoords || !chartFirstNode || !areaOut || !maskSizeOut) { setError("LmPack_ChartMasksDesc: null argument"); return -1; }
    const size_t corners = 3 * (size_t)chartFirstNode[chartCount];
    chartPrepare(reinterpret_cast<V2*>(triCoords), chartFirstNode, chartCount, areaOut, maskSizeOut);
    if (!masksOut) return 0;            // the caller only wanted the moved charts, their areas and mask sizes
    const size_t words = chartMaskWords(maskSizeOut, chartCount);
    size_t maxTexels = 1;
    for (int32_t c = 0; c < chartCount; ++c) maxTexels = std::max(maxTexels, (size_t)maskSizeOut[2 * c] * (size_t)std::max(maskSizeOut[2 * c + 1], 0));
    V2* dTri = nullptr; uint32_t* dBits = nullptr; char* dCharts = nullptr; uint8_t* dBytes = nullptr;
    Lease lease;
    if (!lease.dev(dTri, corners).dev(dBits, words).dev(dCharts, chartInfoBytes() * (size_t)chartCount).dev(dBytes, maxTexels).commit()) return -1;
    cudaStream_t st = nullptr;
    std::vector<int32_t> wordBase;
    bool ok = PT_CUDA(cudaMemcpyAsync(dTri, triCoords, sizeof(V2) * corners, cudaMemcpyHostToDevice, st));
    KernelTimer timer(lease, st);
    ok = ok && launchChartMasks(chartFirstNode, maskSizeOut, chartCount, dCharts, dTri, dBits, wordBase, st);
    timer.stop();
    for (int32_t c = 0; ok && c < chartCount; ++c)
    {
        const int32_t w = maskSizeOut[2 * c], h = maskSizeOut[2 * c + 1];
        if (!masksOut[c] || w <= 0 || h <= 0) continue;
        ok = launchMaskBytes(dCharts, c, w, h, dBits, dBytes, st) &&
             copyOut(masksOut[c], dBytes, (size_t)w * (size_t)h, st) &&
             PT_CUDA(cudaStreamSynchronize(st));
    }
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

int32_t LmPack_AtlasDesc(const pim_float2* triCoords, const int32_t* chartFirstNode, int32_t chartCount, const float* area,
                         const int32_t* maskSize, int32_t atlasSize, int32_t* atlasIndexOut, int32_t* translationOut,
                         int32_t* atlasCountOut)
{
    if (!ensureInit()) return -1;
    if (atlasCountOut) *atlasCountOut = 0;
    if (chartCount <= 0) return 0;
    if (!triCoords || !chartFirstNode || !area || !maskSize || !atlasIndexOut || !translationOut || atlasSize <= 0 || atlasSize > 32768)
    {
        setError("LmPack_AtlasDesc: bad argument");
        return -1;
    }
    // atlas_estimate (lightmap.c:825-842): enough atlases for the summed chart areas
    long long areaRequired = 0;
    for (int32_t c = 0; c < chartCount; ++c) areaRequired += (long long)ceilf(area[c]);
    const long long perAtlas = (long long)atlasSize * atlasSize;
    int32_t atlasCount = (int32_t)std::max<long long>(1, (areaRequired + perAtlas - 1) / perAtlas);
    const size_t corners = 3 * (size_t)chartFirstNode[chartCount];
    const size_t words = chartMaskWords(maskSize, chartCount);
    const size_t atlasWords = (size_t)((atlasSize + 31) / 32) * (size_t)atlasSize * (size_t)atlasCount;
    V2* dTri = nullptr; uint32_t *dBits = nullptr, *dAtlas = nullptr, *dBest = nullptr; char* dCharts = nullptr;
    Lease lease;
    if (!lease.dev(dTri, corners).dev(dBits, words).dev(dCharts, chartInfoBytes() * (size_t)chartCount).dev(dAtlas, atlasWords).dev(dBest, 64).commit()) return -1;
    cudaStream_t st = nullptr;
    std::vector<int32_t> wordBase;
    bool ok = PT_CUDA(cudaMemcpyAsync(dTri, triCoords, sizeof(V2) * corners, cudaMemcpyHostToDevice, st)) &&
              PT_CUDA(cudaMemsetAsync(dAtlas, 0, 4 * atlasWords, st));
    KernelTimer timer(lease, st);
    ok = ok && launchChartMasks(chartFirstNode, maskSize, chartCount, dCharts, dTri, dBits, wordBase, st);
    ok = ok && placeCharts(chartFirstNode, maskSize, area, wordBase, chartCount, atlasSize, atlasCount, dAtlas, dBits, dBest,
                           atlasIndexOut, translationOut, st);
    timer.stop();
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    if (ok && atlasCountOut)
    {
        int32_t used = 0;
        for (int32_t c = 0; c < chartCount; ++c) used = std::max(used, atlasIndexOut[c] + 1);
        *atlasCountOut = used;
    }
    return ok ? 0 : -1;
}

int32_t Pt_BakeSky(Cubemap* cm, const SkyMedium* sky, pim_float3 sunDir, float sunLum, int32_t steps)
{
    if (!ensureInit()) return -1;
    if (!sky) { setError("Pt_BakeSky: null medium"); return -1; }
    if (!cubeValid(cm, "Pt_BakeSky", false)) return -1;
    const size_t len = (size_t)cm->size * cm->size;
    float* dColor = nullptr;
    Lease lease;
    if (!lease.dev(dColor, 3 * len * 6).commit()) return -1;
    SkyParams A;
    A.rCrust = sky->rCrust; A.rAtmos = sky->rAtmos; A.muR = v3(sky->muR.x, sky->muR.y, sky->muR.z);
    A.rhoR = sky->rhoR; A.muM = sky->muM; A.rhoM = sky->rhoM; A.gM = sky->gM;
    cudaStream_t st = nullptr;
    KernelTimer timer(lease, st);
    bool ok = launchSkyBake(A, cm->size, v3(sunDir.x, sunDir.y, sunDir.z), sunLum, steps, dColor, st);
    timer.stop();
    ok = ok && cubeDownload(cm, dColor, nullptr, st);
    if (!ok) cudaStreamSynchronize(st);
    return ok ? 0 : -1;
}

int32_t Pt_Tonemap(const pim_float4* light, int64_t count, float exposure, uint32_t seed, uint32_t pass, uint32_t* rgba8Out)
{
    if (!ensureInit()) return -1;
    if (count <= 0) return 0;
    if (!light || !rgba8Out) { setError("Pt_Tonemap: null argument"); return -1; }
    F4* dLight = nullptr; uint32_t* dOut = nullptr;
    Lease lease;
    if (!lease.dev(dLight, (size_t)count).dev(dOut, (size_t)count).commit()) return -1;
    cudaStream_t st = nullptr;
    bool ok = copyIn(dLight, light, 16 * (size_t)count, st);
    KernelTimer timer(lease, st);
    ok = ok && launchTonemap(dLight, count, exposure, pass, seed, dOut, st);
    timer.stop();
    ok = ok && copyOut(rgba8Out, dOut, 4 * (size_t)count, st);
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

int32_t Pt_RepackRome(const pim_float4* mr, pim_int2 mrSize, const pim_float4* occ, pim_int2 occSize, const pim_float4* em,
                      pim_int2 emSize, int32_t size, uint32_t* romeOut)
{
    if (!ensureInit()) return -1;
    if (!mr || !occ || !em || !romeOut || size <= 0 || mrSize.x <= 0 || mrSize.y <= 0 || occSize.x <= 0 || occSize.y <= 0 ||
        emSize.x <= 0 || emSize.y <= 0)
    {
        setError("Pt_RepackRome: null or empty argument");
        return -1;
    }
    const size_t nMr = (size_t)mrSize.x * mrSize.y, nOcc = (size_t)occSize.x * occSize.y, nEm = (size_t)emSize.x * emSize.y;
    const size_t n = (size_t)size * size;
    F4 *dMr = nullptr, *dOcc = nullptr, *dEm = nullptr; uint32_t* dOut = nullptr;
    Lease lease;
    if (!lease.dev(dMr, nMr).dev(dOcc, nOcc).dev(dEm, nEm).dev(dOut, n).commit()) return -1;
    cudaStream_t st = nullptr;
    bool ok = copyIn(dMr, mr, 16 * nMr, st) &&
              copyIn(dOcc, occ, 16 * nOcc, st) &&
              copyIn(dEm, em, 16 * nEm, st);
    KernelTimer timer(lease, st);
    ok = ok && launchRepackRome(dMr, mrSize.x, mrSize.y, dOcc, occSize.x, occSize.y, dEm, emSize.x, emSize.y, size, dOut, st);
    timer.stop();
    ok = ok && copyOut(romeOut, dOut, 4 * n, st);
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    return ok ? 0 : -1;
}

int32_t PtScene_GetStats(const PtScene* scene, PtSceneStats* stats)
{
    if (!scene || !stats) return -1;
    memset(stats, 0, sizeof(*stats));
    const DScene& V = scene->view;
    stats->vertCount = V.vertCount; stats->matCount = V.matCount; stats->emissiveCount = V.emissiveCount;
    stats->gridSize[0] = V.grid.sx; stats->gridSize[1] = V.grid.sy; stats->gridSize[2] = V.grid.sz;
    stats->gridLo[0] = V.grid.lo.x; stats->gridLo[1] = V.grid.lo.y; stats->gridLo[2] = V.grid.lo.z;
    stats->gridHi[0] = V.grid.hi.x; stats->gridHi[1] = V.grid.hi.y; stats->gridHi[2] = V.grid.hi.z;
    stats->cellsPerMeter = V.grid.cellsPerMeter;
    stats->bvhNodeCount = V.bvh.nodeCount; stats->bvhTriCount = V.bvh.triCount;
    stats->buildMs = scene->buildMs;
    for (int k = 0; k < 6; ++k) stats->phaseMs[k] = scene->phaseMs[k];
    stats->bvhRefitted = scene->refitted ? 1 : 0;
    return 0;
}

int32_t PtScene_Download(const PtScene* scene, pim_float4* positions, pim_float4* normals, pim_float2* uvs, int32_t* matIds,
                         int32_t* vertToEmit, int32_t* emitToVert)
{
    if (!scene) return -1;
    const DScene& V = scene->view;
    const size_t n = (size_t)V.vertCount;
    bool ok = true;
    if (positions && n) ok = ok && PT_CUDA(cudaMemcpy(positions, V.positions, sizeof(F4) * n, cudaMemcpyDeviceToHost));
    if (normals && n) ok = ok && PT_CUDA(cudaMemcpy(normals, V.normals, sizeof(F4) * n, cudaMemcpyDeviceToHost));
    if (uvs && n) ok = ok && PT_CUDA(cudaMemcpy(uvs, V.uvs, sizeof(V2) * n, cudaMemcpyDeviceToHost));
    if (matIds && n) ok = ok && PT_CUDA(cudaMemcpy(matIds, V.matIds, 4 * n, cudaMemcpyDeviceToHost));
    if (vertToEmit && n) ok = ok && PT_CUDA(cudaMemcpy(vertToEmit, V.vertToEmit, 4 * n, cudaMemcpyDeviceToHost));
    if (emitToVert && V.emissiveCount) ok = ok && PT_CUDA(cudaMemcpy(emitToVert, V.emitToVert, 4 * (size_t)V.emissiveCount, cudaMemcpyDeviceToHost));
    return ok ? 0 : -1;
}

int32_t PtScene_GetDist(const PtScene* scene, int32_t cell, float* pdf, float* cdf, uint32_t* live)
{
    if (!scene || cell < 0 || cell >= scene->gridCells) return -1;
    const DScene& V = scene->view;
    int32_t len = 0;
    if (!PT_CUDA(cudaMemcpy(&len, V.distLength + cell, 4, cudaMemcpyDeviceToHost))) return -1;
    const size_t E = (size_t)V.emissiveCount;
    if (len > 0)
    {
        if (pdf && !PT_CUDA(cudaMemcpy(pdf, V.distPdf + cell * E, 4 * E, cudaMemcpyDeviceToHost))) return -1;
        if (cdf && !PT_CUDA(cudaMemcpy(cdf, V.distCdf + cell * (E + 1), 4 * (E + 1), cudaMemcpyDeviceToHost))) return -1;
        if (live && !PT_CUDA(cudaMemcpy(live, V.distLive + cell * E, 4 * E, cudaMemcpyDeviceToHost))) return -1;
    }
    return len;
}

int32_t PtScene_SetDist(PtScene* scene, int32_t cell, int32_t length, const float* pdf)
{
    if (!scene || cell < 0 || cell >= scene->gridCells) return -1;
    for (PtScene_s* r : scene->replicas)
    {
        DeviceGuard guard(r->device);
        if (PtScene_SetDist(r, cell, length, pdf) != 0) return -1;
    }
    const DScene& V = scene->view;
    const int32_t E = V.emissiveCount;
    if (length != 0 && length != E) { setError("PtScene_SetDist: length must be 0 or the emissive count"); return -1; }
    bool ok = PT_CUDA(cudaMemcpy(V.distLength + cell, &length, 4, cudaMemcpyHostToDevice));
    if (ok && length > 0)
    {
        // Dist1D_Bake on the host (dist1d.c:33-73), then upload
        std::vector<float> p(pdf, pdf + E), c((size_t)E + 1);
        const float rcpLen = 1.0f / E;
        c[0] = 0.0f;
        for (int i = 1; i <= E; ++i) c[i] = c[i - 1] + p[i - 1] * rcpLen;
        float integral = c[E];
        if (integral == 0.0f) { for (int i = 1; i <= E; ++i) c[i] = i * rcpLen; }
        else
        {
            float r = 1.0f / integral;
            for (int i = 1; i <= E; ++i) c[i] *= r;
            for (int i = 0; i < E; ++i) p[i] *= r;
        }
        std::vector<uint32_t> zero((size_t)E, 0u);
        uint32_t zsum = 0;
        ok = PT_CUDA(cudaMemcpy(V.distPdf + (size_t)cell * E, p.data(), 4 * (size_t)E, cudaMemcpyHostToDevice)) &&
             PT_CUDA(cudaMemcpy(V.distCdf + (size_t)cell * (E + 1), c.data(), 4 * ((size_t)E + 1), cudaMemcpyHostToDevice)) &&
             PT_CUDA(cudaMemcpy(V.distLive + (size_t)cell * E, zero.data(), 4 * (size_t)E, cudaMemcpyHostToDevice)) &&
             PT_CUDA(cudaMemcpy(V.distSum + cell, &zsum, 4, cudaMemcpyHostToDevice));
    }
    return ok ? 0 : -1;
}

int64_t PtScene_LightCounters(PtScene* scene, void* dCounters, int32_t toScene, void* stream)
{
    if (!scene) { setError("PtScene_LightCounters: null scene"); return -1; }
    const size_t words = (size_t)scene->gridCells * (size_t)scene->view.emissiveCount;
    if (!dCounters || words == 0) return (int64_t)words;
    cudaStream_t st = stream ? (cudaStream_t)stream : scene->stream;
    const bool ok = toScene ? PT_CUDA(cudaMemcpyAsync(scene->view.distLive, dCounters, 4 * words, cudaMemcpyDeviceToDevice, st))
                            : PT_CUDA(cudaMemcpyAsync(dCounters, scene->view.distLive, 4 * words, cudaMemcpyDeviceToDevice, st));
    return ok ? (int64_t)words : -1;
}

int32_t PtScene_GetBrdfLut(const PtScene* scene, float* texels, int32_t* width, int32_t* height)
{
    if (!scene) return -1;
    if (width) *width = scene->view.lutW;
    if (height) *height = scene->view.lutH;
    if (texels && !PT_CUDA(cudaMemcpy(texels, scene->dLut, sizeof(float) * 2 * (size_t)scene->view.lutW * scene->view.lutH, cudaMemcpyDeviceToHost))) return -1;
    return 0;
}

void Pt_GetCounters(PtCounters* out)
{
    if (!out) return;
    out->paths = g_counters.paths; out->rays = g_counters.rays; out->occlusionRays = g_counters.occlusionRays;
    out->kernelLaunches = g_counters.kernelLaunches; out->iterations = g_counters.iterations;
}
void Pt_ResetCounters(void)
{
    g_counters.paths = 0; g_counters.rays = 0; g_counters.occlusionRays = 0; g_counters.kernelLaunches = 0; g_counters.iterations = 0;
}

void Pt_SetKernelTiming(int32_t enable)
{
    g_kernelTiming = enable != 0;
    memset(&g_kernelStats, 0, sizeof(g_kernelStats));
}
void Pt_GetKernelStats(PtKernelStats* out)
{
    if (!out) return;
    for (int k = 0; k < 5; ++k) { out->ms[k] = g_kernelStats.ms[k]; out->launches[k] = g_kernelStats.launches[k]; }
    out->extendRays = g_kernelStats.extendRays;
    out->connectRays = g_kernelStats.connectRays;
}
int32_t Pt_IntersectCountN(const PtScene* scene, const pim_float4* ro, const pim_float4* rd, float tNear, int64_t n,
                           uint64_t* nodeFetches, uint64_t* triFetches)
{
    if (!scene || n <= 0) { setError("Pt_IntersectCountN: bad argument"); return -1; }
    F4 *dRo = nullptr, *dRd = nullptr; unsigned long long* dTotals = nullptr;
    Lease lease;
    if (!lease.dev(dRo, (size_t)n).dev(dRd, (size_t)n).dev(dTotals, 2).commit()) return -1;
    cudaStream_t st = scene->stream;
    unsigned long long a = 0, b = 0;
    bool ok = PT_CUDA(cudaMemcpyAsync(dRo, ro, sizeof(F4) * n, cudaMemcpyHostToDevice, st)) &&
              PT_CUDA(cudaMemcpyAsync(dRd, rd, sizeof(F4) * n, cudaMemcpyHostToDevice, st)) &&
              launchIntersectCount(scene->view, dRo, dRd, tNear, n, dTotals, &a, &b, st);
    if (!ok) cudaStreamSynchronize(st);
    *nodeFetches = a; *triFetches = b;
    return ok ? 0 : -1;
}
float Pt_LastKernelMs(void) { return g_lastKernelMs; }

float Pt_MeasureL2ReadGBs(int64_t bytes, int32_t reps)
{
    if (!ensureInit() || bytes < 4096 || reps < 1) return -1.0f;
    char* buf = nullptr; unsigned* sink = nullptr;
    Lease lease;
    if (!lease.dev(buf, (size_t)bytes).dev(sink, 64).commit()) return -1.0f;
    cudaStream_t st = nullptr;
    bool ok = PT_CUDA(cudaMemsetAsync(buf, 1, (size_t)bytes, st)) && launchL2Read(buf, (size_t)bytes, 2, sink, st);    // warm: the buffer is in L2
    KernelTimer timer(lease, st);
    ok = ok && launchL2Read(buf, (size_t)bytes, reps, sink, st);
    timer.stop();
    ok = PT_CUDA(cudaStreamSynchronize(st)) && ok;
    if (!ok || g_lastKernelMs <= 0.0f) return -1.0f;
    return (float)((double)bytes * reps / (g_lastKernelMs * 1e-3) / 1e9);
}

int32_t Pt_SetOption(const char* name, int32_t value)
{
    if (!name) return -1;
    if (!ensureInit()) return -1;       // the environment's values are read first, then overridden here
    if (!strcmp(name, "persistent")) { g_persistentTraversal = value; return 0; }
    if (!strcmp(name, "finish")) { g_finishThreshold = value; return 0; }
    if (!strcmp(name, "refit")) { g_refitOnUpdate = value != 0; return 0; }
    setError("Pt_SetOption: unknown option '%s'", name);
    return -1;
}

} // extern "C"
