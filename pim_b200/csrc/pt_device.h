// pt_device.h — host-side objects of libpimpt.so and the launch functions each
// .cu file exports to the C-ABI layer (api.cu).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <string>
#include <vector>

#include "integrator.h"

namespace pt {

void setError(const char* fmt, ...);
bool cudaOk(cudaError_t e, const char* what);
#define PT_CUDA(call) ::pt::cudaOk((call), #call)

struct LibCounters      // bumped from every host thread that drives a device (in-process multi-GPU)
{
    std::atomic<unsigned long long> paths{ 0 }, rays{ 0 }, occlusionRays{ 0 }, kernelLaunches{ 0 }, iterations{ 0 };
};
extern thread_local bool t_workerThread;     // a replica's host thread: no per-kernel timing statistics from it
extern LibCounters g_counters;

// per-kernel CUDA-event timing of the wavefront (bench.py's roofline numbers)
struct KernelStats
{
    double ms[5];                    // regen, extend, shadeA, connect, shadeB
    unsigned long long launches[5];
    unsigned long long extendRays, connectRays;
};
extern bool g_kernelTiming;
extern int g_persistentTraversal;
extern int g_finishThreshold;   // 0: one thread per ray, 1: persistent warps that refill idle lanes, -1: by BVH size
bool usePersistent(int32_t bvhNodeCount);
bool usePersistentBatch(int32_t bvhNodeCount);
extern KernelStats g_kernelStats;

constexpr int kSmCount = 148;           // B200: 2 dies x 74 SMs

// ------------------------------------------------------------------ BVH build
struct BvhBuffers
{
    WideNode* nodes = nullptr;
    WideNodeF* nodesF = nullptr;    // fp32 form of the same nodes, small trees only
    PackedTri* tris = nullptr;
    int32_t nodeCount = 0;
    int32_t triCount = 0;
    int32_t depth = 0;              // levels of the wide tree (the traversal stack holds at most 7 entries per level)
    std::vector<int32_t> levelEnd;  // the collapse is breadth first: level l is nodes [levelEnd[l-1], levelEnd[l])
    float sceneLo[3] = { 0, 0, 0 }, sceneHi[3] = { 0, 0, 0 };   // bounds of all vertices
};
// positions: device float4[3*triCount]. Allocates nodes/tris with cudaMalloc.
bool buildBvh(const F4* dPositions, int32_t triCount, BvhBuffers& out, cudaStream_t stream);
// Same topology, moved vertices ("refit path for unchanged topology"): the packed triangles are
// rewritten from the new positions in their leaf order and every wide node's frame and child boxes are recomputed
// bottom up, level by level. Hits are those of a fresh build (the closest hit does not depend on the tree); how
// fast they are found depends on how far the geometry moved.
bool refitBvh(const F4* dPositions, int32_t triCount, BvhBuffers& bvh, cudaStream_t stream);
void freeBvh(BvhBuffers& b);

// LSD radix sort of (64-bit key, 32-bit value) pairs (bvh_build.cu); hist holds 256 * radixSortBlocks(n) words
int radixSortBlocks(int n);
bool radixSortPairs(uint64_t* keysA, uint64_t* keysB, uint32_t* valsA, uint32_t* valsB, int n, int passes,
                    uint32_t* hist, cudaStream_t stream);

// ---------------------------------------------------------------- ray batches
struct RayHitOut      // PtRayHit, 48 bytes
{
    float w, u, v, t;
    float nx, ny, nz, nw;
    int32_t type, iVert;
    uint32_t flags, pad;
};
// dCursor: 8 bytes of device memory for the persistent kernels' ray cursor, private to this call
// until its kernel has finished (the scene hands them out from a ring, api.cu)
bool launchIntersect(const DScene& S, const F4* dRo, const F4* dRd, float tNear, RayHitOut* dOut, int64_t n,
                     unsigned long long* dCursor, cudaStream_t stream);
bool launchOccluded(const DScene& S, const F4* dRo, const F4* dRd, float tNear, uint8_t* dOut, int64_t n,
                    unsigned long long* dCursor, cudaStream_t stream);
bool launchPointQuery(const DScene& S, const F4* dPts, float* dDist, int64_t n, cudaStream_t stream);
// counting variant for the bytes-per-ray roofline: returns total node / triangle fetches
bool launchIntersectCount(const DScene& S, const F4* dRo, const F4* dRd, float tNear, int64_t n,
                          unsigned long long* dTotals, unsigned long long* nodesOut, unsigned long long* trisOut,
                          cudaStream_t stream);

// ---------------------------------------------------------------- scene build
struct EntityUpload
{
    float M[16];        // local to world (column major)
    float IM[9];        // inverse transpose of the upper 3x3 (column major)
    int32_t vertOffset;
    int32_t length;
    int32_t srcOffset;  // into the concatenated source mesh arrays
    int32_t pad;
};
bool launchFlatten(const EntityUpload* dEntities, int32_t entityCount, const F4* dSrcPos, const F4* dSrcNrm,
                   const F4* dSrcUv, F4* dPos, F4* dNrm, V2* dUv, int32_t* dMatIds, int32_t vertCount,
                   cudaStream_t stream);
bool launchEmissionPdf(const DScene& S, int32_t triCount, uint32_t seed, float* dPdfOut, cudaStream_t stream);
bool launchLightGrid(const DScene& S, float ptDistMeters, uint32_t seed, cudaStream_t stream);
bool launchDistUpdate(const DScene& S, cudaStream_t stream);
bool launchBrdfLutBake(float* dLut, int32_t w, int32_t h, int32_t samples, uint32_t seed, cudaStream_t stream);

// ------------------------------------------------------------------ bakes.cu
struct SkyParams          // SkyMedium, src/math/types.h:531-542
{
    float rCrust, rAtmos;
    V3 muR; float rhoR;
    float muM, rhoM, gM;
};
int cubemapMipCount(int size);
int cubemapElemCount(int size);
bool launchCubeRays(int size, V3 origin, uint32_t pass, uint32_t seed, F4* dRo, F4* dRd, cudaStream_t stream);
bool launchCubeConvolve(const float* dColor, float* dConvolved, int size, uint32_t sampleCount, float weight, uint32_t pass,
                        uint32_t seed, cudaStream_t stream);
bool launchAmbCubeFit(F4* dCube, const F4* dDirections, const float* dColors3, int samples, int prevSampleCount,
                      cudaStream_t stream);
bool launchSkyBake(const SkyParams& sky, int size, V3 sunDir, float sunLum, int steps, float* dColor, cudaStream_t stream);
// the lightmapped triangles in the reference's scan order (drawable, then vertex), vertices 3t..3t+2:
// object-space positions and normals and Mesh.uvs (zw = lightmap uv) as EmbedTaskFn reads them;
// page[t] = index of the lightmap whose slot matches texIndices.w, or -1
struct EmbedMesh
{
    const F4* positions; const F4* normals; const F4* uvs;
    const int32_t* page;
    int64_t triCount;
};
bool launchEmbed(const EmbedMesh& mesh, int lmSize, int lmCount, unsigned long long* dKeys,
                 float* dPosition3, float* dNormal3, float* dSampleCounts, cudaStream_t stream);
bool launchRepackRome(const F4* dMr, int mrW, int mrH, const F4* dOcc, int occW, int occH, const F4* dEm, int emW, int emH,
                      int size, uint32_t* dDst, cudaStream_t stream);
bool launchTonemap(const F4* dLight, int64_t n, float exposure, uint32_t pass, uint32_t seed, uint32_t* dRgba8,
                   cudaStream_t stream);

// ------------------------------------------------------------------ wavefront
struct PathPool
{
    int32_t capacity = 0;
    // persistent path state, double buffered (survivors are moved when they are compacted)
    F4* ro[2] = { nullptr, nullptr };       // xyz origin, w = weight
    F4* rd[2] = { nullptr, nullptr };       // xyz direction, w = bits: bounce | prevFlags<<16
    F4* atten[2] = { nullptr, nullptr };    // xyz attenuation, w = bits: rng.x
    F4* lum[2] = { nullptr, nullptr };      // xyz luminance, w = bits: rng.y
    F4* albedo[2] = { nullptr, nullptr };   // xyz albedo accumulator, w = bits: result index
    F4* normal[2] = { nullptr, nullptr };   // xyz normal accumulator
    // per-iteration temporaries
    U4* hit1 = nullptr;     // t, u, v (float bits), primitive index
    F4* ro2 = nullptr;      // secondary ray origin, w = tFar
    F4* rd2 = nullptr;      // secondary ray direction, w = bits: kind | surface flags
    F4* pend = nullptr;     // Pending.a0..a3
    U4* hit2 = nullptr;
    F4* surfN = nullptr;    // stage A's surface for stage B: N + roughness
    F4* surfA = nullptr;    // albedo + metallic
    F4* surfM = nullptr;    // interpolated normal, written only where a normal map changed N
    int32_t* connectList = nullptr;   // slots that need a connect ray this iteration (persistent traversal)
    int32_t* counters = nullptr; // device: see WfCounter
    int32_t* hostCounters = nullptr; // pinned mirror of the per-iteration status
    void* all = nullptr;
    // events of the host loop, created once with the pool (runWavefront creates none per call)
    static constexpr int kRing = 8;
    cudaEvent_t evBegin = nullptr, evEnd = nullptr;
    cudaEvent_t evDone[kRing] = {};
    cudaEvent_t evKernel[kRing][6] = {};    // per-kernel timing, created the first time timing is on
    // growable device scratch owned by the pool for callers that feed it (lightmap bake: picked
    // texels, their rays and colours), so that a steady-state bake pass allocates nothing
    void* aux = nullptr;
    size_t auxBytes = 0;
};
// makes pool.aux at least `bytes` large; when it has to grow, the first keepBytes are carried over
bool poolAux(PathPool& pool, size_t bytes, size_t keepBytes, cudaStream_t stream);
// The counters many CTAs bump with atomics each sit in a 128-byte line of their own, away from
// the values every thread only reads (Wf_NumA, the work cursor): same-address atomics serialise
// in one L2 slice at about one per two clocks, and neighbours in the line queue behind them.
enum WfCounter : int
{
    Wf_NumA = 0,                 // live paths this iteration (read by every kernel)
    Wf_CursorLo = 8,             // 64-bit work cursor and total (ints 8..11)
    Wf_Status = 16,              // 4 ints published per iteration
    Wf_NumNext = 32,             // survivors written to the other buffer (one atomic per CTA of stage B)
    Wf_NumConnect = 64,          // slots asking for a second ray (one atomic per CTA of stage A)
    Wf_ExtendCursor = 96, Wf_ConnectCursor = 128,    // ray fetch cursors of the persistent traversal kernels
    Wf_FinishExtend = 160, Wf_FinishConnect = 161,   // rays traced by the finishing kernel
    Wf_COUNT = 192
};
constexpr size_t kWfCounterBytes = 1024;

bool poolAlloc(PathPool& pool, int32_t capacity);
void poolFree(PathPool& pool);

enum WorkKind : int { Work_Camera = 0, Work_Rays = 1, Work_Sphere = 2 };

struct WorkDesc
{
    int32_t kind;
    CamParams cam;          // Work_Camera: work item = sample * pixelCount + (pixel - pixelBegin)
    int32_t pixelBegin;     // Work_Camera: the pixel range this call renders (a device's band of the frame);
    int32_t pixelCount;     //              0 = the whole frame
    const F4* ro;           // Work_Rays: explicit rays, one path per item
    const F4* rd;
    const uint32_t* rngIds; // Work_Rays: optional RNG work-item id per ray (default: the ray's index)
    V3 origin;              // Work_Sphere: Pt_RayGen, uniform sphere directions from origin
    int64_t count;          // total work items
    uint32_t passBase;      // RNG pass of sample 0
    uint32_t seed;
    uint32_t salt;
    // outputs
    float* sumColor;        // Work_Camera: += per pixel (3 floats); else per item (3 floats)
    float* sumAlbedo;
    float* sumNormal;
    F4* directions;         // Work_Sphere: generated directions per item
    int32_t accumulate;     // 1: atomicAdd into sums (camera), 0: store per item
};
// Runs all work items to completion through the wavefront; blocking on `stream`.
bool runWavefront(const DScene& S, PathPool& pool, const WorkDesc& work, cudaStream_t stream, float* kernelMsOut);

// Same wavefront over explicit rays with its scratch in a static device array (no allocation);
// device pointers in and out, at most noAllocMaxPaths() paths per call.
bool runWavefrontNoAlloc(const DScene& S, const F4* dRo, const F4* dRd, int64_t n, uint32_t pass, uint32_t seed,
                         float* dColor3, float* dAlbedo3, float* dNormal3, cudaStream_t stream);
int32_t noAllocMaxPaths();

// one thread per path, whole path in the thread (small batches, bake)
bool launchTraceRays(const DScene& S, const F4* dRo, const F4* dRd, int64_t n, uint32_t pass, uint32_t seed,
                     float* dOut9, cudaStream_t stream);

struct LightmapView
{
    F4* probes[5];
    const float* position;   // float3
    const float* normal;     // float3
    float* sampleCounts;
    int32_t texelCount;
    int32_t texelBase;       // global work index of texel 0 (for seeding)
};
bool launchLightmapBake(const DScene& S, const LightmapView& lm, float metersPerTexel, float timeSlice, int32_t spp,
                        uint32_t pass, uint32_t seed, cudaStream_t stream);

// the same bake through the path wavefront: all samples of all picked texels trace in parallel
bool launchLightmapBakeWavefront(const DScene& S, PathPool& pool, const LightmapView& lm, float metersPerTexel,
                                 float timeSlice, int32_t spp, uint32_t pass, uint32_t seed, cudaStream_t stream);

// covered texels of a host page to / from a full-size device page (LmPack_BakeDesc)
// one covered texel of a host lightmap page on its way to the device and back (LmPack_BakeDesc)
struct LmTexelRecord { F4 probes[5]; float pos[3]; float nrm[3]; float cnt; int32_t idx; };
static_assert(sizeof(LmTexelRecord) == 112, "LmTexelRecord is seven 16-byte pieces");
bool launchLightmapScatter(const LightmapView& lm, const LmTexelRecord* dRecords, int32_t count, cudaStream_t stream);
bool launchLightmapGather(const LightmapView& lm, LmTexelRecord* dRecords, int32_t count, cudaStream_t stream);

// ------------------------------------------------------------------ lmpack.cu (chart masks, atlas placement)
void chartPrepare(V2* tri, const int32_t* firstNode, int32_t chartCount, float* areaOut, int32_t* maskSizeOut);
size_t chartMaskWords(const int32_t* maskSize, int32_t chartCount);
size_t chartInfoBytes();
bool launchChartMasks(const int32_t* firstNode, const int32_t* maskSize, int32_t chartCount, void* dCharts, const V2* dTri,
                      uint32_t* dBits, std::vector<int32_t>& wordBaseOut, cudaStream_t stream);
bool launchMaskBytes(const void* dCharts, int32_t chart, int32_t w, int32_t h, const uint32_t* dBits, uint8_t* dBytes, cudaStream_t stream);
bool placeCharts(const int32_t* firstNode, const int32_t* maskSize, const float* area, const std::vector<int32_t>& wordBase,
                 int32_t chartCount, int32_t atlasSize, int32_t atlasCount, uint32_t* dAtlas, const uint32_t* dBits, uint32_t* dBest,
                 int32_t* atlasIndexOut, int32_t* translationOut, cudaStream_t stream);

// in-process multi-GPU: dst += src; dst = a - dst; dst += src
bool launchAccumulate(float* dDst, const float* dSrc, int64_t n, cudaStream_t stream);
bool launchU32Sub(uint32_t* dDst, const uint32_t* dA, int64_t n, cudaStream_t stream);
bool launchU32Add(uint32_t* dDst, const uint32_t* dSrc, int64_t n, cudaStream_t stream);

bool launchL2Read(const void* dBuf, size_t bytes, int reps, unsigned* dSink, cudaStream_t stream);

// image helpers
bool launchResolve(float* dImage, const float* dSum, int64_t n, float prevCount, float addCount, cudaStream_t stream);
bool launchLerp(float* dImage, const float* dSample, int64_t n, float weight, cudaStream_t stream);

} // namespace pt
