/* pim_pt.h — C ABI of libpimpt.so, the B200 (sm_100a) replacement for pim's CPU
 * path-tracing hot path.
 *
 * The reference has no FFI layer: its boundary for this path is the plain C
 * header src/rendering/path_tracer.h:67-103 plus LmPack_Bake
 * (src/rendering/lightmap.h:69), linked statically into the engine. This
 * header exports the same names with the same struct layouts, so an engine
 * build can link libpimpt.so in place of path_tracer.c,
 * and adds the batch / device-pointer forms a GPU backend needs.
 *
 * Every entry point is blocking (the stream is synchronised before return)
 * unless it takes a cudaStream_t-like `stream` argument. Nothing aborts: on
 * failure functions return NULL / a negative status and Pt_LastError() holds
 * a message. There is no CPU fallback; without a CUDA device PtSys_Init fails.
 */
#ifndef PIM_PT_H
#define PIM_PT_H

#include <stdint.h>
#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- value types (layouts of src/math/types.h:52-73) ---------------------- */
#if defined(__CUDACC__) || defined(PIM_PT_NO_VECTOR_TYPES)
/* CUDA already defines float2/3/4 with the same layouts */
#else
typedef struct pim_float2_s { float x, y; } pim_float2;
typedef struct pim_float3_s { float x, y, z; } pim_float3;
typedef struct pim_float4_s { float x, y, z, w; } __attribute__((aligned(16))) pim_float4;
typedef struct pim_int2_s { int32_t x, y; } pim_int2;
#define PIM_PT_OWN_VECTOR_TYPES 1
#endif
#ifndef PIM_PT_OWN_VECTOR_TYPES
typedef float2 pim_float2;
typedef float3 pim_float3;
typedef float4 pim_float4;
typedef int2 pim_int2;
#endif

/* src/rendering/material.h:8-20 */
enum
{
    PtMatFlag_Emissive = 1 << 0,
    PtMatFlag_Sky = 1 << 1,
    PtMatFlag_Refractive = 1 << 5,
};

/* src/rendering/path_tracer.h:15-22 */
typedef enum { PtHit_Nothing = 0, PtHit_Backface, PtHit_Triangle, PtHit_COUNT } PtHitType;

/* src/rendering/path_tracer.h:24-31 (48 bytes) */
typedef struct PtRayHit_s
{
    pim_float4 wuvt;     /* barycentrics (w,u,v) and hit distance t */
    pim_float4 normal;   /* unit geometric normal, flipped to face the ray */
    int32_t type;        /* PtHitType */
    int32_t iVert;       /* 3 * triangle index, -1 on miss */
    uint32_t flags;      /* material flags of the hit triangle */
} PtRayHit;

/* src/rendering/path_tracer.h:33-42 (28 bytes) */
typedef struct PtDofInfo_s
{
    float aperture;
    float focalLength;
    int32_t bladeCount;
    float bladeRot;
    float focalPlaneCurvature;
    float autoFocusSpeed;
    bool autoFocus;
} PtDofInfo;

/* src/rendering/path_tracer.h:44-52. The four images are host memory (pinned by
 * PtTrace_New) that the engine's denoiser and blit read after Pt_Trace. */
typedef struct PtTrace_s
{
    pim_float3* color;
    pim_float3* albedo;
    pim_float3* normal;
    pim_float3* denoised;
    pim_int2 imageSize;
    float sampleWeight;
} PtTrace;

/* src/rendering/path_tracer.h:54-59 (36 bytes) */
typedef struct PtResult_s { pim_float3 color, albedo, normal; } PtResult;

/* src/rendering/path_tracer.h:61-65 */
typedef struct PtResults_s { pim_float4* colors; pim_float4* directions; } PtResults;

/* src/rendering/camera.h:8-15 */
typedef struct Camera_s
{
    pim_float4 position;
    pim_float4 rotation;   /* quaternion xyzw */
    float zNear, zFar, fovy;
} Camera;

typedef struct PtScene_s PtScene;

/* ---- scene input ----------------------------------------------------------
 * The reference's PtScene_New(void) reads engine globals (Entities_Get,
 * Mesh_Get, Texture_Get, Cubemaps_Get "sky", g_BrdfLut, cv_pt_dist_meters;
 * src/rendering/path_tracer.c:692-782, 1011-1021, 905). Here the same data is
 * handed over as one flat description; the in-engine adapter that fills it is
 * shown in INTEGRATION.md. */
typedef enum { PtTex_RGBA8 = 0, PtTex_RG16 = 1 } PtTexFormat;

typedef struct PtTextureDesc_s
{
    int32_t width, height;
    int32_t format;         /* PtTexFormat: albedo/rome = RGBA8, normal = RG16 (short2) */
    const void* texels;     /* host memory, row-major */
} PtTextureDesc;

typedef struct PtMaterialDesc_s   /* src/rendering/material.h:22-31 */
{
    int32_t albedo, rome, normal; /* indices into PtSceneDesc.textures, -1 = none */
    uint32_t flags;
    float meanFreePath[4];
    float ior;
    float bumpiness;
} PtMaterialDesc;

typedef struct PtEntityDesc_s     /* one drawable: src/rendering/drawable.h:14-27 + mesh.h:25-33 */
{
    const pim_float4* positions;  /* [length] object space, unindexed triangles */
    const pim_float4* normals;    /* [length] */
    const pim_float4* uvs;        /* [length] xy used */
    int32_t length;
    float matrix[16];             /* local-to-world, column major (float4x4 c0..c3) */
    PtMaterialDesc material;
} PtEntityDesc;

typedef struct PtConfig_s
{
    float ptDistMeters;     /* cv_pt_dist_meters, default 1.5 (src/common/cvars.c:208) */
    uint32_t seed;          /* replaces the wall-clock seeding of src/common/random.c:45-65 */
    int32_t pathPoolSize;   /* resident path slots of the wavefront pool; 0 = default */
    int32_t bakeMode;       /* PtBakeMode for LmPack_BakeDesc */
} PtConfig;

/* How LmPack_BakeDesc orders a texel's samples. Both modes shoot BakeFn's rays (same RNG stream)
 * and fold them into the lobes in sample order. PtBake_Wavefront traces all samples of all texels
 * in parallel, the integrator's RNG seeded per ray. PtBake_Sequential traces a texel's samples one
 * after the other in one thread with the integrator's RNG running on from sample to sample, which
 * reproduces the seeded reference texel for texel (used by the parity tests) but exposes only
 * texel-count parallelism. */
typedef enum { PtBake_Wavefront = 0, PtBake_Sequential = 1 } PtBakeMode;

typedef struct PtSceneDesc_s
{
    const PtEntityDesc* entities;
    int32_t entityCount;
    const PtTextureDesc* textures;
    int32_t textureCount;
    const float* skyFaces;  /* 6 faces (XP,XM,YP,YM,ZP,ZM) of skySize^2 float3, or NULL */
    int32_t skySize;
    const float* brdfLut;   /* lutWidth*lutHeight float2 (g_BrdfLut), or NULL = bake on device */
    int32_t lutWidth, lutHeight;
    PtConfig config;
} PtSceneDesc;

/* ---- the reference's entry points (src/rendering/path_tracer.h:67-103) ---- */
void PtSys_Init(void);
void PtSys_Update(void);
void PtSys_Shutdown(void);

/* PtScene_New(void) (path_tracer.c:1083-1093) builds from the description registered with
 * PtSys_SetSceneDesc, the adapter's stand-in for the engine globals the reference reads.
 * The call copies everything the description points to (entity table, mesh arrays, texels,
 * sky faces, BRDF table), so the caller may fill it from per-frame memory; NULL clears it.
 * Registering again is the equivalent of a changed Entities_Get()->modtime: the next
 * PtScene_Update / Pt_Trace / Pt_RayGen / LmPack_Bake on a scene made by PtScene_New(void)
 * rebuilds that scene in place (path_tracer.c:1028-1032) before it goes on. Scenes made by
 * PtScene_NewFromDesc are never rebuilt; that call reads the description's arrays only
 * until it returns. */
void PtSys_SetSceneDesc(const PtSceneDesc* desc);

/* Multi-GPU inside the one process pim is (src/main.c:27-111 runs everything from a single loop; the
 * reference has no counterpart, its parallel-for is Task_Run over host threads, threading/task.c:178-190).
 * After PtSys_SetDevices(n) every scene built carries a full replica on each of devices 0..n-1 besides the
 * one it is built on (the current device), and these calls spread their work, one host thread per device:
 *   Pt_TraceSpp / Pt_TraceSpp_Device   by sample index (every device renders the whole frame for its
 *                                      samples; the partial sums come to the scene's device by peer copy
 *                                      and are added there), or by pixel band when there are fewer
 *                                      samples than devices
 *   Pt_Trace                           one sample per pixel: by pixel band
 *   LmPack_BakeDesc / LmPack_Bake      contiguous ranges of the covered texels (never samples: the
 *                                      spherical-gaussian fit is sequential per texel)
 * A path's random stream is a function of (pixel or texel, sample, seed), and the light-learning counters
 * every device adds during a call are summed and handed back to all of them afterwards, so results are those
 * of one device up to the order of float additions. Everything else (ray batches, Pt_RayGen, cubemap bakes)
 * runs on the scene's own device. Returns the count in effect, -1 on error; call it before PtScene_New. */
int32_t PtSys_SetDevices(int32_t count);
int32_t PtSys_GetDevices(void);

PtScene* PtScene_New(void);
PtScene* PtScene_NewFromDesc(const PtSceneDesc* desc);
/* rebuild if the registered description changed, then relearn the light-grid distributions
 * (UpdateDists, path_tracer.c:2472-2483) */
void PtScene_Update(PtScene* scene);
void PtScene_Del(PtScene* scene);
void PtScene_Gui(PtScene* scene);   /* no-op */

void PtTrace_New(PtTrace* trace, pim_int2 imageSize);
void PtTrace_Del(PtTrace* trace);

void DofInfo_New(PtDofInfo* dof);
void DofInfo_Gui(PtDofInfo* dof);   /* no-op */

PtRayHit Pt_Intersect(const PtScene* scene, pim_float4 ro, pim_float4 rd, float tNear, float tFar);
PtResult Pt_TraceRay(PtScene* scene, pim_float4 ro, pim_float4 rd);
void Pt_Trace(PtTrace* trace, PtDofInfo* dof, PtScene* scene, const Camera* camera);
PtResults Pt_RayGen(PtScene* scene, pim_float4 origin, int32_t count);

/* ---- batch forms (host buffers; copies are inside the call) ---------------- */
/* ro[i].xyz origin, rd[i].xyz direction, rd[i].w = tFar. Replaces n calls of
 * Pt_Intersect (path_tracer.c:1466-1474 -> rtcIntersect1 :484). */
int32_t Pt_IntersectN(const PtScene* scene, const pim_float4* ro, const pim_float4* rd,
                      float tNear, PtRayHit* hitsOut, int64_t n);
/* any-hit: occluded[i] = 1 when something lies in [tNear, rd[i].w]. Replaces
 * RtcOccluded16 (path_tracer.c:524-553). */
int32_t Pt_OccludedN(const PtScene* scene, const pim_float4* ro, const pim_float4* rd,
                     float tNear, uint8_t* occludedOut, int64_t n);
/* nearest-surface query: points[i] = (x, y, z, radius); distanceOut[i] = unsigned distance to the
 * nearest triangle closer than the radius, 8388608 (kRcpEpsilon) when there is none. Replaces
 * RtcPointQuery / rtcPointQuery (path_tracer.c:597-616, :614), which the light-grid set-up runs per cell. */
int32_t Pt_PointQueryN(const PtScene* scene, const pim_float4* points, float* distanceOut, int64_t n);
/* n calls of Pt_TraceRay; path i is seeded from (i, pass, seed). */
int32_t Pt_TraceRayN(PtScene* scene, const pim_float4* ro, const pim_float4* rd,
                     uint32_t pass, PtResult* resultsOut, int64_t n);
/* sppCount calls of Pt_Trace folded into one: samples sppBegin..sppBegin+sppCount-1
 * of every pixel; images hold the running mean over sppBegin+sppCount samples.
 * The light-grid relearn (PtScene_Update) runs once per call. trace->sampleWeight
 * <= 0 leaves raw per-call sums in the images (multi-GPU partial results). */
int32_t Pt_TraceSpp(PtTrace* trace, PtDofInfo* dof, PtScene* scene, const Camera* camera,
                    int32_t sppBegin, int32_t sppCount);

/* ---- device-resident forms (pointers are CUDA device memory) --------------- */
int32_t Pt_IntersectN_Device(const PtScene* scene, const void* dRo, const void* dRd,
                             float tNear, void* dHitsOut, int64_t n, void* stream);
int32_t Pt_OccludedN_Device(const PtScene* scene, const void* dRo, const void* dRd,
                            float tNear, void* dOccludedOut, int64_t n, void* stream);
/* Adds sppCount samples per pixel into the three float3 sum images (w*h*3 floats each). */
int32_t Pt_TraceSpp_Device(PtScene* scene, const Camera* camera, PtDofInfo* dof,
                           int32_t width, int32_t height, int32_t sppBegin, int32_t sppCount,
                           void* dSumColor, void* dSumAlbedo, void* dSumNormal, void* stream);

/* Pt_TraceRayN for callers that must not allocate. No counterpart in the reference.
 *
 * Some hosts cannot call cudaMalloc where they trace: a render or simulation loop that
 * bans allocation after start-up to keep frame latency flat, a thread holding a lock that
 * an allocating CUDA call could stall behind, an embedding that accounts for all device
 * memory at load time. This entry point performs no cudaMalloc / cudaFree / cudaMallocHost
 * and creates no events: the wavefront's scratch (path state, ray and hit temporaries,
 * counters) is a fixed array inside the library image, about 10 MB, that every GPU
 * loading libpimpt.so holds until the process exits.
 *
 *  - dRo, dRd: device float4[n] (xyz used); outputs are device float3[n] arrays written
 *    per path; dAlbedo3 and dNormal3 may be NULL. Results equal Pt_TraceRayN's for the
 *    same scene, rays and pass.
 *  - n may not exceed PT_NOALLOC_MAX_PATHS; larger batches fail with -1 and a message in
 *    Pt_LastError() — split them, or use Pt_TraceRayN.
 *  - The scene's device must be the current one; otherwise the call fails with a message.
 *  - The first call on each device looks up the array's device address (with that device
 *    current). Make that first call before any CUDA stream capture begins.
 *  - All calls on a device share the one array. It is not cleared between calls (the call
 *    resets the counters it depends on), so two calls that overlap on the GPU — different
 *    streams, different host threads — corrupt each other. Serialise them, first calls
 *    included.
 *  - Blocking: every call synchronises the stream several times and returns after it has
 *    drained. */
#define PT_NOALLOC_MAX_PATHS 32768
int32_t Pt_TraceRayN_NoAlloc(PtScene* scene, const void* dRo, const void* dRd, uint32_t pass,
                             void* dColor3, void* dAlbedo3, void* dNormal3, int64_t n, void* stream);

/* ---- lightmap bake (src/rendering/lightmap.h:69, lightmap.c:1117-1223) ------
 * The reference bakes the global LmPack; here the pack's texel arrays are passed
 * explicitly. Layout per lightmap page as Lightmap_New (lightmap.c:81-123):
 * probes[5][texels] float4, position[texels] float3, normal[texels] float3,
 * sampleCounts[texels] float. All host pointers; probes and sampleCounts are
 * read and written. */
typedef struct PtLightmapDesc_s
{
    pim_float4* probes[5];
    const pim_float3* position;
    const pim_float3* normal;
    float* sampleCounts;
    int32_t size;           /* page is size x size texels */
} PtLightmapDesc;

/* Only the texels a chart covers (sampleCounts != 0; BakeFn skips the others, lightmap.c:1145)
 * cross the bus, compacted, each way. */
int32_t LmPack_BakeDesc(PtScene* scene, PtLightmapDesc* lightmaps, int32_t lmCount,
                        float texelsPerMeter, float timeSlice, int32_t spp, uint32_t pass);

/* LmPack_Bake under the reference's own name and signature (src/rendering/lightmap.h:69, body
 * lightmap.c:1204-1223; caller render_system.c:198). The reference bakes the engine's global
 * pack (LmPack_Get(), lightmap.c:136); PtSys_SetLmPack is the adapter's stand-in for that
 * global: it remembers the pack's pages (the descriptions are copied, the arrays stay the
 * engine's and must live until the pack is registered again or cleared with NULL / 0) and
 * pack->texelsPerMeter. LmPack_Bake then runs PtScene_Update and one bake pass over those
 * pages, spp clamped to >= 1, results written into the engine's arrays before it returns. */
void PtSys_SetLmPack(const PtLightmapDesc* lightmaps, int32_t lmCount, float texelsPerMeter);
void LmPack_Bake(PtScene* scene, float timeSlice, int32_t spp);
/* the same with every pointer of every page in device memory: the pack stays resident between
 * passes (a page is 108 B per texel; over PCIe it costs more than the bake itself) */
int32_t LmPack_BakeDesc_Device(PtScene* scene, const PtLightmapDesc* lightmaps, int32_t lmCount,
                               float texelsPerMeter, float timeSlice, int32_t spp, uint32_t pass);

/* ---- participating medium (PtMediaDesc, src/rendering/path_tracer.c:91-111) -----
 * The reference edits scene->mediaDesc through its UI (media_desc_gui) or a JSON file
 * (media_desc_load) and then runs media_desc_update (:1963-1987); this is the same
 * set of user-facing fields with the same defaults (media_desc_new :1944-1961). The
 * derived values (mu, noise range, majorant) are recomputed by the call. 0 on success. */
typedef struct PtMediaParams_s
{
    pim_float4 constantColor;   /* default (0.5, 0.5, 0.5, 2) */
    pim_float4 noiseColor;      /* default (0.5, 0.5, 0.5, 2) */
    float constantMfp;          /* metres, default 40 km */
    float noiseMfp;             /* metres, default 40 km */
    float absorption;           /* 0.1 */
    int32_t noiseOctaves;       /* 1 */
    float noiseGain;            /* 0.9 */
    float noiseLacunarity;      /* 2.0666 */
    float noiseFreq;            /* 1 */
    float noiseScale;           /* 1 */
    float noiseHeight;          /* 20 */
    float phaseDirA, phaseDirB, phaseBlend;   /* 0, 0, 0.5 */
} PtMediaParams;
void PtMediaParams_Default(PtMediaParams* params);
int32_t PtScene_SetMedia(PtScene* scene, const PtMediaParams* params);

/* ---- lightmap attribute embedding --------------------------
 * EmbedAttributes / EmbedTaskFn (static in src/rendering/lightmap.c:947-1047), the last and
 * O(texels x triangles) step of LmPack_Pack: for every texel of every page find the
 * lightmapped triangle nearest in lightmap space (within 2 texels) and store its
 * interpolated position, normal and a coverage flag. Chart grouping and atlas placement
 * (lightmap.c:527-945) stay host code and run before this call.
 * meshes[i]: one drawable in Entities order — the Mesh's own arrays (mesh.h:25-33; uvs.zw
 * = lightmap uv, texIndices = int4 per vertex, .w = lightmap slot) and its material flags
 * (drawables with PtMatFlag_Sky or lava, 1 << 4, are skipped as in the reference).
 * slotIndices[k] = lightmaps[k].slot.index. Writes position, normal and sampleCounts of
 * every page (probes are left alone); all pages must have the same size. 0 on success. */
typedef struct PtEmbedMesh_s
{
    const pim_float4* positions;
    const pim_float4* normals;
    const pim_float4* uvs;
    const int32_t* texIndices;
    int32_t length;          /* vertices */
    uint32_t materialFlags;
} PtEmbedMesh;
int32_t LmPack_EmbedDesc(const PtEmbedMesh* meshes, int32_t meshCount, PtLightmapDesc* lightmaps,
                         const int32_t* slotIndices, int32_t lmCount);

/* The two data-parallel steps of LmPack_Pack before the embedding. Chart grouping (plane clustering, k-means split:
 * chart_group, lightmap.c:563-642), chart_sort (:644-658) and chartnodes_assign (:867-935) stay host code around them.
 * Charts are runs of nodes: chartFirstNode[c] .. chartFirstNode[c + 1] - 1; triCoords holds the three 2-D corners
 * (chartnode_t.triCoord, texels) of every node, chart after chart.
 *
 * LmPack_ChartMasksDesc = ChartMaskFn (static, lightmap.c:527-561): moves every chart so that its lower corner is at
 * (2, 2) (triCoords is rewritten), returns its bounds area and the size of its mask, and, where masksOut[c] is not
 * NULL, the mask itself (maskSize.x * maskSize.y bytes, 0xff where a triangle of the chart lies within one texel of the
 * texel centre: mask_tri :232-248).
 *
 * LmPack_AtlasDesc = atlases_create / AtlasFn / atlas_search (static, :680-721, :779-865) for charts in the order
 * given (the caller has sorted them): rasterises the masks again on the device, then gives every chart the first
 * position, rows first, of the first atlas where it overlaps nothing placed before, with the reference's row-skipping
 * search state carried from chart to chart as in its single task chunk. atlasIndexOut[c], translationOut[2c..2c+1];
 * *atlasCountOut = atlases in use. A chart that fits nowhere is an error (the reference searches forever). */
int32_t LmPack_ChartMasksDesc(pim_float2* triCoords, const int32_t* chartFirstNode, int32_t chartCount, float* areaOut,
                              int32_t* maskSizeOut, uint8_t* const* masksOut);
int32_t LmPack_AtlasDesc(const pim_float2* triCoords, const int32_t* chartFirstNode, int32_t chartCount, const float* area,
                         const int32_t* maskSize, int32_t atlasSize, int32_t* atlasIndexOut, int32_t* translationOut,
                         int32_t* atlasCountOut);

/* ---- callers either side of the path -------------------
 * The reference functions that feed the path tracer (sky) or consume Pt_TraceRay /
 * Pt_RayGen (cubemap and ambient-cube bakes) or its image (screenshot), same names and
 * struct layouts where the reference exports them. */

/* src/rendering/cubemap.h:28-34. Cubemap_New allocates the twelve arrays as page-locked
 * host memory: color[f] = float3[size*size], convolved[f] = float4[elemCount] (mip chain). */
typedef struct Cubemap_s
{
    int32_t size;
    int32_t mipCount;
    pim_float3* color[6];
    pim_float4* convolved[6];
} Cubemap;
void Cubemap_New(Cubemap* cm, int32_t size);                                          /* cubemap.c:95-114 */
void Cubemap_Del(Cubemap* cm);                                                        /* cubemap.c:116-127 */
/* one jittered path per face texel from `origin`, lerped into color[] by `weight`; the pass
 * index is the scene's running counter (the reference draws from its thread generators) */
void Cubemap_Bake(Cubemap* cm, PtScene* scene, pim_float4 origin, float weight);     /* cubemap.c:162-187 */
/* GGX prefilter of color[] into every mip of convolved[], blended by `weight` */
void Cubemap_Convolve(Cubemap* cm, uint32_t sampleCount, float weight);               /* cubemap.c:262-303 */
/* the same two with the random stream named by the caller (seed, pass); 0 on success */
int32_t Cubemap_BakePass(Cubemap* cm, PtScene* scene, pim_float4 origin, float weight, uint32_t pass);
int32_t Cubemap_ConvolvePass(Cubemap* cm, uint32_t sampleCount, float weight, uint32_t seed, uint32_t pass);

/* src/math/types.h:513-517, src/math/ambcube.c:5-33: fits `samples` Pt_RayGen probes from
 * `origin` into the cube; returns prevSampleCount + 1 */
typedef struct AmbCube_s { pim_float4 Values[6]; } AmbCube_t;
int32_t AmbCube_Bake(PtScene* scene, AmbCube_t* pCube, pim_float4 origin, int32_t samples, int32_t prevSampleCount);

/* src/math/types.h:531-542 */
typedef struct SkyMedium_s
{
    float rCrust;
    float rAtmos;
    pim_float3 muR;
    float rhoR;
    float muM;
    float rhoM;
    float gM;
} SkyMedium;
/* BakeSky's task (static in src/rendering/render_system.c:403-425, 473-491): Atmosphere()
 * raymarch per texel of cm->color; 0 on success */
int32_t Pt_BakeSky(Cubemap* cm, const SkyMedium* sky, pim_float3 sunDir, float sunLum, int32_t steps);

/* CmdScreenshot's pixel loop (static in src/rendering/render_system.c:707-727): exposure,
 * GT tonemap, sRGB inverse EOTF, 1/255 dither, RGBA8. light: host float4[count]; rgba8Out:
 * host uint32[count] (R in the low byte); 0 on success */
int32_t Pt_Tonemap(const pim_float4* light, int64_t count, float exposure, uint32_t seed, uint32_t pass,
                   uint32_t* rgba8Out);

/* CreateRomeTexture's texel loop (static in src/rendering/gltf_model.c:1118-1139): resamples the
 * glTF metallic-roughness, occlusion and emission images (host float4, each of its own size; pass a
 * 1x1 default where the material has none: (0, 0.5, 0, 0), (1, 0, 0, 0), (0, 0, 0, 0)) onto one
 * size x size RGBA8 "ROME" texture (roughness, occlusion, metallic, packed emission) in
 * romeOut[size*size]. `size` is the reference's NextPow2(max side), at most 2048. 0 on success. */
int32_t Pt_RepackRome(const pim_float4* metallicRoughness, pim_int2 mrSize, const pim_float4* occlusion, pim_int2 occSize,
                      const pim_float4* emission, pim_int2 emSize, int32_t size, uint32_t* romeOut);

/* ---- introspection ----------------------------------------------------------- */
typedef struct PtSceneStats_s
{
    int32_t vertCount, matCount, emissiveCount;
    int32_t gridSize[3];
    float gridLo[4], gridHi[4];
    float cellsPerMeter;
    int32_t bvhNodeCount, bvhTriCount;
    float buildMs;
    float phaseMs[6];       /* upload+flatten, BVH, emissive pdf, light grid, BRDF LUT, pool */
    int32_t bvhRefitted;    /* 1: the tree was taken over from the scene this one replaced and refitted ("refit" option) */
} PtSceneStats;

int32_t PtScene_GetStats(const PtScene* scene, PtSceneStats* stats);
/* copies of the flattened scene (path_tracer.c:130-146); any pointer may be NULL */
int32_t PtScene_Download(const PtScene* scene, pim_float4* positions, pim_float4* normals,
                         pim_float2* uvs, int32_t* matIds, int32_t* vertToEmit, int32_t* emitToVert);
/* light-grid cell: returns the cell's distribution length (0 = none) */
int32_t PtScene_GetDist(const PtScene* scene, int32_t cell, float* pdf, float* cdf, uint32_t* live);
int32_t PtScene_SetDist(PtScene* scene, int32_t cell, int32_t length, const float* pdf);
/* The light-learning counters of the whole grid (Dist1D.live of every cell, path_tracer.c:1709-1733: cells x emissives
 * uint32, cell-major) copied device to device on `stream`: toScene = 0 reads them into dCounters, 1 writes them back.
 * Returns the number of counters (also with dCounters = NULL), -1 on error. One process per GPU uses it to sum what
 * every rank's call added (an all-reduce of the differences) so that all replicas relearn from the same counts, as
 * SURVEY.md 8(e) describes; the in-process multi-GPU mode does the same inside the library. */
int64_t PtScene_LightCounters(PtScene* scene, void* dCounters, int32_t toScene, void* stream);
/* BRDF LUT actually in use (lutWidth*lutHeight float2) */
int32_t PtScene_GetBrdfLut(const PtScene* scene, float* texels, int32_t* width, int32_t* height);

typedef struct PtCounters_s
{
    uint64_t paths;          /* camera paths / probe paths completed */
    uint64_t rays;           /* closest-hit BVH queries */
    uint64_t occlusionRays;  /* any-hit BVH queries */
    uint64_t kernelLaunches;
    uint64_t iterations;     /* wavefront iterations */
} PtCounters;
void Pt_GetCounters(PtCounters* out);
void Pt_ResetCounters(void);
/* milliseconds the last batch call spent in kernels, measured with CUDA events on
 * the library's stream (0 when events were not recorded) */
float Pt_LastKernelMs(void);

/* Per-kernel timing of the wavefront, measured with CUDA events on the launching
 * stream while enabled (index: 0 regen, 1 extend, 2 shadeA, 3 connect, 4 shadeB). */
typedef struct PtKernelStats_s
{
    double ms[5];
    uint64_t launches[5];
    uint64_t extendRays;     /* rays traced by the extend kernel */
    uint64_t connectRays;    /* NEE / MIS rays traced by the connect kernel */
} PtKernelStats;
void Pt_SetKernelTiming(int32_t enable);   /* also clears the accumulated stats */
void Pt_GetKernelStats(PtKernelStats* out);
/* closest-hit batch that only counts wide-node and triangle fetches (bytes-per-ray bookkeeping) */
int32_t Pt_IntersectCountN(const PtScene* scene, const pim_float4* ro, const pim_float4* rd, float tNear,
                           int64_t n, uint64_t* nodeFetches, uint64_t* triFetches);

/* Read bandwidth of the L2 cache as this library's kernels see it: `reps` read-only sweeps (128-bit loads) over a
 * buffer of `bytes` that was just brought into L2; GB/s, negative on error. bench.py uses it as the denominator of
 * the L2 roofline of the configs whose BVH never leaves the caches. */
float Pt_MeasureL2ReadGBs(int64_t bytes, int32_t reps);

/* Tuning switches, also readable from the environment at PtSys_Init (PIMPT_PERSISTENT, PIMPT_FINISH):
 *   "persistent"  -1 choose the traversal kernels by BVH size (default), 0 one thread per ray, 1 persistent
 *                 warps with postponed leaves
 *   "finish"      live paths below which the tail of a wavefront call finishes in one launch (0 = never)
 *   "refit"       1: when PtScene_Update / Pt_Trace / LmPack_Bake find a newer registered description whose
 *                 geometry has the shape of the old one (same entities, same vertex counts), the scene keeps its
 *                 BVH topology: packed triangles and node boxes are recomputed bottom up from the moved vertices
 *                 instead of a new SAH build. 0 (default): always rebuild, as the reference
 *                 does (path_tracer.c:1028-1032). Hits are the same either way; a refitted tree gets slower the
 *                 further the geometry has moved from the pose it was built for.
 * Results do not depend on any of them. 0 on success. */
int32_t Pt_SetOption(const char* name, int32_t value);

const char* Pt_LastError(void);

#ifdef __cplusplus
}
#endif
#endif /* PIM_PT_H */
