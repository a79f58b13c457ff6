/* TEST INFRASTRUCTURE — C API of oracle/_ref/libpim_ref.so.
 *
 * libpim_ref.so = the reference's own src/rendering/path_tracer.c,
 * src/rendering/lightmap.c, src/math/{dist1d,sphgauss,pcg,lighting}.c compiled
 * unmodified from /root/reference, over the CPU ray-query restatement
 * (cpu_bvh.c, standing in for the absent Embree 3.13.5 binary) and the engine
 * stubs (ref_stubs.c). Only tests/, __graft_entry__.smoke() and bench.py's CPU
 * baseline may load it. Everything here is plain pointers and sizes for ctypes.
 */
#ifndef ORACLE_REF_HARNESS_H
#define ORACLE_REF_HARNESS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* RNG streams: work item i of Task_Run call number c inside oracle_begin(stream, pass)
 * starts from Prng_Get() state Pcg4(i, pass, seed, (stream<<4)|c) and path-tracer
 * context state Pcg4(i, pass, seed, ((stream<<4)|c) ^ 0x80000000).            */
enum
{
    ORACLE_STREAM_SCENE = 1,   /* call 0: CalcEmissionPdfFn, call 1: SetupLightGridFn */
    ORACLE_STREAM_TRACE = 2,   /* call 0: UpdateDistsFn, call 1: TraceFn */
    ORACLE_STREAM_RAYGEN = 3,  /* call 0: UpdateDistsFn, call 1: RayGenFn */
    ORACLE_STREAM_BAKE = 4,    /* call 0: UpdateDistsFn, call 1: BakeFn */
    ORACLE_STREAM_LUT = 5,     /* call 0: BakeBrdfFn */
    ORACLE_STREAM_RAYS = 6,    /* call 0: harness Pt_TraceRay batch */
    ORACLE_STREAM_LMPACK = 7,  /* LmPack_Pack's parallel loops */
    ORACLE_STREAM_CUBEBAKE = 8, /* call 0: cubemap.c BakeFn */
    ORACLE_STREAM_CONVOLVE = 9, /* call m: cubemap.c PrefilterFn of mip m */
    ORACLE_STREAM_SKY = 10,     /* call 0: BakeSkyFn (no random numbers) */
    ORACLE_STREAM_TONEMAP = 11, /* call 0: screenshot pixel loop (dither) */
};

void oracle_set_threads(int n);
int oracle_get_threads(void);
void oracle_set_seed(uint32_t seed);
/* 0: one call and one seeded random state per work item (parity); 1: the reference's chunking, task.c:35-64 (timing) */
void oracle_set_task_mode(int mode);
int oracle_get_task_mode(void);
void oracle_begin(uint32_t stream, uint32_t pass);
void oracle_frame_end(void);
void oracle_set_pt_dist_meters(float v);

void oracle_tables_clear(void);
int oracle_add_texture(int w, int h, const void* texels, int bytesPerTexel, int vkFormat);
int oracle_add_mesh(const float* pos4, const float* nrm4, const float* uv4, int length);
int oracle_add_entity(int meshId, const float* matrix16,
                      int albedoTex, int romeTex, int normalTex,
                      uint32_t flags, const float* meanFreePath4, float ior, float bumpiness);
void oracle_set_sky(int size, const float* faces);
void oracle_set_brdf_lut(int w, int h, const float* texels2);
void oracle_bake_brdf_lut(int w, int h, int samples, float* texelsOut2);

/* restatement of CmdCornellBox (src/rendering/render_system.c:1109-1306) on the
 * reference's own math headers; fills the tables above. */
void oracle_cornell_box(int spheres);

int oracle_entity_count(void);
void oracle_entity_get(int i, int* meshId, float* matrix16, int* texIds3,
                       uint32_t* flags, float* meanFreePath4, float* ior, float* bumpiness);
int oracle_mesh_get(int meshId, const float** pos4, const float** nrm4, const float** uv4);
int oracle_texture_get(int texId, int* w, int* h, const void** texels, int* vkFormat);
const int32_t* oracle_mesh_texindices(int meshId);

typedef struct OracleSceneInfo
{
    int32_t vertCount, matCount, emissiveCount;
    int32_t gridSize[3];
    float gridLo[4], gridHi[4];
    float cellsPerMeter;
    float rcpMajorant;
    const float* positions;     /* float4[vertCount] */
    const float* normals;       /* float4[vertCount] */
    const float* uvs;           /* float2[vertCount] */
    const int32_t* matIds;      /* [vertCount] */
    const int32_t* vertToEmit;  /* [vertCount] */
    const int32_t* emitToVert;  /* [emissiveCount] */
} OracleSceneInfo;

int oracle_sys_init(void);
void* oracle_scene_new(void);
void oracle_scene_del(void* scene);
void oracle_scene_info(const void* scene, OracleSceneInfo* info);
void oracle_scene_set_media(void* scene, const float* params14, int noiseOctaves, const float* phase3);
/* copies cell `cell`'s distribution; returns its length (0 = no distribution) */
int oracle_scene_dist(const void* scene, int cell, float* pdf, float* cdf, uint32_t* live);
/* overwrite a cell's pdf (length must match or be 0) and re-bake its cdf */
void oracle_scene_set_dist(void* scene, int cell, int length, const float* pdf);
void oracle_scene_set_live(void* scene, int cell, const uint32_t* live);

typedef struct OracleCamera { float position[4]; float rotation[4]; float zNear, zFar, fovy; } OracleCamera;
typedef struct OracleDof
{
    float aperture, focalLength; int32_t bladeCount; float bladeRot;
    float focalPlaneCurvature, autoFocusSpeed; int32_t autoFocus;
} OracleDof;

void oracle_dof_default(OracleDof* dof);
void oracle_camera_lookat(OracleCamera* cam, const float* eye3, const float* at3, float fovy);

/* one call of the reference's Pt_Trace (1 spp) on host float3 images */
void oracle_trace(void* scene, int w, int h, float* color3, float* albedo3, float* normal3,
                  float sampleWeight, const OracleCamera* cam, OracleDof* dof, uint32_t pass);
/* Pt_Intersect over n rays: ro4[i].xyz, rd4[i].xyz, rd4[i].w = tFar; out = PtRayHit[n] (48 B each) */
void oracle_intersect_n(const void* scene, const float* ro4, const float* rd4, float tNear,
                        int64_t n, void* hitsOut);
/* Pt_TraceRay over n rays, item i seeded as stream RAYS; out = PtResult[n] (9 floats each) */
void oracle_trace_rays(void* scene, const float* ro4, const float* rd4, int64_t n,
                       uint32_t pass, float* resultsOut9);
/* Pt_RayGen */
void oracle_raygen(void* scene, const float* origin4, int count, uint32_t pass,
                   float* colors4, float* directions4);

/* fetch counters of the canonical BVH, summed over worker threads since reset */
/* RtcOccluded16 (path_tracer.c:524-553) and RtcPointQuery (:597-616) as the reference calls them */
void oracle_occluded_n(const void* scene, const float* ro4, const float* rd4, float tNear, int64_t n, uint8_t* out);
void oracle_point_query_n(const void* scene, const float* pts4, int64_t n, float* distOut, int32_t* primOut);

void oracle_counters_reset(void);
void oracle_counters_get(uint64_t* nodes, uint64_t* tris, uint64_t* rays);

/* brute-force fp64 ground truth on the scene's triangles; flags[i]: 0 miss, 1 hit, 2 hit with near-tie */
void oracle_brute_n(const void* scene, const float* ro4, const float* rd4, float tNear,
                    int64_t n, float* tuv3, int32_t* prim, int32_t* flags);

/* lightmaps: LmPack_Pack (reference code, CPU pre-step) fills the global pack from the entity
 * tables; pages follow Lightmap_New's layout; LmPack_Bake is the path under test */
int oracle_lm_pack(int atlasSize, float texelsPerUnit, float distThresh, float degThresh);
int oracle_lm_size(void);
void oracle_lm_page(int i, float** probes5, float** position3, float** normal3, float** sampleCounts);
void oracle_lm_embed(void);    /* EmbedAttributes alone, on the pack oracle_lm_pack left */
int oracle_lm_slot(int i);      /* Lightmap.slot.index: what EmbedTaskFn matches texIndices.w against */
void oracle_lm_bake(void* scene, float timeSlice, int spp, uint32_t pass);
/* LmPack_Pack with its intermediate state recorded (ChartMaskFn's input and output, AtlasFn's input order and
 * output): one task thread, the reference's chunking, so AtlasFn's row-skipping search state runs through all
 * charts as it does in the engine's single chunk. Returns the chart count; the getters copy chart c. */
int oracle_lm_pack_capture(int atlasSize, float texelsPerUnit, float distThresh, float degThresh);
int oracle_lm_cap_atlas_count(void);
int oracle_lm_cap_chart(int c, int* nodeCount, float* area, int* maskSize2, int* atlasIndex, int* translation2);
void oracle_lm_cap_chart_data(int c, float* triBefore, float* triAfter, unsigned char* mask);
int oracle_lm_cap_sorted_index(int c);

/* SURVEY 8(f) rows (ref_cm_wrap.c). color: 6 x size*size float3; convolved: 6 x elemCount float4 */
int oracle_cubemap_elem_count(int size);
int oracle_cubemap_mip_count(int size);
void oracle_cubemap_bake(void* scene, int size, float* color, const float* origin4, float weight, uint32_t pass);
void oracle_cubemap_convolve(int size, float* color, float* convolved, uint32_t sampleCount, float weight, uint32_t pass);
int oracle_ambcube_bake(void* scene, float* cube24, const float* origin4, int samples, int prevSampleCount, uint32_t pass);
void oracle_bake_sky(int size, float* color, const float* sky9, const float* sunDir3, float sunLum, int steps);
void oracle_tonemap(const float* light4, int count, float exposure, uint32_t pass, uint32_t* rgba8);
void oracle_repack_rome(const float* mr4, int mrW, int mrH, const float* occ4, int occW, int occH,
                        const float* em4, int emW, int emH, int size, uint32_t* rgba8);

/* known-answer vectors of the 8(a) helpers (ref_kat.c) */
int oracle_kat_count(void);
const char* oracle_kat_name(int k, int* kin, int* kout);
void oracle_kat_run(int k, const float* in, int n, float* out);
void oracle_sg_axii(float* out20);

/* CalcStdDev (src/rendering/render_system.c:1374-1394) */
float oracle_stddev(const float* color3, int w, int h);

#ifdef __cplusplus
}
#endif
#endif
