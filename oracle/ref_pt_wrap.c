/* TEST INFRASTRUCTURE — compiles the reference's src/rendering/path_tracer.c
 * where it lies (no copy) and adds accessors to its file-static state so the
 * harness can seed the per-thread contexts and read the flattened scene. */
#include "rendering/path_tracer.c"

#include "ref_harness.h"
#include "cpu_bvh.h"

void oracle_pt_set_ctx_rng(i32 tid, uint4 state)
{
    ms_contexts[tid].rng.state = state;
}

const float* oracle_rtc_scene_verts(RTCScene hs, int* stride, int64_t* ntris);

int oracle_sys_init(void)
{
    PtSys_Init();
    return ms_device != NULL;
}

void* oracle_scene_new(void)
{
    oracle_begin(ORACLE_STREAM_SCENE, 0);
    PtScene* scene = PtScene_New();
    oracle_frame_end();
    return scene;
}

void oracle_scene_del(void* scene) { PtScene_Del((PtScene*)scene); }

void oracle_scene_info(const void* pscene, OracleSceneInfo* info)
{
    const PtScene* scene = (const PtScene*)pscene;
    memset(info, 0, sizeof(*info));
    info->vertCount = scene->vertCount;
    info->matCount = scene->matCount;
    info->emissiveCount = scene->emissiveCount;
    info->gridSize[0] = scene->lightGrid.size.x;
    info->gridSize[1] = scene->lightGrid.size.y;
    info->gridSize[2] = scene->lightGrid.size.z;
    memcpy(info->gridLo, &scene->lightGrid.bounds.lo, sizeof(float4));
    memcpy(info->gridHi, &scene->lightGrid.bounds.hi, sizeof(float4));
    info->cellsPerMeter = scene->lightGrid.cellsPerMeter;
    info->rcpMajorant = scene->mediaDesc.rcpMajorant;
    info->positions = (const float*)scene->positions;
    info->normals = (const float*)scene->normals;
    info->uvs = (const float*)scene->uvs;
    info->matIds = scene->matIds;
    info->vertToEmit = scene->vertToEmit;
    info->emitToVert = scene->emitToVert;
}

/* what media_desc_gui / media_desc_load do to scene->mediaDesc (path_tracer.c:1989-2026, 2080-2110):
 * set the user-facing fields, then media_desc_update. params14: constantColor.xyz, noiseColor.xyz,
 * constantMfp, noiseMfp, absorption, noiseGain, noiseLacunarity, noiseFreq, noiseScale, noiseHeight;
 * phase3: phaseDirA, phaseDirB, phaseBlend */
void oracle_scene_set_media(void* pscene, const float* params14, int noiseOctaves, const float* phase3)
{
    PtScene* scene = (PtScene*)pscene;
    PtMediaDesc* d = &scene->mediaDesc;
    d->constantColor = f4_v(params14[0], params14[1], params14[2], 2.0f);
    d->noiseColor = f4_v(params14[3], params14[4], params14[5], 2.0f);
    d->constantMfp = params14[6]; d->noiseMfp = params14[7]; d->absorption = params14[8];
    d->noiseOctaves = noiseOctaves;
    d->noiseGain = params14[9]; d->noiseLacunarity = params14[10]; d->noiseFreq = params14[11];
    d->noiseScale = params14[12]; d->noiseHeight = params14[13];
    d->phaseDirA = phase3[0]; d->phaseDirB = phase3[1]; d->phaseBlend = phase3[2];
    media_desc_update(d);
}

int oracle_scene_dist(const void* pscene, int cell, float* pdf, float* cdf, uint32_t* live)
{
    const PtScene* scene = (const PtScene*)pscene;
    const Dist1D* d = &scene->lightDists[cell];
    if (d->length > 0)
    {
        if (pdf) memcpy(pdf, d->pdf, sizeof(float) * d->length);
        if (cdf) memcpy(cdf, d->cdf, sizeof(float) * (d->length + 1));
        if (live) memcpy(live, d->live, sizeof(u32) * d->length);
    }
    return d->length;
}

void oracle_scene_set_dist(void* pscene, int cell, int length, const float* pdf)
{
    PtScene* scene = (PtScene*)pscene;
    Dist1D* d = &scene->lightDists[cell];
    if (d->length != length)
    {
        Dist1D_Del(d);
        if (length > 0) Dist1D_New(d, length);
    }
    if (length > 0)
    {
        memcpy(d->pdf, pdf, sizeof(float) * length);
        memset(d->live, 0, sizeof(u32) * length);
        d->sum = 0;
        Dist1D_Bake(d);
    }
}

void oracle_scene_set_live(void* pscene, int cell, const uint32_t* live)
{
    PtScene* scene = (PtScene*)pscene;
    Dist1D* d = &scene->lightDists[cell];
    if (d->length > 0) memcpy(d->live, live, sizeof(u32) * d->length);
}

void oracle_dof_default(OracleDof* dof)
{
    PtDofInfo d;
    memset(&d, 0, sizeof(d));
    DofInfo_New(&d);
    dof->aperture = d.aperture; dof->focalLength = d.focalLength;
    dof->bladeCount = d.bladeCount; dof->bladeRot = d.bladeRot;
    dof->focalPlaneCurvature = d.focalPlaneCurvature;
    dof->autoFocusSpeed = d.autoFocusSpeed; dof->autoFocus = d.autoFocus ? 1 : 0;
}

/* CmdTeleport + CmdLookat (src/rendering/render_system.c:1308-1346) */
void oracle_camera_lookat(OracleCamera* cam, const float* eye3, const float* at3, float fovy)
{
    float4 ro = { eye3[0], eye3[1], eye3[2], 0.0f };
    float4 at = { at3[0], at3[1], at3[2], 0.0f };
    float4 rd = f4_normalize3(f4_sub(at, ro));
    const float4 up = { 0.0f, 1.0f, 0.0f, 0.0f };
    quat rot = quat_lookat(rd, up);
    memcpy(cam->position, &ro, sizeof(float4));
    memcpy(cam->rotation, &rot, sizeof(float4));
    cam->zNear = 0.1f; cam->zFar = 500.0f; cam->fovy = fovy;
}

static void to_dof(PtDofInfo* d, const OracleDof* dof)
{
    memset(d, 0, sizeof(*d));
    d->aperture = dof->aperture; d->focalLength = dof->focalLength;
    d->bladeCount = dof->bladeCount; d->bladeRot = dof->bladeRot;
    d->focalPlaneCurvature = dof->focalPlaneCurvature;
    d->autoFocusSpeed = dof->autoFocusSpeed; d->autoFocus = dof->autoFocus != 0;
}

void oracle_trace(void* pscene, int w, int h, float* color3, float* albedo3, float* normal3,
                  float sampleWeight, const OracleCamera* ocam, OracleDof* odof, uint32_t pass)
{
    PtTrace trace;
    memset(&trace, 0, sizeof(trace));
    trace.color = (float3*)color3;
    trace.albedo = (float3*)albedo3;
    trace.normal = (float3*)normal3;
    trace.imageSize.x = w; trace.imageSize.y = h;
    trace.sampleWeight = sampleWeight;
    Camera cam;
    memset(&cam, 0, sizeof(cam));
    memcpy(&cam.position, ocam->position, sizeof(float4));
    memcpy(&cam.rotation, ocam->rotation, sizeof(float4));
    cam.zNear = ocam->zNear; cam.zFar = ocam->zFar; cam.fovy = ocam->fovy;
    PtDofInfo dof; to_dof(&dof, odof);
    oracle_begin(ORACLE_STREAM_TRACE, pass);
    Pt_Trace(&trace, &dof, (PtScene*)pscene, &cam);
    odof->focalLength = dof.focalLength;
    oracle_frame_end();
}

typedef struct IsectTask { Task task; const PtScene* scene; const float4* ro; const float4* rd; float tNear; PtRayHit* out; i64 base; } IsectTask;
static void IsectFn(void* p, i32 begin, i32 end)
{
    IsectTask* t = (IsectTask*)p;
    for (i32 i = begin; i < end; ++i)
    {
        i64 j = t->base + i;
        float4 rd = t->rd[j];
        float tFar = rd.w; rd.w = 0.0f;
        t->out[j] = Pt_Intersect(t->scene, t->ro[j], rd, t->tNear, tFar);
    }
}

static void* aligned_copy(const void* src, size_t bytes)
{
    void* p = aligned_alloc(64, (bytes + 63) & ~(size_t)63);
    memcpy(p, src, bytes);
    return p;
}

void oracle_intersect_n(const void* pscene, const float* ro4, const float* rd4, float tNear,
                        int64_t n, void* hitsOut)
{
    float4* ro = aligned_copy(ro4, sizeof(float4) * (size_t)n);
    float4* rd = aligned_copy(rd4, sizeof(float4) * (size_t)n);
    PtRayHit* out = aligned_alloc(64, (sizeof(PtRayHit) * (size_t)n + 63) & ~(size_t)63);
    oracle_begin(ORACLE_STREAM_RAYS, 0);
    for (i64 base = 0; base < n; base += (1 << 30))
    {
        IsectTask t; memset(&t, 0, sizeof(t));
        t.scene = (const PtScene*)pscene; t.ro = ro; t.rd = rd; t.tNear = tNear; t.out = out; t.base = base;
        i64 cnt = n - base; if (cnt > (1 << 30)) cnt = (1 << 30);
        Task_Run(&t, IsectFn, (i32)cnt);
    }
    memcpy(hitsOut, out, sizeof(PtRayHit) * (size_t)n);
    free(ro); free(rd); free(out);
}

typedef struct RaysTask { Task task; PtScene* scene; const float4* ro; const float4* rd; float* out; } RaysTask;
static void RaysFn(void* p, i32 begin, i32 end)
{
    RaysTask* t = (RaysTask*)p;
    for (i32 i = begin; i < end; ++i)
    {
        float4 rd = t->rd[i]; rd.w = 0.0f;
        PtResult r = Pt_TraceRay(t->scene, t->ro[i], rd);
        memcpy(t->out + 9 * (size_t)i, &r, sizeof(float) * 9);
    }
}

void oracle_trace_rays(void* pscene, const float* ro4, const float* rd4, int64_t n,
                       uint32_t pass, float* resultsOut9)
{
    float4* ro = aligned_copy(ro4, sizeof(float4) * (size_t)n);
    float4* rd = aligned_copy(rd4, sizeof(float4) * (size_t)n);
    RaysTask t; memset(&t, 0, sizeof(t));
    t.scene = (PtScene*)pscene; t.ro = ro; t.rd = rd; t.out = resultsOut9;
    oracle_begin(ORACLE_STREAM_RAYS, pass);
    Task_Run(&t, RaysFn, (i32)n);
    free(ro); free(rd);
}

void oracle_raygen(void* pscene, const float* origin4, int count, uint32_t pass,
                   float* colors4, float* directions4)
{
    float4 origin; memcpy(&origin, origin4, sizeof(float4));
    oracle_begin(ORACLE_STREAM_RAYGEN, pass);
    PtResults r = Pt_RayGen((PtScene*)pscene, origin, count);
    memcpy(colors4, r.colors, sizeof(float4) * (size_t)count);
    memcpy(directions4, r.directions, sizeof(float4) * (size_t)count);
    oracle_frame_end();
}

typedef struct BruteTask { Task task; const float* verts; int stride; int64_t ntris; const float* ro4; const float* rd4; float tNear; float* tuv; int32_t* prim; int32_t* flags; } BruteTask;
static void BruteFn(void* p, i32 begin, i32 end)
{
    BruteTask* t = (BruteTask*)p;
    for (i32 i = begin; i < end; ++i)
    {
        CpuHit hit;
        int r = brute_intersect_f64(t->verts, t->stride, t->ntris, t->ro4 + 4 * (size_t)i, t->rd4 + 4 * (size_t)i,
                                    t->tNear, t->rd4[4 * (size_t)i + 3], &hit);
        t->flags[i] = r; t->prim[i] = hit.prim;
        t->tuv[3 * (size_t)i + 0] = hit.t; t->tuv[3 * (size_t)i + 1] = hit.u; t->tuv[3 * (size_t)i + 2] = hit.v;
    }
}

void oracle_brute_n(const void* pscene, const float* ro4, const float* rd4, float tNear,
                    int64_t n, float* tuv3, int32_t* prim, int32_t* flags)
{
    const PtScene* scene = (const PtScene*)pscene;
    BruteTask t; memset(&t, 0, sizeof(t));
    t.verts = oracle_rtc_scene_verts(scene->rtcScene, &t.stride, &t.ntris);
    t.ro4 = ro4; t.rd4 = rd4; t.tNear = tNear; t.tuv = tuv3; t.prim = prim; t.flags = flags;
    oracle_begin(ORACLE_STREAM_RAYS, 0);
    Task_Run(&t, BruteFn, (i32)n);
}

/* the reference's RtcOccluded16 (path_tracer.c:524-553) over the shim's rtcOccluded16, 16 rays at a
 * time: ro4.w unused (tNear is a parameter), rd4.w = tFar; out[i] = 1 when occluded */
void oracle_occluded_n(const void* pscene, const float* ro4, const float* rd4, float tNear, int64_t n, uint8_t* out)
{
    const PtScene* scene = (const PtScene*)pscene;
    for (int64_t base = 0; base < n; base += 16)
    {
        float4 ros[16]; float4 rds[16]; bool visibles[16];
        for (int i = 0; i < 16; ++i)
        {
            const int64_t j = (base + i < n) ? base + i : n - 1;
            ros[i] = f4_v(ro4[4 * j], ro4[4 * j + 1], ro4[4 * j + 2], tNear);
            rds[i] = f4_v(rd4[4 * j], rd4[4 * j + 1], rd4[4 * j + 2], rd4[4 * j + 3]);
        }
        RtcOccluded16(scene->rtcScene, ros, rds, visibles);
        for (int i = 0; i < 16 && base + i < n; ++i) out[base + i] = visibles[i] ? 0 : 1;
    }
}

/* the reference's RtcPointQuery (path_tracer.c:597-616): pts4 = (x, y, z, radius); distOut = usr.distance
 * (kRcpEpsilon when nothing lies within the radius), primOut = usr.primID (-1 for none) */
void oracle_point_query_n(const void* pscene, const float* pts4, int64_t n, float* distOut, int32_t* primOut)
{
    const PtScene* scene = (const PtScene*)pscene;
    for (int64_t i = 0; i < n; ++i)
    {
        PointQueryUserData usr = RtcPointQuery(scene, f4_v(pts4[4 * i], pts4[4 * i + 1], pts4[4 * i + 2], pts4[4 * i + 3]));
        distOut[i] = usr.distance;
        primOut[i] = (usr.primID == RTC_INVALID_GEOMETRY_ID) ? -1 : (int32_t)usr.primID;
    }
}
