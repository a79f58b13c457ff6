/* TEST INFRASTRUCTURE — not part of the shipped library.
 *
 * Oracle for the SURVEY.md 8(f) rows that sit either side of the path tracer:
 *   N2  Cubemap_Bake / Cubemap_Convolve  src/rendering/cubemap.c:138-303 — the reference's own
 *       file, compiled where it lies (the engine-table functions in it are renamed so the
 *       harness tables of ref_stubs.c stay the ones the path tracer sees);
 *       AmbCube_Bake                     src/math/ambcube.c:5-33 (compiled by the Makefile)
 *   N3  BakeSkyFn                        src/rendering/render_system.c:403-425. That file is the
 *       whole renderer (Vulkan, window, UI) and does not compile here; the loop below restates
 *       its five lines around the reference's own Atmosphere() (src/math/atmosphere.h:79-181)
 *       and Cubemap_CalcDir (src/rendering/cubemap.h:190-207), both used as they are.
 *   N4  CmdScreenshot's pixel loop       src/rendering/render_system.c:707-727, restated around
 *       the reference's own Color_SceneToSDR / f4_GTTonemap / f4_sRGB_InverseEOTF / f4_rgba8
 *       (src/math/color.h). The 1/255 dither term reads the caller's serial Prng chain in the
 *       reference; here it is seeded per pixel like every other work item (ORACLE_SEED).
 * parity: the reference has no tests or golden vectors for these; the functions above are the
 * reference's code itself, the two restated loops are pinned by reading them side by side.
 */
#define Cubemap_kForwards refcm_kForwards
#define Cubemap_kUps refcm_kUps
#define Cubemap_kRights refcm_kRights
#define Cubemaps_Get refcm_Cubemaps_Get
#define Cubemaps_Add refcm_Cubemaps_Add
#define Cubemaps_Rm refcm_Cubemaps_Rm
#define Cubemaps_Find refcm_Cubemaps_Find
#include "rendering/cubemap.c"

#include "math/ambcube.h"
#include "math/atmosphere.h"
#include "math/color.h"
#include "rendering/texture.h"
#include "rendering/sampler.h"
#include "ref_harness.h"

/* layout of one cubemap handed across the harness: color = 6 faces of size*size float3,
 * convolved = 6 faces of elemCount float4 (mip chain, cubemap.c:102-113) */
int oracle_cubemap_elem_count(int size)
{
    const int2 s = { size, size };
    return CalcMipOffset(s, CalcMipCount(s)) + 1;
}
int oracle_cubemap_mip_count(int size)
{
    const int2 s = { size, size };
    return CalcMipCount(s);
}

static void cm_wrap(Cubemap* cm, int size, float* color, float* convolved)
{
    const int2 s = { size, size };
    cm->size = size;
    cm->mipCount = CalcMipCount(s);
    const i32 elemCount = CalcMipOffset(s, cm->mipCount) + 1;
    for (i32 f = 0; f < Cubeface_COUNT; ++f)
    {
        cm->color[f] = (float3*)(color + (size_t)f * size * size * 3);
        cm->convolved[f] = convolved ? (float4*)(convolved + (size_t)f * elemCount * 4) : NULL;
    }
}

void oracle_cubemap_bake(void* scene, int size, float* color, const float* origin4, float weight, uint32_t pass)
{
    Cubemap cm; cm_wrap(&cm, size, color, NULL);
    float4 origin; memcpy(&origin, origin4, sizeof(origin));
    oracle_begin(ORACLE_STREAM_CUBEBAKE, pass);
    Cubemap_Bake(&cm, (PtScene*)scene, origin, weight);
    oracle_frame_end();
}

void oracle_cubemap_convolve(int size, float* color, float* convolved, uint32_t sampleCount, float weight, uint32_t pass)
{
    Cubemap cm; cm_wrap(&cm, size, color, convolved);
    oracle_begin(ORACLE_STREAM_CONVOLVE, pass);
    Cubemap_Convolve(&cm, sampleCount, weight);
    oracle_frame_end();
}

int oracle_ambcube_bake(void* scene, float* cube24, const float* origin4, int samples, int prevSampleCount, uint32_t pass)
{
    AmbCube_t cube; memcpy(&cube, cube24, sizeof(cube));
    float4 origin; memcpy(&origin, origin4, sizeof(origin));
    oracle_begin(ORACLE_STREAM_RAYGEN, pass);
    int r = AmbCube_Bake((PtScene*)scene, &cube, origin, samples, prevSampleCount);
    oracle_frame_end();
    memcpy(cube24, &cube, sizeof(cube));
    return r;
}

/* sky9: rCrust, rAtmos, muR.xyz, rhoR, muM, rhoM, gM (SkyMedium, math/types.h) */
typedef struct SkyTask { Task task; SkyMedium sky; Cubemap* cm; float3 sunDir; float3 sunLum; i32 steps; } SkyTask;
static void SkyFn(void* p, i32 begin, i32 end)       /* render_system.c:403-425 */
{
    SkyTask* task = (SkyTask*)p;
    const i32 size = task->cm->size, len = size * size;
    for (i32 iWork = begin; iWork < end; ++iWork)
    {
        i32 iFace = iWork / len, iTexel = iWork % len;
        i32 x = iTexel % size, y = iTexel / size;
        float4 rd = Cubemap_CalcDir(size, iFace, i2_v(x, y), f2_0);
        float3 ro = f3_v(0.0f, task->sky.rCrust, 0.0f);
        task->cm->color[iFace][iTexel] = Atmosphere(&task->sky, ro, f4_f3(rd), task->sunDir, task->sunLum, task->steps);
    }
}
void oracle_bake_sky(int size, float* color, const float* sky9, const float* sunDir3, float sunLum, int steps)
{
    Cubemap cm; cm_wrap(&cm, size, color, NULL);
    SkyTask t; memset(&t, 0, sizeof(t));
    t.sky.rCrust = sky9[0]; t.sky.rAtmos = sky9[1];
    t.sky.muR = f3_v(sky9[2], sky9[3], sky9[4]); t.sky.rhoR = sky9[5];
    t.sky.muM = sky9[6]; t.sky.rhoM = sky9[7]; t.sky.gM = sky9[8];
    t.cm = &cm;
    t.sunDir = f3_v(sunDir3[0], sunDir3[1], sunDir3[2]);
    t.sunLum = f3_s(sunLum);
    t.steps = steps;
    oracle_begin(ORACLE_STREAM_SKY, 0);
    Task_Run(&t.task, SkyFn, Cubeface_COUNT * size * size);
    oracle_frame_end();
}

typedef struct ToneTask { Task task; const float4* light; R8G8B8A8_t* color; float exposure; } ToneTask;
static void ToneFn(void* p, i32 begin, i32 end)      /* render_system.c:707-727 */
{
    ToneTask* t = (ToneTask*)p;
    const GTTonemapParams gtp = { .P = 1.0f, .a = 1.0f, .m = 0.5f, .l = 0.4f, .c = 1.33f, .b = 0.0f };
    Prng* rng = Prng_Get();
    for (i32 i = begin; i < end; ++i)
    {
        float4 v = t->light[i];
        v = f4_mulvs(v, t->exposure);
        v = Color_SceneToSDR(v);
        v = f4_maxvs(v, 0.0f);
        v = f4_GTTonemap(v, gtp);
        v = f4_sRGB_InverseEOTF(v);
        v = f4_lerpvs(v, f4_rand(rng), 1.0f / 255.0f);
        v.w = 1.0f;
        t->color[i] = f4_rgba8(v);
    }
}
void oracle_tonemap(const float* light4, int count, float exposure, uint32_t pass, uint32_t* rgba8)
{
    ToneTask t; memset(&t, 0, sizeof(t));
    t.light = (const float4*)light4; t.color = (R8G8B8A8_t*)rgba8; t.exposure = exposure;
    oracle_begin(ORACLE_STREAM_TONEMAP, pass);
    Task_Run(&t.task, ToneFn, count);
    oracle_frame_end();
}

/* CreateRomeTexture's texel loop (src/rendering/gltf_model.c:1118-1139; the file itself needs cgltf and
 * stb_image and does not compile here), restated around the reference's own UvBilinearClamp_f4
 * (sampler.h:327-331), PackEmission (color.h:582-586) and f4_rgba8 (color.h:102-112). */
typedef struct RomeTask { Task task; Texture mr, occ, em; i32 size; R8G8B8A8_t* dst; } RomeTask;
static void RomeFn(void* p, i32 begin, i32 end)
{
    RomeTask* t = (RomeTask*)p;
    const i32 size = t->size;
    const float rcpSize = 1.0f / (size - 1);
    for (i32 index = begin; index < end; ++index)
    {
        i32 x = index % size, y = index / size;
        float2 uv = { x * rcpSize, y * rcpSize };
        float4 rome = f4_0;
        {
            float4 mr = UvBilinearClamp_f4(t->mr.texels, t->mr.size, uv);
            rome.x = mr.y;
            rome.z = mr.x;
        }
        rome.y = UvBilinearClamp_f4(t->occ.texels, t->occ.size, uv).x;
        rome.w = PackEmission(UvBilinearClamp_f4(t->em.texels, t->em.size, uv));
        t->dst[index] = f4_rgba8(rome);
    }
}
void oracle_repack_rome(const float* mr4, int mrW, int mrH, const float* occ4, int occW, int occH,
                        const float* em4, int emW, int emH, int size, uint32_t* rgba8)
{
    RomeTask t; memset(&t, 0, sizeof(t));
    t.mr.texels = (void*)mr4; t.mr.size = i2_v(mrW, mrH);
    t.occ.texels = (void*)occ4; t.occ.size = i2_v(occW, occH);
    t.em.texels = (void*)em4; t.em.size = i2_v(emW, emH);
    t.size = size; t.dst = (R8G8B8A8_t*)rgba8;
    oracle_begin(ORACLE_STREAM_SKY, 1);
    Task_Run(&t.task, RomeFn, size * size);
    oracle_frame_end();
}
