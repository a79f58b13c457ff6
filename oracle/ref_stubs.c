/* TEST INFRASTRUCTURE — not part of the shipped library.
 *
 * Engine stubs that let the reference's own path_tracer.c / lightmap.c /
 * math/lighting.c (compiled unmodified from /root/reference, see Makefile)
 * link into oracle/_ref/libpim_ref.so without the engine around them:
 * allocator -> malloc, task system -> a pthread parallel-for with the
 * reference's chunk rule (src/threading/task.c: chunk = worksize / T^2),
 * Prng_Get -> deterministic per-work-item seeding (the reference seeds from
 * the wall clock, src/common/random.c:45-65, so it is not reproducible),
 * cvars/console/profiler/ImGui/serializer -> constants and no-ops, and the
 * Entities / Mesh / Texture / Cubemaps tables -> plain arrays filled by the
 * harness (ref_harness.c).
 */
#include "common/macro.h"
#include "math/types.h"
#include "common/random.h"
#include "common/cvar.h"
#include "common/console.h"
#include "common/profiler.h"
#include "common/guid.h"
#include "common/serialize.h"
#include "allocator/allocator.h"
#include "threading/task.h"
#include "rendering/mesh.h"
#include "rendering/texture.h"
#include "rendering/material.h"
#include "rendering/drawable.h"
#include "rendering/cubemap.h"
#include "math/lighting.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <stdarg.h>
#include <xmmintrin.h>
#include <pmmintrin.h>

#include "ref_harness.h"
#include "cpu_bvh.h"

/* ---------------------------------------------------------------- allocator */

typedef struct AllocHdr { i32 kind; i32 bytes; struct AllocHdr* next; u64 pad[6]; } AllocHdr;
_Static_assert(sizeof(AllocHdr) == 64, "header keeps 64-byte alignment");

static pthread_mutex_t s_tempMtx = PTHREAD_MUTEX_INITIALIZER;
static AllocHdr* s_tempList;

void* Mem_Alloc(EAlloc allocator, i32 bytes)
{
    if (bytes <= 0) return NULL;
    AllocHdr* h = (AllocHdr*)aligned_alloc(64, ((size_t)bytes + sizeof(AllocHdr) + 63) & ~(size_t)63);
    h->kind = (i32)allocator; h->bytes = bytes; h->next = NULL;
    if (allocator == EAlloc_Temp)
    {
        pthread_mutex_lock(&s_tempMtx);
        h->next = s_tempList; s_tempList = h;
        pthread_mutex_unlock(&s_tempMtx);
    }
    return h + 1;
}
void* Mem_Calloc(EAlloc allocator, i32 bytes)
{
    void* p = Mem_Alloc(allocator, bytes);
    if (p) memset(p, 0, (size_t)bytes);
    return p;
}
void Mem_Free(void* ptr)
{
    if (!ptr) return;
    AllocHdr* h = (AllocHdr*)ptr - 1;
    if (h->kind == EAlloc_Temp) return;   /* released at frame end */
    free(h);
}
void* Mem_Realloc(EAlloc allocator, void* prev, i32 bytes)
{
    if (!prev) return Mem_Alloc(allocator, bytes);
    if (bytes <= 0) { Mem_Free(prev); return NULL; }
    AllocHdr* h = (AllocHdr*)prev - 1;
    void* p = Mem_Alloc(allocator, bytes);
    memcpy(p, prev, (size_t)(h->bytes < bytes ? h->bytes : bytes));
    Mem_Free(prev);
    return p;
}
void oracle_frame_end(void)
{
    pthread_mutex_lock(&s_tempMtx);
    AllocHdr* h = s_tempList; s_tempList = NULL;
    pthread_mutex_unlock(&s_tempMtx);
    while (h) { AllocHdr* n = h->next; free(h); h = n; }
}

/* -------------------------------------------------------------- task system */

static i32 s_threads = 1;
static __thread i32 t_tid;
static Prng s_prngs[kMaxThreads];

static u32 s_seed = 1, s_pass = 0, s_stream = 0, s_call = 0;

/* How Task_Run hands out work. 0 (default, parity runs): one fn(i, i + 1) call per item, each with its own
 * seeded random state, so results do not depend on the thread count or on scheduling. 1 (timing runs, and
 * code whose behaviour depends on the chunking, like AtlasFn's per-call search state): the reference's own rule,
 * src/threading/task.c:35-64 — chunks of max(1, worksize / T^2) items claimed atomically, fn(begin, end) once per
 * chunk, one random state per thread seeded at the start of the call. */
static int s_taskMode = 0;
void oracle_set_task_mode(int mode) { s_taskMode = mode; }
int oracle_get_task_mode(void) { return s_taskMode; }

/* test hook around every Task_Run: phase 0 before, 1 after (ref_lm_wrap.c captures the packer's intermediate state) */
static void (*s_taskHook)(void* task, TaskExecuteFn fn, i32 worksize, int phase);
void oracle_set_task_hook(void (*hook)(void*, TaskExecuteFn, i32, int)) { s_taskHook = hook; }

void oracle_set_threads(int n)
{
    if (n < 1) n = 1;
    if (n > kMaxThreads) n = kMaxThreads;
    s_threads = n;
}
int oracle_get_threads(void) { return s_threads; }
void oracle_set_seed(uint32_t seed) { s_seed = seed; }
void oracle_begin(uint32_t stream, uint32_t pass) { s_stream = stream; s_pass = pass; s_call = 0; }

i32 Task_ThreadId(void) { return t_tid; }
i32 Task_ThreadCount(void) { return s_threads; }
Prng* Prng_Get(void) { return &s_prngs[t_tid]; }

/* defined beside the reference's static ms_contexts (ref_pt_wrap.c) */
void oracle_pt_set_ctx_rng(i32 tid, uint4 state);

static inline void seed_item(i32 tid, u32 item, u32 salt)
{
    /* ORACLE_SEED: both sides derive a work item's RNG state like this */
    s_prngs[tid].state = Pcg4(item, s_pass, s_seed, salt);
    oracle_pt_set_ctx_rng(tid, Pcg4(item, s_pass, s_seed, salt ^ 0x80000000u));
}

typedef struct RunCtx
{
    void* task; TaskExecuteFn fn; i32 worksize; i32 chunk; u32 salt;
    volatile i32 head;
} RunCtx;
typedef struct ThreadArg { RunCtx* run; i32 tid; } ThreadArg;

static u64 s_cntNodes, s_cntTris, s_cntRays;
void oracle_counters_reset(void) { s_cntNodes = s_cntTris = s_cntRays = 0; cpubvh_counters_reset(); }
void oracle_counters_get(uint64_t* nodes, uint64_t* tris, uint64_t* rays)
{
    *nodes = s_cntNodes; *tris = s_cntTris; *rays = s_cntRays;
}

static void run_worker(RunCtx* r, i32 tid)
{
    t_tid = tid;
    if (s_taskMode == 1) seed_item(tid, 0x40000000u + (u32)tid, r->salt);     /* one stream per thread and call */
    cpubvh_counters_reset();
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);          /* task.c:73-74 */
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    for (;;)
    {
        i32 b = __atomic_fetch_add(&r->head, r->chunk, __ATOMIC_RELAXED);
        if (b >= r->worksize) break;
        i32 e = b + r->chunk; if (e > r->worksize) e = r->worksize;
        if (s_taskMode == 1)
        {
            r->fn(r->task, b, e);
            continue;
        }
        for (i32 i = b; i < e; ++i)
        {
            seed_item(tid, (u32)i, r->salt);
            r->fn(r->task, i, i + 1);
        }
    }
    u64 n, t, y;
    cpubvh_counters_get(&n, &t, &y);
    __atomic_fetch_add(&s_cntNodes, n, __ATOMIC_RELAXED);
    __atomic_fetch_add(&s_cntTris, t, __ATOMIC_RELAXED);
    __atomic_fetch_add(&s_cntRays, y, __ATOMIC_RELAXED);
    cpubvh_counters_reset();
}
static void* thread_main(void* p)
{
    ThreadArg* a = (ThreadArg*)p;
    run_worker(a->run, a->tid);
    return NULL;
}

void Task_Run(void* task, TaskExecuteFn fn, i32 worksize)
{
    if (worksize <= 0) return;
    if (s_taskHook) s_taskHook(task, fn, worksize, 0);
    RunCtx r;
    r.task = task; r.fn = fn; r.worksize = worksize; r.head = 0;
    r.salt = (s_stream << 4) | (s_call & 0xf);
    s_call++;
    i32 T = s_threads;
    r.chunk = worksize / (T * T); if (r.chunk < 1) r.chunk = 1;
    i32 savedTid = t_tid;
    unsigned savedCsr = _mm_getcsr();   /* the caller may be a Python thread */
    if (T == 1 || savedTid != 0)
    {
        run_worker(&r, savedTid);
    }
    else
    {
        pthread_t th[kMaxThreads]; ThreadArg args[kMaxThreads];
        for (i32 i = 1; i < T; ++i)
        {
            args[i].run = &r; args[i].tid = i;
            pthread_create(&th[i], NULL, thread_main, &args[i]);
        }
        run_worker(&r, 0);
        for (i32 i = 1; i < T; ++i) pthread_join(th[i], NULL);
    }
    t_tid = savedTid;
    _mm_setcsr(savedCsr);
    if (s_taskHook) s_taskHook(task, fn, worksize, 1);
}

/* Task_Submit + TaskSys_Schedule + Task_Await (cubemap.c:271-300 queues one task per mip):
 * run each task when it is submitted, so call number c is the c-th submitted task */
void Task_Submit(void* task, TaskExecuteFn fn, i32 worksize) { Task_Run(task, fn, worksize); }
void TaskSys_Schedule(void) {}
void Task_Await(void* task) { (void)task; }

/* ------------------------------------------------- cvars / console / misc */

ConVar cv_pt_dist_meters;
ConVar cv_r_brdflut_spf;
static float s_ptDistMeters = 1.5f;      /* cvars.c:208-216 */
void oracle_set_pt_dist_meters(float v) { s_ptDistMeters = v; }
float ConVar_GetFloat(const ConVar* ptr) { return (ptr == &cv_pt_dist_meters) ? s_ptDistMeters : 0.0f; }
i32 ConVar_GetInt(const ConVar* ptr) { return (ptr == &cv_r_brdflut_spf) ? 10 : 0; }

void Con_Logf(LogSev sev, const char* tag, const char* fmt, ...)
{
    if (sev > LogSev_Warning) return;
    va_list ap; va_start(ap, fmt);
    fprintf(stderr, "[ref:%s] ", tag); vfprintf(stderr, fmt, ap); fputc('\n', stderr);
    va_end(ap);
}
void _ProfileBegin(ProfMark *const mark) { (void)mark; }
void _ProfileEnd(ProfMark *const mark) { (void)mark; }
double Time_Deltaf(void) { return 1.0 / 60.0; }

i32 SPrintf(char* dst, i32 size, const char* fmt, ...)
{
    va_list ap; va_start(ap, fmt);
    i32 n = vsnprintf(dst, (size_t)size, fmt, ap);
    va_end(ap);
    return n;
}

Guid VEC_CALL Guid_FromStr(char const *const pim_noalias str)
{
    Guid g = { 0 };
    u64 a = 1469598103934665603ull, b = 0x9E3779B97F4A7C15ull;
    for (const char* p = str; p && *p; ++p)
    {
        a = (a ^ (u8)*p) * 1099511628211ull;
        b = (b + (u8)*p) * 0xD6E8FEB86659FD93ull;
    }
    g.a = a; g.b = b;
    return g;
}

bool VEC_CALL Guid_IsNull(Guid x) { return !(x.a | x.b); }
i32 Guid_Find(Guid const *const pim_noalias ptr, i32 count, Guid key)
{
    for (i32 i = 0; i < count; ++i) { if (ptr[i].a == key.a && ptr[i].b == key.b) return i; }
    return -1;
}

/* serializer / ImGui: unreachable from the harness, present only to link */
SerObj* Ser_ReadFile(const char* f) { (void)f; return NULL; }
bool Ser_WriteFile(const char* f, SerObj* o) { (void)f; (void)o; return false; }
SerObj* SerObj_Dict(void) { return NULL; }
void SerObj_Del(SerObj* o) { (void)o; }
float Ser_GetFloat(SerObj* o, const char* k) { (void)o; (void)k; return 0.0f; }
i32 Ser_GetInt(SerObj* o, const char* k) { (void)o; (void)k; return 0; }
bool Ser_SetFloat(SerObj* o, const char* k, float v) { (void)o; (void)k; (void)v; return false; }
bool Ser_SetInt(SerObj* o, const char* k, i32 v) { (void)o; (void)k; (void)v; return false; }

bool igCheckbox() { return false; }
bool igColorEdit3() { return false; }
bool igExButton() { return false; }
bool igExCollapsingHeader1() { return false; }
bool igExSliderFloat() { return false; }
bool igExSliderInt() { return false; }
void igIndent() {}
void igUnindent() {}
bool igInputText() { return false; }
bool igSliderFloat() { return false; }
void igText() {}

/* ------------------------------------------------------------- scene tables */

#define MAX_TEX  4096
#define MAX_MESH 65536

static Texture s_textures[MAX_TEX];
static i32 s_texCount;
static Mesh s_meshes[MAX_MESH];
static i32 s_meshCount;
static Entities s_ents;
static i32 s_entCap;
static Cubemaps s_cubemaps;
static Cubemap s_sky;
static Guid s_skyName;

const float4 Cubemap_kForwards[6] =
{
    {1.0f, 0.0f, 0.0f}, {-1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f},
    {0.0f, -1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}, {0.0f, 0.0f, -1.0f},
};
const float4 Cubemap_kUps[6] =
{
    {0.0f, 1.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, -1.0f},
    {0.0f, 0.0f, -1.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 1.0f, 0.0f},
};
const float4 Cubemap_kRights[6] =
{
    {0.0f, 0.0f, -1.0f}, {0.0f, 0.0f, 1.0f}, {1.0f, 0.0f, 0.0f},
    {-1.0f, 0.0f, 0.0f}, {1.0f, 0.0f, 0.0f}, {-1.0f, 0.0f, 0.0f},
};

/* ids are index+1 so that a zeroed id means "no texture", as in the engine */
Texture* Texture_Get(TextureId id)
{
    i32 i = (i32)id.index - 1;
    return (i >= 0 && i < s_texCount) ? &s_textures[i] : NULL;
}
Mesh *const Mesh_Get(MeshId id)
{
    i32 i = (i32)id.index - 1;
    return (i >= 0 && i < s_meshCount) ? &s_meshes[i] : NULL;
}
Entities *const Entities_Get(void) { return &s_ents; }
Cubemaps* Cubemaps_Get(void) { return &s_cubemaps; }
i32 Cubemaps_Find(const Cubemaps* maps, Guid name)
{
    for (i32 i = 0; i < maps->count; ++i)
    {
        if (maps->names[i].a == name.a && maps->names[i].b == name.b) return i;
    }
    return -1;
}

bool Texture_New(Texture* tex, VkFormat format, VkSamplerAddressMode clamp, Guid name, TextureId* idOut)
{
    (void)clamp; (void)name;
    if (s_texCount >= MAX_TEX) return false;
    tex->format = format;
    s_textures[s_texCount] = *tex;
    TextureId id = { 0 }; id.version = 1; id.index = (u32)(s_texCount + 1);
    s_texCount++;
    *idOut = id;
    return true;
}
void Texture_Release(TextureId id) { (void)id; }
bool Texture_Upload(TextureId id) { (void)id; return true; }

void oracle_tables_clear(void)
{
    for (i32 i = 0; i < s_texCount; ++i) Mem_Free(s_textures[i].texels);
    s_texCount = 0;
    for (i32 i = 0; i < s_meshCount; ++i)
    {
        Mem_Free(s_meshes[i].positions); Mem_Free(s_meshes[i].normals);
        Mem_Free(s_meshes[i].uvs); Mem_Free(s_meshes[i].texIndices);
    }
    s_meshCount = 0;
    s_ents.count = 0;
    s_ents.modtime++;
}

int oracle_add_texture(int w, int h, const void* texels, int bytesPerTexel, int vkFormat)
{
    Texture tex = { 0 };
    tex.size.x = w; tex.size.y = h;
    tex.texels = Mem_Alloc(EAlloc_Texture, w * h * bytesPerTexel);
    memcpy(tex.texels, texels, (size_t)w * h * bytesPerTexel);
    TextureId id;
    Guid name = { 0 };
    if (!Texture_New(&tex, (VkFormat)vkFormat, 0, name, &id)) return 0;
    return (int)id.index;
}

int oracle_add_mesh(const float* pos4, const float* nrm4, const float* uv4, int length)
{
    if (s_meshCount >= MAX_MESH) return 0;
    Mesh m = { 0 };
    m.length = length;
    m.positions = Mem_Alloc(EAlloc_Perm, sizeof(float4) * length);
    m.normals = Mem_Alloc(EAlloc_Perm, sizeof(float4) * length);
    m.uvs = Mem_Alloc(EAlloc_Perm, sizeof(float4) * length);
    m.texIndices = Mem_Calloc(EAlloc_Perm, sizeof(int4) * length);
    memcpy(m.positions, pos4, sizeof(float4) * length);
    memcpy(m.normals, nrm4, sizeof(float4) * length);
    memcpy(m.uvs, uv4, sizeof(float4) * length);
    s_meshes[s_meshCount++] = m;
    return s_meshCount;
}

int oracle_add_entity(int meshId, const float* matrix16,
                      int albedoTex, int romeTex, int normalTex,
                      uint32_t flags, const float* meanFreePath4, float ior, float bumpiness)
{
    if (s_ents.count == s_entCap)
    {
        s_entCap = s_entCap ? s_entCap * 2 : 64;
        s_ents.meshes = Mem_Realloc(EAlloc_Perm, s_ents.meshes, sizeof(MeshId) * s_entCap);
        s_ents.materials = Mem_Realloc(EAlloc_Perm, s_ents.materials, sizeof(Material) * s_entCap);
        s_ents.matrices = Mem_Realloc(EAlloc_Perm, s_ents.matrices, sizeof(float4x4) * s_entCap);
    }
    i32 i = s_ents.count++;
    MeshId mid = { 0 }; mid.version = 1; mid.index = (u32)meshId;
    s_ents.meshes[i] = mid;
    Material mat;
    memset(&mat, 0, sizeof(mat));
    if (albedoTex) { mat.albedo.version = 1; mat.albedo.index = (u32)albedoTex; }
    if (romeTex) { mat.rome.version = 1; mat.rome.index = (u32)romeTex; }
    if (normalTex) { mat.normal.version = 1; mat.normal.index = (u32)normalTex; }
    mat.flags = flags;
    if (meanFreePath4) memcpy(&mat.meanFreePath, meanFreePath4, sizeof(float4));
    mat.ior = ior;
    mat.bumpiness = bumpiness;
    s_ents.materials[i] = mat;
    memcpy(&s_ents.matrices[i], matrix16, sizeof(float4x4));
    s_ents.modtime++;
    return i;
}

void oracle_set_sky(int size, const float* faces /* 6 * size*size * float3, or NULL */)
{
    for (int f = 0; f < 6; ++f) { Mem_Free(s_sky.color[f]); s_sky.color[f] = NULL; }
    s_cubemaps.count = 0;
    if (!faces || size <= 0) return;
    s_sky.size = size; s_sky.mipCount = 1;
    for (int f = 0; f < 6; ++f)
    {
        s_sky.color[f] = Mem_Alloc(EAlloc_Texture, sizeof(float3) * size * size);
        memcpy(s_sky.color[f], faces + (size_t)f * size * size * 3, sizeof(float3) * size * size);
    }
    s_skyName = Guid_FromStr("sky");
    s_cubemaps.count = 1;
    s_cubemaps.names = &s_skyName;
    s_cubemaps.cubemaps = &s_sky;
}

/* ------------------------------------------------------------------ BRDF LUT */

static float2* s_lutTexels;

void oracle_set_brdf_lut(int w, int h, const float* texels2)
{
    Mem_Free(s_lutTexels);
    s_lutTexels = Mem_Alloc(EAlloc_Texture, sizeof(float2) * w * h);
    memcpy(s_lutTexels, texels2, sizeof(float2) * w * h);
    g_BrdfLut.size.x = w; g_BrdfLut.size.y = h;
    g_BrdfLut.texels = s_lutTexels;
}

/* Bakes with the reference's own BakeBrdfFn (src/math/lighting.c:31-86). */
void oracle_bake_brdf_lut(int w, int h, int samples, float* texelsOut2)
{
    oracle_begin(5, 0);
    int2 size = { w, h };
    BrdfLut lut = BrdfLut_New(size, samples);
    memcpy(texelsOut2, lut.texels, sizeof(float2) * w * h);
    BrdfLut_Del(&lut);
}

i32 StrLen(const char* x) { return x ? (i32)strlen(x) : 0; }

/* ---------------------------------------------- stubs only lightmap.c links against */
#include "threading/mutex.h"
#include "common/cmd.h"
#include "assets/crate.h"
void Mutex_New(Mutex* mut) { pthread_mutex_init((pthread_mutex_t*)mut, NULL); }
void Mutex_Del(Mutex* mut) { pthread_mutex_destroy((pthread_mutex_t*)mut); }
void Mutex_Lock(Mutex* mut) { pthread_mutex_lock((pthread_mutex_t*)mut); }
void Mutex_Unlock(Mutex* mut) { pthread_mutex_unlock((pthread_mutex_t*)mut); }
_Static_assert(sizeof(Mutex) >= sizeof(pthread_mutex_t), "reference Mutex can hold a pthread mutex");
void cmd_reg(const char* name, const char* argDocs, const char* helpText, cmd_fn_t fn) { (void)name; (void)argDocs; (void)helpText; (void)fn; }
bool Mesh_Upload(MeshId id) { (void)id; return true; }
int stbi_write_png(char const* filename, int w, int h, int comp, const void* data, int stride) { (void)filename; (void)w; (void)h; (void)comp; (void)data; (void)stride; return 0; }
bool Crate_Get(Crate *const crate, Guid id, void* dst, i32 size) { (void)crate; (void)id; (void)dst; (void)size; return false; }
bool Crate_Set(Crate *const crate, Guid id, const void* src, i32 size) { (void)crate; (void)id; (void)src; (void)size; return false; }
vkrTextureId vkrTexTable_Alloc(VkImageViewType type, VkFormat format, VkSamplerAddressMode clamp, i32 width, i32 height, i32 depth, i32 layers, bool mips)
{
    (void)type; (void)format; (void)clamp; (void)width; (void)height; (void)depth; (void)layers; (void)mips;
    /* a fresh slot per call, as the engine's table gives: EmbedTaskFn (lightmap.c:968,991) tells the
     * pages apart by it */
    static u32 s_nextSlot = 1;
    vkrTextureId id = { 0 };
    id.index = s_nextSlot++;
    return id;
}
bool vkrTexTable_Free(vkrTextureId id) { (void)id; return true; }
bool vkrTexTable_Upload(vkrTextureId id, i32 layer, void const *const data, i32 bytes) { (void)id; (void)layer; (void)data; (void)bytes; return true; }

/* Mem_Push / Mem_Pop: the reference's alloca substitute (LIFO per thread) */
static __thread void* t_pushStack[64];
static __thread int t_pushTop;
void* Mem_Push(i32 bytes)
{
    void* p = aligned_alloc(64, ((size_t)(bytes > 0 ? bytes : 1) + 63) & ~(size_t)63);
    if (t_pushTop < 64) t_pushStack[t_pushTop++] = p;
    return p;
}
void Mem_Pop(i32 bytes)
{
    (void)bytes;
    if (t_pushTop > 0) free(t_pushStack[--t_pushTop]);
}
