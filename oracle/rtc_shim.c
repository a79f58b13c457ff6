/* TEST INFRASTRUCTURE — the 16 Embree 3 link symbols the reference's
 * path_tracer.c uses (call sites: src/rendering/path_tracer.c:400-406, 443, 484,
 * 548, 614, 620-687, 1056), implemented on the CPU restatement in cpu_bvh.c
 * because the Embree 3.13.5 binary is absent. Semantics per the Embree manual:
 * rtcIntersect1 leaves the hit untouched on a miss and writes tfar/u/v/Ng/ids
 * on a hit; rtcOccluded* sets tfar = -inf on a hit. */
#include <embree3/rtcore.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "cpu_bvh.h"

typedef struct ShimGeom { int refs; float* verts; size_t vstride; size_t nverts; void* index; size_t ntris; } ShimGeom;
typedef struct ShimScene { int refs; ShimGeom* geom; CpuBvh* bvh; } ShimScene;
static int s_device;

RTCDevice rtcNewDevice(const char* config) { (void)config; return (RTCDevice)&s_device; }
void rtcReleaseDevice(RTCDevice d) { (void)d; }
enum RTCError rtcGetDeviceError(RTCDevice d) { (void)d; return RTC_ERROR_NONE; }
void rtcSetDeviceErrorFunction(RTCDevice d, RTCErrorFunction fn, void* user) { (void)d; (void)fn; (void)user; }

RTCScene rtcNewScene(RTCDevice d) { (void)d; ShimScene* s = calloc(1, sizeof(*s)); s->refs = 1; return (RTCScene)s; }
void rtcSetSceneBuildQuality(RTCScene s, enum RTCBuildQuality q) { (void)s; (void)q; }

static void geom_release(ShimGeom* g)
{
    if (g && --g->refs == 0) { free(g->verts); free(g->index); free(g); }
}
RTCGeometry rtcNewGeometry(RTCDevice d, enum RTCGeometryType type)
{
    (void)d; (void)type;
    ShimGeom* g = calloc(1, sizeof(*g)); g->refs = 1; return (RTCGeometry)g;
}
void* rtcSetNewGeometryBuffer(RTCGeometry hg, enum RTCBufferType type, unsigned slot,
                              enum RTCFormat format, size_t byteStride, size_t itemCount)
{
    (void)slot; (void)format;
    ShimGeom* g = (ShimGeom*)hg;
    if (type == RTC_BUFFER_TYPE_VERTEX)
    {
        g->verts = malloc(byteStride * itemCount + 16);
        g->vstride = byteStride / sizeof(float); g->nverts = itemCount;
        return g->verts;
    }
    g->index = malloc(byteStride * itemCount + 16);
    g->ntris = itemCount;
    return g->index;
}
void rtcCommitGeometry(RTCGeometry g) { (void)g; }
unsigned rtcAttachGeometry(RTCScene hs, RTCGeometry hg)
{
    ShimScene* s = (ShimScene*)hs; ShimGeom* g = (ShimGeom*)hg;
    g->refs++; s->geom = g; return 0;
}
void rtcReleaseGeometry(RTCGeometry g) { geom_release((ShimGeom*)g); }
void rtcCommitScene(RTCScene hs)
{
    ShimScene* s = (ShimScene*)hs;
    if (s->bvh) { cpubvh_free(s->bvh); s->bvh = NULL; }
    if (s->geom)
    {
        /* the reference always supplies an identity index buffer (path_tracer.c:676-679) */
        s->bvh = cpubvh_build(s->geom->verts, (int)s->geom->vstride, (int64_t)s->geom->ntris);
    }
}
void rtcReleaseScene(RTCScene hs)
{
    ShimScene* s = (ShimScene*)hs;
    if (s && --s->refs == 0) { cpubvh_free(s->bvh); geom_release(s->geom); free(s); }
}

void rtcIntersect1(RTCScene hs, struct RTCIntersectContext* ctx, struct RTCRayHit* rh)
{
    (void)ctx;
    ShimScene* s = (ShimScene*)hs;
    if (!s->bvh) return;
    float ro[3] = { rh->ray.org_x, rh->ray.org_y, rh->ray.org_z };
    float rd[3] = { rh->ray.dir_x, rh->ray.dir_y, rh->ray.dir_z };
    CpuHit hit;
    if (cpubvh_intersect(s->bvh, ro, rd, rh->ray.tnear, rh->ray.tfar, &hit))
    {
        rh->ray.tfar = hit.t;
        rh->hit.u = hit.u; rh->hit.v = hit.v;
        rh->hit.Ng_x = hit.ng[0]; rh->hit.Ng_y = hit.ng[1]; rh->hit.Ng_z = hit.ng[2];
        rh->hit.primID = (unsigned)hit.prim;
        rh->hit.geomID = 0;
        rh->hit.instID[0] = RTC_INVALID_GEOMETRY_ID;
    }
}

void rtcOccluded16(const int* valid, RTCScene hs, struct RTCIntersectContext* ctx, struct RTCRay16* r)
{
    (void)ctx;
    ShimScene* s = (ShimScene*)hs;
    if (!s->bvh) return;
    for (int i = 0; i < 16; ++i)
    {
        if (!valid[i]) continue;
        float ro[3] = { r->org_x[i], r->org_y[i], r->org_z[i] };
        float rd[3] = { r->dir_x[i], r->dir_y[i], r->dir_z[i] };
        if (cpubvh_occluded(s->bvh, ro, rd, r->tnear[i], r->tfar[i])) r->tfar[i] = -INFINITY;
    }
}

typedef struct PqAdapter { RTCPointQueryFunction fn; struct RTCPointQuery* q; struct RTCPointQueryContext* ctx; void* user; } PqAdapter;
static int pq_adapter(void* user, uint32_t prim, float* radius)
{
    PqAdapter* a = (PqAdapter*)user;
    struct RTCPointQueryFunctionArguments args;
    memset(&args, 0, sizeof(args));
    args.query = a->q; args.userPtr = a->user; args.primID = prim; args.geomID = 0;
    args.context = a->ctx; args.similarityScale = 1.0f;
    a->q->radius = *radius;
    bool changed = a->fn(&args);
    *radius = a->q->radius;
    return changed ? 1 : 0;
}
bool rtcPointQuery(RTCScene hs, struct RTCPointQuery* q, struct RTCPointQueryContext* ctx,
                   RTCPointQueryFunction fn, void* user)
{
    ShimScene* s = (ShimScene*)hs;
    if (!s->bvh || !fn) return false;
    PqAdapter a = { fn, q, ctx, user };
    float p[3] = { q->x, q->y, q->z };
    cpubvh_point_query(s->bvh, p, q->radius, pq_adapter, &a);
    return true;
}

/* exposed to the harness for brute-force checks */
const float* oracle_rtc_scene_verts(RTCScene hs, int* stride, int64_t* ntris)
{
    ShimScene* s = (ShimScene*)hs;
    *stride = (int)s->geom->vstride; *ntris = (int64_t)s->geom->ntris;
    return s->geom->verts;
}
int64_t oracle_rtc_scene_nodes(RTCScene hs) { return cpubvh_node_count(((ShimScene*)hs)->bvh); }
