/* TEST INFRASTRUCTURE — see cpu_bvh.h for what this restates and why. */
#include "cpu_bvh.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define LEAF_MAX 4
#define NBINS 16
#define WIDTH 8

typedef struct Box { float lo[3], hi[3]; } Box;

typedef struct Bin2
{
    Box box;
    int32_t left, right;   /* children (internal) */
    int64_t first;         /* leaf: first index into order[] */
    int32_t count;         /* leaf: >0 */
} Bin2;

typedef struct Wide
{
    float lox[WIDTH], loy[WIDTH], loz[WIDTH];
    float hix[WIDTH], hiy[WIDTH], hiz[WIDTH];
    int32_t child[WIDTH];  /* >= 0: wide node index; < 0: leaf, first = ~child */
    int32_t count[WIDTH];  /* leaf triangle count, 0 for internal */
    int32_t n;
} Wide;

struct CpuBvh
{
    const float* verts;
    int stride;
    int64_t ntris;
    float* tri;            /* 9 floats per triangle, in leaf order */
    int32_t* prim;         /* leaf order -> primitive id */
    Wide* nodes;
    int64_t nnodes;
};

static __thread uint64_t t_nodes, t_tris, t_rays;

void cpubvh_counters_reset(void) { t_nodes = t_tris = t_rays = 0; }
void cpubvh_counters_get(uint64_t* n, uint64_t* t, uint64_t* r)
{
    *n = t_nodes; *t = t_tris; *r = t_rays;
}

static inline void box_empty(Box* b)
{
    for (int a = 0; a < 3; ++a) { b->lo[a] = FLT_MAX; b->hi[a] = -FLT_MAX; }
}
static inline void box_grow_pt(Box* b, const float* p)
{
    for (int a = 0; a < 3; ++a)
    {
        if (p[a] < b->lo[a]) b->lo[a] = p[a];
        if (p[a] > b->hi[a]) b->hi[a] = p[a];
    }
}
static inline void box_grow(Box* b, const Box* o)
{
    for (int a = 0; a < 3; ++a)
    {
        if (o->lo[a] < b->lo[a]) b->lo[a] = o->lo[a];
        if (o->hi[a] > b->hi[a]) b->hi[a] = o->hi[a];
    }
}
static inline float box_area(const Box* b)
{
    float dx = b->hi[0] - b->lo[0], dy = b->hi[1] - b->lo[1], dz = b->hi[2] - b->lo[2];
    if (dx < 0.0f) return 0.0f;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}

typedef struct Builder
{
    const Box* tbox;
    const float* cent;     /* 3 per tri */
    int32_t* order;
    Bin2* nodes;
    int64_t nnodes, cap;
} Builder;

static int32_t build_rec(Builder* B, int64_t first, int64_t count)
{
    if (B->nnodes == B->cap)
    {
        B->cap *= 2;
        B->nodes = (Bin2*)realloc(B->nodes, sizeof(Bin2) * (size_t)B->cap);
    }
    int32_t me = (int32_t)B->nnodes++;
    Box box, cbox;
    box_empty(&box); box_empty(&cbox);
    for (int64_t i = first; i < first + count; ++i)
    {
        int32_t t = B->order[i];
        box_grow(&box, &B->tbox[t]);
        box_grow_pt(&cbox, &B->cent[3 * t]);
    }
    B->nodes[me].box = box;
    B->nodes[me].first = first;
    B->nodes[me].count = (int32_t)count;
    B->nodes[me].left = B->nodes[me].right = -1;
    if (count <= LEAF_MAX)
    {
        return me;
    }

    /* binned SAH over the three axes */
    float bestCost = FLT_MAX;
    int bestAxis = -1, bestSplit = -1;
    for (int a = 0; a < 3; ++a)
    {
        float lo = cbox.lo[a], ext = cbox.hi[a] - cbox.lo[a];
        if (!(ext > 0.0f)) continue;
        float scale = (float)NBINS / ext;
        Box bb[NBINS]; int64_t bc[NBINS];
        for (int b = 0; b < NBINS; ++b) { box_empty(&bb[b]); bc[b] = 0; }
        for (int64_t i = first; i < first + count; ++i)
        {
            int32_t t = B->order[i];
            int b = (int)((B->cent[3 * t + a] - lo) * scale);
            if (b < 0) b = 0;
            if (b > NBINS - 1) b = NBINS - 1;
            bc[b]++;
            box_grow(&bb[b], &B->tbox[t]);
        }
        float rarea[NBINS]; int64_t rcnt[NBINS];
        Box acc; box_empty(&acc); int64_t c = 0;
        for (int b = NBINS - 1; b > 0; --b)
        {
            box_grow(&acc, &bb[b]); c += bc[b];
            rarea[b] = box_area(&acc); rcnt[b] = c;
        }
        box_empty(&acc); c = 0;
        for (int b = 0; b < NBINS - 1; ++b)
        {
            box_grow(&acc, &bb[b]); c += bc[b];
            if (c == 0 || rcnt[b + 1] == 0) continue;
            float cost = box_area(&acc) * (float)c + rarea[b + 1] * (float)rcnt[b + 1];
            if (cost < bestCost) { bestCost = cost; bestAxis = a; bestSplit = b; }
        }
    }

    int64_t mid;
    if (bestAxis < 0)
    {
        mid = first + count / 2;   /* all centroids coincide: median split */
    }
    else
    {
        float lo = cbox.lo[bestAxis];
        float scale = (float)NBINS / (cbox.hi[bestAxis] - lo);
        int64_t i = first, j = first + count - 1;
        while (i <= j)
        {
            int32_t t = B->order[i];
            int b = (int)((B->cent[3 * t + bestAxis] - lo) * scale);
            if (b < 0) b = 0;
            if (b > NBINS - 1) b = NBINS - 1;
            if (b <= bestSplit) { ++i; }
            else { int32_t tmp = B->order[i]; B->order[i] = B->order[j]; B->order[j] = tmp; --j; }
        }
        mid = i;
        if (mid == first || mid == first + count) mid = first + count / 2;
    }
    int32_t l = build_rec(B, first, mid - first);
    int32_t r = build_rec(B, mid, first + count - mid);
    B->nodes[me].left = l;
    B->nodes[me].right = r;
    B->nodes[me].count = 0;
    return me;
}

typedef struct Collapse
{
    const Bin2* bin;
    Wide* wide;
    int64_t n, cap;
} Collapse;

static int32_t collapse_rec(Collapse* C, int32_t root)
{
    if (C->n == C->cap)
    {
        C->cap *= 2;
        C->wide = (Wide*)realloc(C->wide, sizeof(Wide) * (size_t)C->cap);
    }
    int32_t me = (int32_t)C->n++;
    int32_t kids[WIDTH];
    int n = 0;
    const Bin2* b = &C->bin[root];
    if (b->count > 0) { kids[n++] = root; }
    else { kids[n++] = b->left; kids[n++] = b->right; }
    while (n < WIDTH)
    {
        int best = -1; float bestArea = -1.0f;
        for (int i = 0; i < n; ++i)
        {
            const Bin2* k = &C->bin[kids[i]];
            if (k->count > 0) continue;
            float a = box_area(&k->box);
            if (a > bestArea) { bestArea = a; best = i; }
        }
        if (best < 0) break;
        const Bin2* k = &C->bin[kids[best]];
        kids[best] = k->left;
        kids[n++] = k->right;
    }
    Wide w;
    memset(&w, 0, sizeof(w));
    w.n = n;
    for (int i = 0; i < WIDTH; ++i)
    {
        w.lox[i] = w.loy[i] = w.loz[i] = FLT_MAX;
        w.hix[i] = w.hiy[i] = w.hiz[i] = -FLT_MAX;
        w.child[i] = 0; w.count[i] = 0;
    }
    for (int i = 0; i < n; ++i)
    {
        const Bin2* k = &C->bin[kids[i]];
        w.lox[i] = k->box.lo[0]; w.loy[i] = k->box.lo[1]; w.loz[i] = k->box.lo[2];
        w.hix[i] = k->box.hi[0]; w.hiy[i] = k->box.hi[1]; w.hiz[i] = k->box.hi[2];
    }
    C->wide[me] = w;
    for (int i = 0; i < n; ++i)
    {
        const Bin2* k = &C->bin[kids[i]];
        if (k->count > 0)
        {
            C->wide[me].child[i] = ~(int32_t)k->first;
            C->wide[me].count[i] = k->count;
        }
        else
        {
            int32_t c = collapse_rec(C, kids[i]);
            C->wide[me].child[i] = c;
        }
    }
    return me;
}

CpuBvh* cpubvh_build(const float* verts, int stride, int64_t ntris)
{
    CpuBvh* bvh = (CpuBvh*)calloc(1, sizeof(*bvh));
    bvh->verts = verts; bvh->stride = stride; bvh->ntris = ntris;
    if (ntris <= 0) return bvh;

    Box* tbox = (Box*)malloc(sizeof(Box) * (size_t)ntris);
    float* cent = (float*)malloc(sizeof(float) * 3 * (size_t)ntris);
    int32_t* order = (int32_t*)malloc(sizeof(int32_t) * (size_t)ntris);
    for (int64_t t = 0; t < ntris; ++t)
    {
        Box b; box_empty(&b);
        for (int k = 0; k < 3; ++k) box_grow_pt(&b, verts + (size_t)(3 * t + k) * stride);
        tbox[t] = b;
        for (int a = 0; a < 3; ++a) cent[3 * t + a] = 0.5f * (b.lo[a] + b.hi[a]);
        order[t] = (int32_t)t;
    }
    Builder B = { tbox, cent, order, NULL, 0, 0 };
    B.cap = 2 * ntris + 16;
    B.nodes = (Bin2*)malloc(sizeof(Bin2) * (size_t)B.cap);
    build_rec(&B, 0, ntris);

    Collapse C = { B.nodes, NULL, 0, 0 };
    C.cap = ntris / 2 + 16;
    C.wide = (Wide*)malloc(sizeof(Wide) * (size_t)C.cap);
    collapse_rec(&C, 0);

    bvh->nodes = C.wide;
    bvh->nnodes = C.n;
    bvh->prim = order;
    bvh->tri = (float*)malloc(sizeof(float) * 9 * (size_t)ntris);
    for (int64_t i = 0; i < ntris; ++i)
    {
        int32_t t = order[i];
        for (int k = 0; k < 3; ++k)
            for (int a = 0; a < 3; ++a)
                bvh->tri[9 * i + 3 * k + a] = verts[(size_t)(3 * t + k) * stride + a];
    }
    free(B.nodes); free(tbox); free(cent);
    return bvh;
}

void cpubvh_free(CpuBvh* bvh)
{
    if (bvh)
    {
        free(bvh->tri); free(bvh->prim); free(bvh->nodes); free(bvh);
    }
}

int64_t cpubvh_node_count(const CpuBvh* bvh) { return bvh->nnodes; }

/* two-sided Moller-Trumbore, fp32 */
static inline int tri_test(const float* T, const float ro[3], const float rd[3],
                           float tnear, float tfar, float* tOut, float* uOut, float* vOut)
{
    float e1x = T[3] - T[0], e1y = T[4] - T[1], e1z = T[5] - T[2];
    float e2x = T[6] - T[0], e2y = T[7] - T[1], e2z = T[8] - T[2];
    float px = rd[1] * e2z - rd[2] * e2y;
    float py = rd[2] * e2x - rd[0] * e2z;
    float pz = rd[0] * e2y - rd[1] * e2x;
    float det = e1x * px + e1y * py + e1z * pz;
    if (det == 0.0f) return 0;
    float inv = 1.0f / det;
    float tx = ro[0] - T[0], ty = ro[1] - T[1], tz = ro[2] - T[2];
    float u = (tx * px + ty * py + tz * pz) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return 0;
    float qx = ty * e1z - tz * e1y;
    float qy = tz * e1x - tx * e1z;
    float qz = tx * e1y - ty * e1x;
    float v = (rd[0] * qx + rd[1] * qy + rd[2] * qz) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return 0;
    float t = (e2x * qx + e2y * qy + e2z * qz) * inv;
    if (!(t >= tnear && t <= tfar)) return 0;
    *tOut = t; *uOut = u; *vOut = v;
    return 1;
}

static inline void tri_normal(const float* T, float ng[3])
{
    float e1x = T[3] - T[0], e1y = T[4] - T[1], e1z = T[5] - T[2];
    float e2x = T[6] - T[0], e2y = T[7] - T[1], e2z = T[8] - T[2];
    ng[0] = e1y * e2z - e1z * e2y;
    ng[1] = e1z * e2x - e1x * e2z;
    ng[2] = e1x * e2y - e1y * e2x;
}

#define STACK_MAX 256

static int traverse(const CpuBvh* bvh, const float ro[3], const float rd[3],
                    float tnear, float tfar, CpuHit* hit, int anyHit)
{
    t_rays++;
    if (hit) { hit->prim = -1; hit->t = tfar; }
    if (bvh->ntris <= 0) return 0;
    float idx = 1.0f / rd[0], idy = 1.0f / rd[1], idz = 1.0f / rd[2];
    int32_t stack[STACK_MAX];
    float stackT[STACK_MAX];
    int sp = 0;
    stack[sp] = 0; stackT[sp] = tnear; ++sp;
    int found = 0;
    while (sp > 0)
    {
        --sp;
        if (stackT[sp] > tfar) continue;
        const Wide* w = &bvh->nodes[stack[sp]];
        t_nodes++;
        float tmin[WIDTH]; int hitIdx[WIDTH]; int nh = 0;
        float te[WIDTH]; int ok[WIDTH];
        for (int i = 0; i < WIDTH; ++i)
        {
            float t0x = (w->lox[i] - ro[0]) * idx, t1x = (w->hix[i] - ro[0]) * idx;
            float t0y = (w->loy[i] - ro[1]) * idy, t1y = (w->hiy[i] - ro[1]) * idy;
            float t0z = (w->loz[i] - ro[2]) * idz, t1z = (w->hiz[i] - ro[2]) * idz;
            float nx = fminf(t0x, t1x), fx = fmaxf(t0x, t1x);
            float ny = fminf(t0y, t1y), fy = fmaxf(t0y, t1y);
            float nz = fminf(t0z, t1z), fz = fmaxf(t0z, t1z);
            float tn = fmaxf(fmaxf(nx, ny), fmaxf(nz, tnear));
            float tf = fminf(fminf(fx, fy), fminf(fz, tfar));
            /* conservative: widen by a few ulps so fp32 slabs never cull a true hit */
            te[i] = tn;
            ok[i] = (i < w->n) && (tn <= tf * 1.0000004f + 1e-30f);
        }
        for (int i = 0; i < w->n; ++i)
        {
            if (ok[i]) { tmin[nh] = te[i]; hitIdx[nh] = i; ++nh; }
        }
        /* insertion sort, ascending entry distance */
        for (int i = 1; i < nh; ++i)
        {
            float kt = tmin[i]; int ki = hitIdx[i]; int j = i - 1;
            while (j >= 0 && tmin[j] > kt) { tmin[j + 1] = tmin[j]; hitIdx[j + 1] = hitIdx[j]; --j; }
            tmin[j + 1] = kt; hitIdx[j + 1] = ki;
        }
        /* leaves first (in order), internal children pushed far-to-near */
        for (int k = 0; k < nh; ++k)
        {
            int i = hitIdx[k];
            if (w->count[i] > 0)
            {
                if (tmin[k] > tfar) continue;
                int32_t first = ~w->child[i];
                for (int32_t j = 0; j < w->count[i]; ++j)
                {
                    t_tris++;
                    float t, u, v;
                    const float* T = &bvh->tri[9 * (size_t)(first + j)];
                    if (tri_test(T, ro, rd, tnear, tfar, &t, &u, &v))
                    {
                        if (anyHit) return 1;
                        int32_t prim = bvh->prim[first + j];
                        /* ties resolve to the lowest primitive id, like the brute force */
                        if (!found || t < hit->t || (t == hit->t && prim < hit->prim))
                        {
                            tfar = t; found = 1;
                            hit->t = t; hit->u = u; hit->v = v;
                            hit->prim = prim;
                            tri_normal(T, hit->ng);
                        }
                    }
                }
            }
        }
        for (int k = nh - 1; k >= 0; --k)
        {
            int i = hitIdx[k];
            if (w->count[i] == 0 && sp < STACK_MAX)
            {
                stack[sp] = w->child[i]; stackT[sp] = tmin[k]; ++sp;
            }
        }
    }
    return found;
}

int cpubvh_intersect(const CpuBvh* bvh, const float ro[3], const float rd[3],
                     float tnear, float tfar, CpuHit* hit)
{
    return traverse(bvh, ro, rd, tnear, tfar, hit, 0);
}

int cpubvh_occluded(const CpuBvh* bvh, const float ro[3], const float rd[3],
                    float tnear, float tfar)
{
    return traverse(bvh, ro, rd, tnear, tfar, NULL, 1);
}

void cpubvh_point_query(const CpuBvh* bvh, const float p[3], float radius,
                        CpuPointQueryFn fn, void* user)
{
    if (bvh->ntris <= 0) return;
    int32_t stack[STACK_MAX];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0)
    {
        const Wide* w = &bvh->nodes[stack[--sp]];
        for (int i = 0; i < w->n; ++i)
        {
            float dx = fmaxf(fmaxf(w->lox[i] - p[0], p[0] - w->hix[i]), 0.0f);
            float dy = fmaxf(fmaxf(w->loy[i] - p[1], p[1] - w->hiy[i]), 0.0f);
            float dz = fmaxf(fmaxf(w->loz[i] - p[2], p[2] - w->hiz[i]), 0.0f);
            float d2 = dx * dx + dy * dy + dz * dz;
            if (d2 > radius * radius) continue;
            if (w->count[i] > 0)
            {
                int32_t first = ~w->child[i];
                for (int32_t j = 0; j < w->count[i]; ++j)
                {
                    fn(user, (uint32_t)bvh->prim[first + j], &radius);
                }
            }
            else if (sp < STACK_MAX)
            {
                stack[sp++] = w->child[i];
            }
        }
    }
}

int brute_intersect_f64(const float* verts, int stride, int64_t ntris,
                        const float ro[3], const float rd[3],
                        float tnear, float tfar, CpuHit* hit)
{
    double bestT = (double)tfar;
    double second = DBL_MAX;
    int found = 0;
    hit->prim = -1; hit->t = tfar; hit->u = hit->v = 0.0f;
    for (int64_t i = 0; i < ntris; ++i)
    {
        const float* a = verts + (size_t)(3 * i + 0) * stride;
        const float* b = verts + (size_t)(3 * i + 1) * stride;
        const float* c = verts + (size_t)(3 * i + 2) * stride;
        double e1x = (double)b[0] - a[0], e1y = (double)b[1] - a[1], e1z = (double)b[2] - a[2];
        double e2x = (double)c[0] - a[0], e2y = (double)c[1] - a[1], e2z = (double)c[2] - a[2];
        double px = rd[1] * e2z - rd[2] * e2y;
        double py = rd[2] * e2x - rd[0] * e2z;
        double pz = rd[0] * e2y - rd[1] * e2x;
        double det = e1x * px + e1y * py + e1z * pz;
        if (det == 0.0) continue;
        double inv = 1.0 / det;
        double tx = (double)ro[0] - a[0], ty = (double)ro[1] - a[1], tz = (double)ro[2] - a[2];
        double u = (tx * px + ty * py + tz * pz) * inv;
        if (!(u >= 0.0 && u <= 1.0)) continue;
        double qx = ty * e1z - tz * e1y;
        double qy = tz * e1x - tx * e1z;
        double qz = tx * e1y - ty * e1x;
        double v = (rd[0] * qx + rd[1] * qy + rd[2] * qz) * inv;
        if (!(v >= 0.0 && u + v <= 1.0)) continue;
        double t = (e2x * qx + e2y * qy + e2z * qz) * inv;
        if (!(t >= (double)tnear && t <= (double)tfar)) continue;
        if (!found || t < bestT)
        {
            if (found) second = bestT;
            bestT = t; found = 1;
            hit->t = (float)t; hit->u = (float)u; hit->v = (float)v;
            hit->prim = (int32_t)i;
            hit->ng[0] = (float)(e1y * e2z - e1z * e2y);
            hit->ng[1] = (float)(e1z * e2x - e1x * e2z);
            hit->ng[2] = (float)(e1x * e2y - e1y * e2x);
        }
        else if (t < second)
        {
            second = t;
        }
    }
    if (!found) return 0;
    if (second < DBL_MAX && (second - bestT) <= 1e-6 * fmax(1.0, fabs(bestT))) return 2;
    return 1;
}

/* Multi-threaded closest-hit batch for timing the CPU baseline and for the canonical
 * fetch counts: ro4/rd4 are float4 per ray, rd4.w = tFar. */
#include <pthread.h>

typedef struct BatchArg
{
    const CpuBvh* bvh; const float* ro4; const float* rd4; float tnear; int64_t begin, end;
    int32_t* prim; uint64_t nodes, tris;
} BatchArg;

static void* batch_main(void* p)
{
    BatchArg* a = (BatchArg*)p;
    cpubvh_counters_reset();
    for (int64_t i = a->begin; i < a->end; ++i)
    {
        CpuHit hit;
        int found = cpubvh_intersect(a->bvh, a->ro4 + 4 * i, a->rd4 + 4 * i, a->tnear, a->rd4[4 * i + 3], &hit);
        a->prim[i] = found ? hit.prim : -1;
    }
    uint64_t rays;
    cpubvh_counters_get(&a->nodes, &a->tris, &rays);
    return NULL;
}

void cpubvh_intersect_batch(const CpuBvh* bvh, const float* ro4, const float* rd4, float tnear, int64_t n,
                            int threads, int32_t* prim, uint64_t* nodes, uint64_t* tris)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256]; BatchArg args[256];
    for (int t = 0; t < threads; ++t)
    {
        args[t].bvh = bvh; args[t].ro4 = ro4; args[t].rd4 = rd4; args[t].tnear = tnear; args[t].prim = prim;
        args[t].begin = n * t / threads; args[t].end = n * (t + 1) / threads;
        pthread_create(&th[t], NULL, batch_main, &args[t]);
    }
    *nodes = *tris = 0;
    for (int t = 0; t < threads; ++t)
    {
        pthread_join(th[t], NULL);
        *nodes += args[t].nodes; *tris += args[t].tris;
    }
}

/* ---- fp64 ground truth at scale --------------------------------------------------------------
 * The all-triangles loop above is O(rays x triangles): fine for a few thousand rays against
 * 10 M triangles, not for 100 K. f64grid_* is a second ground truth that shares no code or
 * structure with any BVH: triangles are binned (by their slightly padded boxes) into every cell
 * of a uniform grid they overlap, a ray walks the cells it crosses in order (3-D DDA in fp64)
 * and runs the same fp64 Moller-Trumbore on each cell's triangles, stopping once the next cell
 * starts beyond the closest hit. The tests check it ray for ray against the all-triangles loop
 * on a subsample before using it as the reference for the rest. */

static int mt_f64(const float* a, const float* b, const float* c, const float ro[3], const float rd[3],
                  double tnear, double tfar, double* tOut, double* uOut, double* vOut)
{
    double e1x = (double)b[0] - a[0], e1y = (double)b[1] - a[1], e1z = (double)b[2] - a[2];
    double e2x = (double)c[0] - a[0], e2y = (double)c[1] - a[1], e2z = (double)c[2] - a[2];
    double px = rd[1] * e2z - rd[2] * e2y;
    double py = rd[2] * e2x - rd[0] * e2z;
    double pz = rd[0] * e2y - rd[1] * e2x;
    double det = e1x * px + e1y * py + e1z * pz;
    if (det == 0.0) return 0;
    double inv = 1.0 / det;
    double tx = (double)ro[0] - a[0], ty = (double)ro[1] - a[1], tz = (double)ro[2] - a[2];
    double u = (tx * px + ty * py + tz * pz) * inv;
    if (!(u >= 0.0 && u <= 1.0)) return 0;
    double qx = ty * e1z - tz * e1y;
    double qy = tz * e1x - tx * e1z;
    double qz = tx * e1y - ty * e1x;
    double v = (rd[0] * qx + rd[1] * qy + rd[2] * qz) * inv;
    if (!(v >= 0.0 && u + v <= 1.0)) return 0;
    double t = (e2x * qx + e2y * qy + e2z * qz) * inv;
    if (!(t >= tnear && t <= tfar)) return 0;
    *tOut = t; *uOut = u; *vOut = v;
    return 1;
}

typedef struct F64Grid
{
    const float* verts; int stride; int64_t ntris;
    double lo[3], cell; int n[3];
    int64_t* start;     /* n[0]*n[1]*n[2] + 1 */
    int32_t* items;
} F64Grid;

static void tri_cells(const F64Grid* g, int64_t i, int c0[3], int c1[3])
{
    const float* a = g->verts + (size_t)(3 * i + 0) * g->stride;
    const float* b = g->verts + (size_t)(3 * i + 1) * g->stride;
    const float* c = g->verts + (size_t)(3 * i + 2) * g->stride;
    const double pad = 1e-3 * g->cell;
    for (int k = 0; k < 3; ++k)
    {
        double lo = fmin(a[k], fmin(b[k], c[k])) - pad, hi = fmax(a[k], fmax(b[k], c[k])) + pad;
        int i0 = (int)floor((lo - g->lo[k]) / g->cell), i1 = (int)floor((hi - g->lo[k]) / g->cell);
        c0[k] = i0 < 0 ? 0 : (i0 >= g->n[k] ? g->n[k] - 1 : i0);
        c1[k] = i1 < 0 ? 0 : (i1 >= g->n[k] ? g->n[k] - 1 : i1);
    }
}

F64Grid* f64grid_build(const float* verts, int stride, int64_t ntris, double cell)
{
    F64Grid* g = (F64Grid*)calloc(1, sizeof(F64Grid));
    g->verts = verts; g->stride = stride; g->ntris = ntris; g->cell = cell;
    double lo[3] = { DBL_MAX, DBL_MAX, DBL_MAX }, hi[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };
    for (int64_t v = 0; v < 3 * ntris; ++v)
        for (int k = 0; k < 3; ++k)
        {
            double x = verts[(size_t)v * stride + k];
            if (x < lo[k]) lo[k] = x;
            if (x > hi[k]) hi[k] = x;
        }
    for (int k = 0; k < 3; ++k)
    {
        g->lo[k] = lo[k] - 0.01 * cell;
        int n = (int)ceil((hi[k] + 0.01 * cell - g->lo[k]) / cell);
        g->n[k] = n < 1 ? 1 : n;
    }
    const int64_t cells = (int64_t)g->n[0] * g->n[1] * g->n[2];
    g->start = (int64_t*)calloc((size_t)cells + 1, sizeof(int64_t));
    for (int64_t i = 0; i < ntris; ++i)
    {
        int c0[3], c1[3];
        tri_cells(g, i, c0, c1);
        for (int z = c0[2]; z <= c1[2]; ++z) for (int y = c0[1]; y <= c1[1]; ++y) for (int x = c0[0]; x <= c1[0]; ++x)
            g->start[((int64_t)z * g->n[1] + y) * g->n[0] + x + 1]++;
    }
    for (int64_t c = 0; c < cells; ++c) g->start[c + 1] += g->start[c];
    g->items = (int32_t*)malloc(sizeof(int32_t) * (size_t)(g->start[cells] > 0 ? g->start[cells] : 1));
    int64_t* fill = (int64_t*)malloc(sizeof(int64_t) * (size_t)cells);
    memcpy(fill, g->start, sizeof(int64_t) * (size_t)cells);
    for (int64_t i = 0; i < ntris; ++i)
    {
        int c0[3], c1[3];
        tri_cells(g, i, c0, c1);
        for (int z = c0[2]; z <= c1[2]; ++z) for (int y = c0[1]; y <= c1[1]; ++y) for (int x = c0[0]; x <= c1[0]; ++x)
            g->items[fill[((int64_t)z * g->n[1] + y) * g->n[0] + x]++] = (int32_t)i;
    }
    free(fill);
    return g;
}

void f64grid_free(F64Grid* g)
{
    if (!g) return;
    free(g->start); free(g->items); free(g);
}

/* same contract as brute_intersect_f64 */
int f64grid_intersect(const F64Grid* g, const float ro[3], const float rd[3], float tnear, float tfar, CpuHit* hit)
{
    hit->prim = -1; hit->t = tfar; hit->u = hit->v = 0.0f;
    double bestT = (double)tfar, second = DBL_MAX;
    int found = 0;
    /* clip the ray to the grid's box */
    double t0 = (double)tnear, t1 = (double)tfar;
    for (int k = 0; k < 3; ++k)
    {
        const double lo = g->lo[k], hi = g->lo[k] + g->cell * g->n[k];
        if (rd[k] == 0.0f) { if (ro[k] < lo || ro[k] > hi) return 0; continue; }
        double a = (lo - ro[k]) / rd[k], b = (hi - ro[k]) / rd[k];
        if (a > b) { double s = a; a = b; b = s; }
        if (a > t0) t0 = a;
        if (b < t1) t1 = b;
    }
    if (t0 > t1) return 0;
    int c[3], step[3]; double tMax[3], tDelta[3];
    for (int k = 0; k < 3; ++k)
    {
        const double p = (double)ro[k] + t0 * (double)rd[k];
        int i = (int)floor((p - g->lo[k]) / g->cell);
        c[k] = i < 0 ? 0 : (i >= g->n[k] ? g->n[k] - 1 : i);
        if (rd[k] > 0.0f) { step[k] = 1; tDelta[k] = g->cell / rd[k]; tMax[k] = (g->lo[k] + (c[k] + 1) * g->cell - ro[k]) / rd[k]; }
        else if (rd[k] < 0.0f) { step[k] = -1; tDelta[k] = -g->cell / rd[k]; tMax[k] = (g->lo[k] + c[k] * g->cell - ro[k]) / rd[k]; }
        else { step[k] = 0; tDelta[k] = DBL_MAX; tMax[k] = DBL_MAX; }
    }
    double tEnter = t0;
    for (;;)
    {
        if (tEnter > bestT + 1e-6 * fmax(1.0, fabs(bestT))) break;     /* nothing closer (or tying) can start here */
        const int64_t cellIdx = ((int64_t)c[2] * g->n[1] + c[1]) * g->n[0] + c[0];
        for (int64_t j = g->start[cellIdx]; j < g->start[cellIdx + 1]; ++j)
        {
            const int32_t i = g->items[j];
            if (found && i == hit->prim) continue;
            const float* a = g->verts + (size_t)(3 * (int64_t)i + 0) * g->stride;
            const float* b = g->verts + (size_t)(3 * (int64_t)i + 1) * g->stride;
            const float* cc = g->verts + (size_t)(3 * (int64_t)i + 2) * g->stride;
            double t, u, v;
            if (!mt_f64(a, b, cc, ro, rd, (double)tnear, (double)tfar, &t, &u, &v)) continue;
            if (!found || t < bestT || (t == bestT && i < hit->prim))
            {
                if (found) second = bestT;
                bestT = t; found = 1;
                hit->t = (float)t; hit->u = (float)u; hit->v = (float)v; hit->prim = i;
                double e1x = (double)b[0] - a[0], e1y = (double)b[1] - a[1], e1z = (double)b[2] - a[2];
                double e2x = (double)cc[0] - a[0], e2y = (double)cc[1] - a[1], e2z = (double)cc[2] - a[2];
                hit->ng[0] = (float)(e1y * e2z - e1z * e2y);
                hit->ng[1] = (float)(e1z * e2x - e1x * e2z);
                hit->ng[2] = (float)(e1x * e2y - e1y * e2x);
            }
            else if (t < second) second = t;
        }
        /* next cell */
        int k = (tMax[0] <= tMax[1]) ? ((tMax[0] <= tMax[2]) ? 0 : 2) : ((tMax[1] <= tMax[2]) ? 1 : 2);
        if (tMax[k] > t1 || step[k] == 0) break;
        tEnter = tMax[k];
        c[k] += step[k];
        if (c[k] < 0 || c[k] >= g->n[k]) break;
        tMax[k] += tDelta[k];
    }
    if (!found) return 0;
    if (second < DBL_MAX && (second - bestT) <= 1e-6 * fmax(1.0, fabs(bestT))) return 2;
    return 1;
}

typedef struct TruthArg
{
    const F64Grid* grid; const CpuBvh* bvh;
    const float* verts; int stride; int64_t ntris;
    const float* ro4; const float* rd4; float tnear; int64_t begin, end;
    CpuHit* hits; int32_t* flags; uint8_t* occ;
    uint64_t nodes, tris;       /* wide-node / triangle fetches of this thread's occlusion rays */
} TruthArg;

static void* truth_main(void* p)
{
    TruthArg* a = (TruthArg*)p;
    cpubvh_counters_reset();
    for (int64_t i = a->begin; i < a->end; ++i)
    {
        const float* ro = a->ro4 + 4 * i; const float* rd = a->rd4 + 4 * i;
        if (a->occ) a->occ[i] = (uint8_t)(cpubvh_occluded(a->bvh, ro, rd, a->tnear, rd[3]) ? 1 : 0);
        else if (a->grid) a->flags[i] = f64grid_intersect(a->grid, ro, rd, a->tnear, rd[3], &a->hits[i]);
        else a->flags[i] = brute_intersect_f64(a->verts, a->stride, a->ntris, ro, rd, a->tnear, rd[3], &a->hits[i]);
    }
    uint64_t rays;
    cpubvh_counters_get(&a->nodes, &a->tris, &rays);
    return NULL;
}

static uint64_t g_truthNodes, g_truthTris;     /* summed over the threads of the last batch */

static void truth_run(TruthArg proto, int64_t n, int threads)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256]; TruthArg args[256];
    for (int t = 0; t < threads; ++t)
    {
        args[t] = proto;
        args[t].begin = n * t / threads; args[t].end = n * (t + 1) / threads;
        pthread_create(&th[t], NULL, truth_main, &args[t]);
    }
    g_truthNodes = g_truthTris = 0;
    for (int t = 0; t < threads; ++t)
    {
        pthread_join(th[t], NULL);
        g_truthNodes += args[t].nodes; g_truthTris += args[t].tris;
    }
}

/* rd4[i].w = tFar; flags[i] = 0 miss, 1 hit, 2 hit with a tie */
void brute_intersect_f64_batch(const float* verts, int stride, int64_t ntris, const float* ro4, const float* rd4,
                               float tnear, int64_t n, int threads, CpuHit* hits, int32_t* flags)
{
    TruthArg a; memset(&a, 0, sizeof(a));
    a.verts = verts; a.stride = stride; a.ntris = ntris; a.ro4 = ro4; a.rd4 = rd4; a.tnear = tnear; a.hits = hits; a.flags = flags;
    truth_run(a, n, threads);
}

void f64grid_intersect_batch(const F64Grid* g, const float* ro4, const float* rd4, float tnear, int64_t n, int threads,
                             CpuHit* hits, int32_t* flags)
{
    TruthArg a; memset(&a, 0, sizeof(a));
    a.grid = g; a.ro4 = ro4; a.rd4 = rd4; a.tnear = tnear; a.hits = hits; a.flags = flags;
    truth_run(a, n, threads);
}

/* cpubvh_occluded (what the shim's rtcOccluded16 calls per lane) over a batch; rd4[i].w = tFar */
void cpubvh_occluded_batch(const CpuBvh* bvh, const float* ro4, const float* rd4, float tnear, int64_t n, int threads,
                           uint8_t* occluded)
{
    TruthArg a; memset(&a, 0, sizeof(a));
    a.bvh = bvh; a.ro4 = ro4; a.rd4 = rd4; a.tnear = tnear; a.occ = occluded;
    truth_run(a, n, threads);
}
void cpubvh_last_batch_counts(uint64_t* nodes, uint64_t* tris) { *nodes = g_truthNodes; *tris = g_truthTris; }
