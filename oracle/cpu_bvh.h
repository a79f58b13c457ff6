/* TEST INFRASTRUCTURE — not part of the shipped library.
 *
 * CPU restatement of the ray-query backend the reference links against:
 * Embree 3.13.5 (prebuilt libembree3.so.3, pinned by the reference at
 * submodules/embree/embree-3.13.5.x86_64.linux; the binary is absent from the
 * mount, only headers and man pages exist). The reference's call sites are
 * src/rendering/path_tracer.c:484 (rtcIntersect1), :548 (rtcOccluded16),
 * :614 (rtcPointQuery) and :618-690 (scene build, one triangle geometry with
 * an identity index buffer, RTC_BUILD_QUALITY_HIGH).
 *
 * Semantics follow the Embree 3 manual (share/man/man3):
 *   - triangle hit point  = (1-u-v)*v0 + u*v1 + v*v2
 *   - Ng = (v1-v0) x (v2-v0), unnormalised, front face = counter-clockwise
 *   - two-sided test, closest t in [tnear, tfar]
 *   - occlusion: any hit in [tnear, tfar]
 *   - point query: callback per candidate primitive, radius may shrink
 *
 * The BVH is the "canonical reference BVH" of SURVEY.md §8(d): binned SAH with
 * 16 bins, leaves <= 4 triangles, collapsed 8-wide, children visited front to
 * back. It also counts wide-node and triangle fetches per ray; those counts
 * are the N_node / N_tri of the bytes-per-ray roofline.
 *
 * parity: no Embree golden vectors exist in the reference (it has no tests);
 * this restatement is pinned against the fp64 brute-force intersector below.
 */
#ifndef ORACLE_CPU_BVH_H
#define ORACLE_CPU_BVH_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CpuBvh CpuBvh;

typedef struct CpuHit
{
    float t, u, v;
    int32_t prim;       /* -1 on miss */
    float ng[3];        /* unnormalised geometric normal */
} CpuHit;

/* verts: 3*ntris vertices, `stride` floats apart (3 for float3, 4 for float4). */
CpuBvh* cpubvh_build(const float* verts, int stride, int64_t ntris);
void cpubvh_free(CpuBvh* bvh);

int cpubvh_intersect(const CpuBvh* bvh, const float ro[3], const float rd[3],
                     float tnear, float tfar, CpuHit* hit);
int cpubvh_occluded(const CpuBvh* bvh, const float ro[3], const float rd[3],
                    float tnear, float tfar);

/* Calls fn(user, prim, &radius) for every triangle whose leaf box is within
 * `radius` of p; fn may shrink *radius and returns nonzero when it did. */
typedef int (*CpuPointQueryFn)(void* user, uint32_t prim, float* radius);
void cpubvh_point_query(const CpuBvh* bvh, const float p[3], float radius,
                        CpuPointQueryFn fn, void* user);

/* per-thread fetch counters (wide nodes visited, triangles tested) */
void cpubvh_counters_reset(void);
void cpubvh_counters_get(uint64_t* nodes, uint64_t* tris, uint64_t* rays);

int64_t cpubvh_node_count(const CpuBvh* bvh);

/* Ground truth independent of any BVH: all-triangle loop, Moller-Trumbore in
 * fp64. Ties within 1e-9 relative t are flagged (returns 2) so parity tests
 * can exempt exact-edge cases. Returns 0 on miss. */
int brute_intersect_f64(const float* verts, int stride, int64_t ntris,
                        const float ro[3], const float rd[3],
                        float tnear, float tfar, CpuHit* hit);

/* closest hit over n rays on `threads` host threads (ro4/rd4: float4 per ray, rd4.w = tFar);
 * prim[i] = primitive or -1; returns summed wide-node and triangle fetches */
void cpubvh_intersect_batch(const CpuBvh* bvh, const float* ro4, const float* rd4, float tnear, int64_t n,
                            int threads, int32_t* prim, uint64_t* nodes, uint64_t* tris);


/* threaded batches of the two ground truths and of the occlusion query; rd4[i].w = tFar,
 * flags[i] = 0 miss, 1 hit, 2 hit whose runner-up lies within 1e-6 relative (tie, exempt) */
void brute_intersect_f64_batch(const float* verts, int stride, int64_t ntris, const float* ro4, const float* rd4,
                               float tnear, int64_t n, int threads, CpuHit* hits, int32_t* flags);
/* fp64 ground truth over a uniform grid instead of all triangles (no BVH involved): see cpu_bvh.c */
typedef struct F64Grid F64Grid;
F64Grid* f64grid_build(const float* verts, int stride, int64_t ntris, double cell);
void f64grid_free(F64Grid* g);
int f64grid_intersect(const F64Grid* g, const float ro[3], const float rd[3], float tnear, float tfar, CpuHit* hit);
void f64grid_intersect_batch(const F64Grid* g, const float* ro4, const float* rd4, float tnear, int64_t n, int threads,
                             CpuHit* hits, int32_t* flags);
void cpubvh_occluded_batch(const CpuBvh* bvh, const float* ro4, const float* rd4, float tnear, int64_t n, int threads,
                           uint8_t* occluded);
void cpubvh_last_batch_counts(uint64_t* nodes, uint64_t* tris);   /* fetches of the last occlusion batch */

#ifdef __cplusplus
}
#endif
#endif
