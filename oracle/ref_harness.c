/* TEST INFRASTRUCTURE — scene generator and metric of the reference's console
 * commands, restated on the reference's own math headers:
 *   cornell_box  src/rendering/render_system.c:745-1306
 *   pt_stddev    src/rendering/render_system.c:1374-1394 */
#include "common/macro.h"
#include "math/types.h"
#include "math/float4_funcs.h"
#include "math/float4x4_funcs.h"
#include "math/quat_funcs.h"
#include "math/color.h"
#include "rendering/material.h"
#include "rendering/cubemap.h"
#include "rendering/mesh.h"
#include "rendering/texture.h"
#include "rendering/drawable.h"
#include <string.h>
#include <stdlib.h>
#include <stdio.h>

#include "ref_harness.h"

static int s_boxMesh, s_sphereMesh;

static int gen_box_mesh(void)
{
    static const float4 v[] =
    {
        {  1.0f,  1.0f, -1.0f }, {  1.0f, -1.0f, -1.0f }, {  1.0f,  1.0f,  1.0f }, {  1.0f, -1.0f,  1.0f },
        { -1.0f,  1.0f, -1.0f }, { -1.0f, -1.0f, -1.0f }, { -1.0f,  1.0f,  1.0f }, { -1.0f, -1.0f,  1.0f },
    };
    static const float2 vt[] =
    {
        { 0.875f, 0.5f }, { 0.625f, 0.75f }, { 0.625f, 0.5f }, { 0.375f, 1.0f }, { 0.375f, 0.75f },
        { 0.625f, 0.0f }, { 0.375f, 0.25f }, { 0.375f, 0.0f }, { 0.375f, 0.5f }, { 0.125f, 0.75f },
        { 0.125f, 0.5f }, { 0.625f, 0.25f }, { 0.875f, 0.75f }, { 0.625f, 1.0f },
    };
    static const float4 vn[] =
    {
        { 0.0f, 1.0f, 0.0f }, { 0.0f, 0.0f, 1.0f }, { -1.0f, 0.0f, 0.0f },
        { 0.0f, -1.0f, 0.0f }, { 1.0f, 0.0f, 0.0f }, { 0.0f, 0.0f, -1.0f },
    };
    static const int3 f[] =
    {
        { 5, 1, 1 }, { 3, 2, 1 }, { 1, 3, 1 },   { 3, 2, 2 }, { 8, 4, 2 }, { 4, 5, 2 },
        { 7, 6, 3 }, { 6, 7, 3 }, { 8, 8, 3 },   { 2, 9, 4 }, { 8, 10, 4 }, { 6, 11, 4 },
        { 1, 3, 5 }, { 4, 5, 5 }, { 2, 9, 5 },   { 5, 12, 6 }, { 2, 9, 6 }, { 6, 7, 6 },
        { 5, 1, 1 }, { 7, 13, 1 }, { 3, 2, 1 },  { 3, 2, 2 }, { 7, 14, 2 }, { 8, 4, 2 },
        { 7, 6, 3 }, { 5, 12, 3 }, { 6, 7, 3 },  { 2, 9, 4 }, { 4, 5, 4 }, { 8, 10, 4 },
        { 1, 3, 5 }, { 3, 2, 5 }, { 4, 5, 5 },   { 5, 12, 6 }, { 1, 3, 6 }, { 2, 9, 6 },
    };
    enum { N = 36 };
    float4 pos[N], nrm[N], uv[N];
    for (int i = 0; i < N; ++i)
    {
        pos[i] = f4_mulvs(v[f[i].x - 1], 0.5f);
        uv[i] = f4_v(vt[f[i].y - 1].x, vt[f[i].y - 1].y, 0.0f, 0.0f);
        nrm[i] = vn[f[i].z - 1];
    }
    return oracle_add_mesh((const float*)pos, (const float*)nrm, (const float*)uv, N);
}

static int gen_sphere_mesh(void)
{
    const i32 vsteps = 24, hsteps = 48;
    const float dv = kPi / vsteps, dh = kTau / hsteps;
    const i32 maxlen = 6 * vsteps * hsteps;
    float4* P = malloc(sizeof(float4) * maxlen);
    float4* U = malloc(sizeof(float4) * maxlen);
    i32 len = 0;
    for (i32 v = 0; v < vsteps; ++v)
    {
        float theta1 = v * dv, theta2 = (v + 1) * dv;
        float st1 = sinf(theta1), ct1 = cosf(theta1), st2 = sinf(theta2), ct2 = cosf(theta2);
        for (i32 h = 0; h < hsteps; ++h)
        {
            float phi1 = h * dh, phi2 = (h + 1) * dh;
            float sp1 = sinf(phi1), cp1 = cosf(phi1), sp2 = sinf(phi2), cp2 = cosf(phi2);
            float4 u1 = f4_v(phi1 / kTau, 1.0f - theta1 / kPi, 0.0f, 0.0f);
            float4 u2 = f4_v(phi2 / kTau, 1.0f - theta1 / kPi, 0.0f, 0.0f);
            float4 u3 = f4_v(phi2 / kTau, 1.0f - theta2 / kPi, 0.0f, 0.0f);
            float4 u4 = f4_v(phi1 / kTau, 1.0f - theta2 / kPi, 0.0f, 0.0f);
            float4 n1 = f4_v(st1 * cp1, ct1, st1 * sp1, 0.0f);
            float4 n2 = f4_v(st1 * cp2, ct1, st1 * sp2, 0.0f);
            float4 n3 = f4_v(st2 * cp2, ct2, st2 * sp2, 0.0f);
            float4 n4 = f4_v(st2 * cp1, ct2, st2 * sp1, 0.0f);
            if (v == 0)
            {
                P[len] = n1; U[len++] = u1; P[len] = n3; U[len++] = u3; P[len] = n4; U[len++] = u4;
            }
            else if (v + 1 == vsteps)
            {
                P[len] = n3; U[len++] = u3; P[len] = n1; U[len++] = u1; P[len] = n2; U[len++] = u2;
            }
            else
            {
                P[len] = n1; U[len++] = u1; P[len] = n2; U[len++] = u2; P[len] = n4; U[len++] = u4;
                P[len] = n2; U[len++] = u2; P[len] = n3; U[len++] = u3; P[len] = n4; U[len++] = u4;
            }
        }
    }
    int id = oracle_add_mesh((const float*)P, (const float*)P, (const float*)U, len);
    free(P); free(U);
    return id;
}

static int flat_texture(float4 value)
{
    R8G8B8A8_t texel = GammaEncode_rgba8(value);
    return oracle_add_texture(1, 1, &texel, 4, VK_FORMAT_R8G8B8A8_SRGB);
}

static int create(int mesh, float4 t, quat r, float4 s, float4 albedo, float4 rome, u32 flags, float ior)
{
    float4x4 M = f4x4_trs(t, r, s);
    albedo = Color_SDRToScene(albedo);
    int ta = flat_texture(albedo);
    int tr = flat_texture(rome);
    if (rome.w > 0.0f) flags |= MatFlag_Emissive;
    float4 mfp = f4_0;
    return oracle_add_entity(mesh, (const float*)&M, ta, tr, 0, flags, (const float*)&mfp, ior, 0.0f);
}

void oracle_cornell_box(int spheres)
{
    oracle_tables_clear();
    s_boxMesh = gen_box_mesh();
    s_sphereMesh = spheres ? gen_sphere_mesh() : 0;

    const float wallExtents = 5.0f;
    const float4 wallScale = f4_v(2.0f * wallExtents, 2.0f * wallExtents, kDeci, 0.0f);
    const float lightScale = 1.0f;
    const float4 right = Cubemap_kForwards[Cubeface_XP], right_up = Cubemap_kUps[Cubeface_XP];
    const float4 left = Cubemap_kForwards[Cubeface_XM], left_up = Cubemap_kUps[Cubeface_XM];
    const float4 up = Cubemap_kForwards[Cubeface_YP], up_up = Cubemap_kUps[Cubeface_YP];
    const float4 down = Cubemap_kForwards[Cubeface_YM], down_up = Cubemap_kUps[Cubeface_YM];
    const float4 fwd = Cubemap_kForwards[Cubeface_ZM], fwd_up = Cubemap_kUps[Cubeface_ZM];
    const float4 back = Cubemap_kForwards[Cubeface_ZP], back_up = Cubemap_kUps[Cubeface_ZP];
    const float cHi = 0.9f, cLo = 1.0f - cHi;
    const float4 red = { cHi, cLo, cLo, 1.0f }, green = { cLo, cHi, cLo, 1.0f };
    const float4 blue = { cLo, cLo, cHi, 1.0f }, white = { cHi, cHi, cHi, 1.0f };
    const float4 plastic = { 0.9f, 1.0f, 0.0f, 0.0f }, metal = { 0.1f, 1.0f, 1.0f, 0.0f };
    const float4 light = { 0.9f, 1.0f, 0.0f, 1.0f };
    const int B = s_boxMesh;

    create(B, f4_mulvs(down, wallExtents), quat_lookat(up, up_up), wallScale, white, plastic, 0, 1.0f);
    create(B, f4_mulvs(up, wallExtents), quat_lookat(down, down_up), wallScale, white, plastic, 0, 1.0f);
    create(B, f4_mulvs(up, wallExtents - kDeci * 2.0f), quat_lookat(down, down_up),
           f4_v(lightScale, lightScale, kDeci, 0.0f), f4_s(1.0f), light, 0, 1.0f);
    create(B, f4_mulvs(left, wallExtents), quat_lookat(right, right_up), wallScale, green, plastic, 0, 1.0f);
    create(B, f4_mulvs(right, wallExtents), quat_lookat(left, left_up), wallScale, red, plastic, 0, 1.0f);
    create(B, f4_mulvs(back, wallExtents), quat_lookat(back, back_up), wallScale, white, plastic, 0, 1.0f);
    create(B, f4_mulvs(fwd, wallExtents), quat_lookat(fwd, fwd_up), wallScale, blue, plastic, 0, 1.0f);

    if (spheres)
    {
        const float sphereScale = 0.75f, margin = sphereScale * 1.5f;
        const float lo = -wallExtents + margin, hi = wallExtents - margin;
        const u32 kMatFlags[] = { 0x0, 0x0, MatFlag_Refractive };
        const float kMetallics[] = { 1.0f, 0.0f, 0.0f };
        const float kIors[] = { 1.0f, 1.0f, 1.5f };
        for (i32 iRow = 0; iRow < 3; ++iRow)
        {
            const float z = f1_lerp(lo, hi, (iRow + 0.5f) / 3);
            for (i32 iCol = 0; iCol < 5; ++iCol)
            {
                float tCol = (iCol + 0.5f) / 5;
                float roughness = f1_lerp(0.0f, 1.0f, tCol);
                float x = f1_lerp(lo, hi, tCol);
                create(s_sphereMesh, f4_v(x, lo, z, 0.0f), quat_id, f4_s(sphereScale), white,
                       f4_v(roughness, 1.0f, kMetallics[iRow], 0.0f), kMatFlags[iRow], kIors[iRow]);
            }
        }
    }
    else
    {
        const float boxScale = 2.0f, margin = boxScale * 0.5f;
        const float lo = -wallExtents + margin, hi = wallExtents - margin;
        create(B, f4_v(f1_lerp(lo, hi, 0.2f), -wallExtents + boxScale, f1_lerp(lo, hi, 0.2f), 0.0f),
               quat_lookat(f4_normalize3(f4_v(0.2f, 0.0f, 1.0f, 0.0f)), up),
               f4_v(boxScale, boxScale * 2.0f, boxScale, 0.0f), white, metal, 0, 1.0f);
        create(B, f4_v(f1_lerp(lo, hi, 0.8f), -wallExtents + boxScale * 0.5f, f1_lerp(lo, hi, 0.8f), 0.0f),
               quat_lookat(f4_normalize3(f4_v(-0.2f, 0.0f, 1.0f, 0.0f)), up),
               f4_v(boxScale, boxScale, boxScale, 0.0f), white, plastic, 0, 1.0f);
    }
}

int oracle_entity_count(void) { return Entities_Get()->count; }

void oracle_entity_get(int i, int* meshId, float* matrix16, int* texIds3,
                       uint32_t* flags, float* meanFreePath4, float* ior, float* bumpiness)
{
    const Entities* e = Entities_Get();
    *meshId = (int)e->meshes[i].index;
    memcpy(matrix16, &e->matrices[i], sizeof(float4x4));
    const Material* m = &e->materials[i];
    texIds3[0] = (int)m->albedo.index; texIds3[1] = (int)m->rome.index; texIds3[2] = (int)m->normal.index;
    *flags = m->flags;
    memcpy(meanFreePath4, &m->meanFreePath, sizeof(float4));
    *ior = m->ior; *bumpiness = m->bumpiness;
}

int oracle_mesh_get(int meshId, const float** pos4, const float** nrm4, const float** uv4)
{
    MeshId id = { 0 }; id.version = 1; id.index = (u32)meshId;
    const Mesh* m = Mesh_Get(id);
    if (!m) return 0;
    *pos4 = (const float*)m->positions; *nrm4 = (const float*)m->normals; *uv4 = (const float*)m->uvs;
    return m->length;
}

/* int4 per vertex; .w = lightmap slot index after LmPack_Pack (lightmap.c:892-921) */
const int32_t* oracle_mesh_texindices(int meshId)
{
    MeshId id = { 0 }; id.version = 1; id.index = (u32)meshId;
    const Mesh* m = Mesh_Get(id);
    return m ? (const int32_t*)m->texIndices : NULL;
}

int oracle_texture_get(int texId, int* w, int* h, const void** texels, int* vkFormat)
{
    TextureId id = { 0 }; id.version = 1; id.index = (u32)texId;
    const Texture* t = Texture_Get(id);
    if (!t) return 0;
    *w = t->size.x; *h = t->size.y; *texels = t->texels; *vkFormat = (int)t->format;
    return 1;
}

float oracle_stddev(const float* color3, int w, int h)
{
    const i32 len = w * h;
    const float meanWeight = 1.0f / len;
    const float varianceWeight = 1.0f / (len - 1);
    float mean = 0.0f;
    for (i32 i = 0; i < len; ++i)
    {
        float lum = (color3[3 * i] + color3[3 * i + 1] + color3[3 * i + 2]) * (1.0f / 3.0f);
        mean += lum * meanWeight;
    }
    float variance = 0.0f;
    for (i32 i = 0; i < len; ++i)
    {
        float lum = (color3[3 * i] + color3[3 * i + 1] + color3[3 * i + 2]) * (1.0f / 3.0f);
        float err = lum - mean;
        variance += varianceWeight * (err * err);
    }
    return sqrtf(variance);
}
