/* TEST INFRASTRUCTURE — compiles the reference's src/rendering/lightmap.c where it lies and
 * adds the harness entry points around LmPack_Pack (CPU pre-step, shared input of both sides)
 * and LmPack_Bake (the path under test, lightmap.c:1117-1223). */
#include "rendering/lightmap.c"

#include "ref_harness.h"

int oracle_lm_pack(int atlasSize, float texelsPerUnit, float distThresh, float degThresh)
{
    LmPack* dst = LmPack_Get();
    LmPack_Del(dst);
    oracle_begin(7, 0);
    *dst = LmPack_Pack(atlasSize, texelsPerUnit, distThresh, degThresh);
    oracle_frame_end();
    return dst->lmCount;
}

int oracle_lm_size(void) { return LmPack_Get()->lmSize; }

/* pointers into page i (layout of Lightmap_New, lightmap.c:81-123) */
void oracle_lm_page(int i, float** probes5, float** position3, float** normal3, float** sampleCounts)
{
    Lightmap* lm = &LmPack_Get()->lightmaps[i];
    for (int k = 0; k < kGiDirections; ++k) probes5[k] = (float*)lm->probes[k];
    *position3 = (float*)lm->position;
    *normal3 = (float*)lm->normal;
    *sampleCounts = lm->sampleCounts;
}

/* EmbedAttributes (lightmap.c:1036-1047) again on the current pack: the step LmPack_EmbedDesc replaces, alone */
void oracle_lm_embed(void)
{
    LmPack* pack = LmPack_Get();
    oracle_begin(ORACLE_STREAM_LMPACK, 1);
    EmbedAttributes(pack->lightmaps, pack->lmCount, pack->texelsPerMeter);
    oracle_frame_end();
}

int oracle_lm_slot(int i) { return (int)LmPack_Get()->lightmaps[i].slot.index; }

void oracle_lm_bake(void* scene, float timeSlice, int spp, uint32_t pass)
{
    oracle_begin(ORACLE_STREAM_BAKE, pass);
    LmPack_Bake((PtScene*)scene, timeSlice, spp);
    oracle_frame_end();
}
