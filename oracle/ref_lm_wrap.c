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

/* ---- capture of the packer's intermediate state (tests of LmPack_ChartMasksDesc / LmPack_AtlasDesc) ---- */
void oracle_set_task_hook(void (*hook)(void*, TaskExecuteFn, i32, int));

typedef struct CapChart
{
    i32 nodeCount;
    float* triBefore;      /* [nodeCount][3][2] as chart_group hands them to ChartMaskFn */
    float* triAfter;       /* translated by ChartMaskFn */
    float area; int2 maskSize; u8* mask;
    i32 atlasIndex; int2 translation;
    const chartnode_t* key; /* nodes pointer: identifies the chart across chart_sort */
    i32 sortedIndex;       /* position in AtlasFn's order */
} CapChart;
static CapChart* s_cap; static i32 s_capCount; static i32 s_capAtlasCount;

static void cap_free(void)
{
    for (i32 i = 0; i < s_capCount; ++i) { free(s_cap[i].triBefore); free(s_cap[i].triAfter); free(s_cap[i].mask); }
    free(s_cap); s_cap = NULL; s_capCount = 0;
}
static float* cap_tris(const chart_t* ch)
{
    float* t = (float*)malloc(sizeof(float) * 6 * (size_t)(ch->nodeCount > 0 ? ch->nodeCount : 1));
    for (i32 n = 0; n < ch->nodeCount; ++n)
    {
        const Tri2D tri = ch->nodes[n].triCoord;
        t[6 * n + 0] = tri.a.x; t[6 * n + 1] = tri.a.y; t[6 * n + 2] = tri.b.x; t[6 * n + 3] = tri.b.y;
        t[6 * n + 4] = tri.c.x; t[6 * n + 5] = tri.c.y;
    }
    return t;
}
static void cap_hook(void* task, TaskExecuteFn fn, i32 worksize, int phase)
{
    if (fn == ChartMaskFn)
    {
        chartmask_t* t = (chartmask_t*)task;
        if (phase == 0)
        {
            cap_free();
            s_capCount = t->chartCount;
            s_cap = (CapChart*)calloc((size_t)(s_capCount > 0 ? s_capCount : 1), sizeof(CapChart));
            for (i32 i = 0; i < s_capCount; ++i)
            {
                s_cap[i].nodeCount = t->charts[i].nodeCount;
                s_cap[i].triBefore = cap_tris(&t->charts[i]);
                s_cap[i].key = t->charts[i].nodes;
            }
        }
        else
        {
            for (i32 i = 0; i < s_capCount; ++i)
            {
                const chart_t* ch = &t->charts[i];
                s_cap[i].triAfter = cap_tris(ch);
                s_cap[i].area = ch->area; s_cap[i].maskSize = ch->mask.size;
                const size_t len = (size_t)ch->mask.size.x * (size_t)ch->mask.size.y;
                s_cap[i].mask = (u8*)malloc(len ? len : 1);
                memcpy(s_cap[i].mask, ch->mask.ptr, len);
            }
        }
    }
    else if (fn == AtlasFn)
    {
        atlastask_t* t = (atlastask_t*)task;
        for (i32 j = 0; j < t->chartCount; ++j)
        {
            for (i32 i = 0; i < s_capCount; ++i)
            {
                if (s_cap[i].key != t->charts[j].nodes) continue;
                if (phase == 0) s_cap[i].sortedIndex = j;
                else { s_cap[i].atlasIndex = t->charts[j].atlasIndex; s_cap[i].translation = t->charts[j].translation; }
                break;
            }
        }
        if (phase == 0) s_capAtlasCount = t->atlasCount;
    }
}

int oracle_lm_pack_capture(int atlasSize, float texelsPerUnit, float distThresh, float degThresh)
{
    const int threads = oracle_get_threads(), mode = oracle_get_task_mode();
    oracle_set_threads(1);
    oracle_set_task_mode(1);
    oracle_set_task_hook(cap_hook);
    oracle_lm_pack(atlasSize, texelsPerUnit, distThresh, degThresh);
    oracle_set_task_hook(NULL);
    oracle_set_task_mode(mode);
    oracle_set_threads(threads);
    return s_capCount;
}
int oracle_lm_cap_atlas_count(void) { return s_capAtlasCount; }
int oracle_lm_cap_chart(int c, int* nodeCount, float* area, int* maskSize2, int* atlasIndex, int* translation2)
{
    if (c < 0 || c >= s_capCount) return -1;
    const CapChart* ch = &s_cap[c];
    *nodeCount = ch->nodeCount; *area = ch->area; maskSize2[0] = ch->maskSize.x; maskSize2[1] = ch->maskSize.y;
    *atlasIndex = ch->atlasIndex; translation2[0] = ch->translation.x; translation2[1] = ch->translation.y;
    return 0;
}
void oracle_lm_cap_chart_data(int c, float* triBefore, float* triAfter, unsigned char* mask)
{
    const CapChart* ch = &s_cap[c];
    if (triBefore) memcpy(triBefore, ch->triBefore, sizeof(float) * 6 * (size_t)ch->nodeCount);
    if (triAfter) memcpy(triAfter, ch->triAfter, sizeof(float) * 6 * (size_t)ch->nodeCount);
    if (mask) memcpy(mask, ch->mask, (size_t)ch->maskSize.x * (size_t)ch->maskSize.y);
}
int oracle_lm_cap_sorted_index(int c) { return s_cap[c].sortedIndex; }
