// lmpack.cu — the two data-parallel steps of LmPack_Pack ahead of the attribute embedding
// (src/rendering/lightmap.c):
//
//   chart masks   ChartMaskFn / mask_tri (:225-248, :527-561): every chart is moved to (2, 2) and rasterised
//                 into a mask of its own — a texel is set when any of the chart's triangles lies within
//                 kMaskPadding = 1 texel of its centre. texels x triangles per chart.
//   atlas         AtlasFn / atlas_search / mask_find / mask_fits / mask_write (:172-222, :262-282, :680-721,
//                 :779-823): charts, in the order given, take the first position (rows, then columns) of the
//                 first atlas where their mask meets no set texel.
//
// Masks are bit rows here (32 texels per word): a fit test is a few shifted ANDs per mask row, one thread per
// candidate position, all positions of an atlas in one launch, the first fitting one in scan order by atomicMin.
// The charts themselves are placed one after the other (each placement changes the atlas), as the reference's
// single task chunk does; its row-skipping search state (prevAtlas, prevRow, prevArea) is kept, so placements
// are the reference's with one task thread. What stays host code: plane clustering and the k-means split of
// chart_group (sequential, random), chart_sort, chartnodes_assign.
#include "pt_device.h"

#include <algorithm>
#include <vector>

namespace pt {

namespace {

// sdTriangle2D (src/math/sdf.h:125-156) with every product, sum and quotient rounded on its own, as the
// reference's CPU build computes it (no fused multiply-add on baseline x86-64): the texels on a chart's rim
// fall on the same side of the <= 1 test.
__device__ __forceinline__ float dot2r(float ax, float ay, float bx, float by) { return addr(mulr(ax, bx), mulr(ay, by)); }
__device__ __forceinline__ float satr(float x) { return x > 0.0f ? (x < 1.0f ? x : 1.0f) : 0.0f; }      // f1_saturate: NaN -> 0
__device__ float sdTriangle2Dr(V2 a, V2 b, V2 c, V2 p)
{
    const float bax = subr(b.x, a.x), bay = subr(b.y, a.y), cbx = subr(c.x, b.x), cby = subr(c.y, b.y);
    const float acx = subr(a.x, c.x), acy = subr(a.y, c.y);
    const float pax = subr(p.x, a.x), pay = subr(p.y, a.y), pbx = subr(p.x, b.x), pby = subr(p.y, b.y);
    const float pcx = subr(p.x, c.x), pcy = subr(p.y, c.y);
    const float h0 = satr(divr(dot2r(pax, pay, bax, bay), dot2r(bax, bay, bax, bay)));
    const float h1 = satr(divr(dot2r(pbx, pby, cbx, cby), dot2r(cbx, cby, cbx, cby)));
    const float h2 = satr(divr(dot2r(pcx, pcy, acx, acy), dot2r(acx, acy, acx, acy)));
    const float q0x = subr(pax, mulr(bax, h0)), q0y = subr(pay, mulr(bay, h0));
    const float q1x = subr(pbx, mulr(cbx, h1)), q1y = subr(pby, mulr(cby, h1));
    const float q2x = subr(pcx, mulr(acx, h2)), q2y = subr(pcy, mulr(acy, h2));
    const float w = subr(mulr(bax, acy), mulr(bay, acx));
    const float s = (w > 0.0f) ? 1.0f : ((w < 0.0f) ? -1.0f : 0.0f);
    const float dx = fminf(dot2r(q0x, q0y, q0x, q0y), fminf(dot2r(q1x, q1y, q1x, q1y), dot2r(q2x, q2y, q2x, q2y)));
    const float dy = fminf(mulr(s, subr(mulr(pax, bay), mulr(pay, bax))),
                           fminf(mulr(s, subr(mulr(pbx, cby), mulr(pby, cbx))), mulr(s, subr(mulr(pcx, acy), mulr(pcy, acx)))));
    const float dySign = (dy >= 0.0f) ? 1.0f : -1.0f;
    return mulr(-sqrtr(fmaxf(dx, kEpsilonSq)), dySign);
}

struct ChartInfo
{
    int32_t firstNode, nodeCount;
    int32_t w, h;           // mask size in texels
    int32_t rowWords;       // 32-texel words per mask row
    int32_t wordBase;       // first word of the chart's mask
};

// one thread per mask word: 32 texels against all triangles of the chart
__global__ void __launch_bounds__(128) k_chartMask(const ChartInfo* __restrict__ charts, int32_t chartCount, const V2* __restrict__ tri,
                                                   uint32_t* __restrict__ bits, int32_t totalWords)
{
    const int32_t word = blockIdx.x * blockDim.x + threadIdx.x;
    if (word >= totalWords) return;
    int lo = 0, hi = chartCount - 1;
    while (lo < hi)
    {
        const int mid = (lo + hi + 1) >> 1;
        if (charts[mid].wordBase <= word) lo = mid; else hi = mid - 1;
    }
    const ChartInfo c = charts[lo];
    const int32_t local = word - c.wordBase;
    const int32_t y = local / c.rowWords, x0 = (local % c.rowWords) * 32;
    uint32_t out = 0u;
    for (int32_t n = 0; n < c.nodeCount && out != 0xffffffffu; ++n)
    {
        const V2 A = tri[3 * (size_t)(c.firstNode + n)], B = tri[3 * (size_t)(c.firstNode + n) + 1], C = tri[3 * (size_t)(c.firstNode + n) + 2];
        // the 32 texels lie on one row: skip the triangle when its padded box misses the span
        const float ylo = fminf(A.y, fminf(B.y, C.y)) - 1.5f, yhi = fmaxf(A.y, fmaxf(B.y, C.y)) + 1.5f;
        const float xlo = fminf(A.x, fminf(B.x, C.x)) - 1.5f, xhi = fmaxf(A.x, fmaxf(B.x, C.x)) + 1.5f;
        if (y + 0.5f < ylo || y + 0.5f > yhi || x0 + 32.0f < xlo || (float)x0 > xhi) continue;
        for (int k = 0; k < 32; ++k)
        {
            const int32_t x = x0 + k;
            if (x >= c.w || ((out >> k) & 1u)) continue;
            if (sdTriangle2Dr(A, B, C, v2(x + 0.5f, y + 0.5f)) <= 1.0f) out |= 1u << k;      // TriTest, kMaskPadding
        }
    }
    bits[word] = out;
}

__global__ void __launch_bounds__(256) k_maskBytes(const ChartInfo* __restrict__ charts, int32_t chart, const uint32_t* __restrict__ bits,
                                                   uint8_t* __restrict__ bytes)
{
    const ChartInfo c = charts[chart];
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c.w * c.h) return;
    const int32_t x = i % c.w, y = i / c.w;
    bytes[i] = ((bits[c.wordBase + y * c.rowWords + (x >> 5)] >> (x & 31)) & 1u) ? 0xff : 0x00;
}

// one thread per candidate translation (x, y), y >= y0: mask_fits over bit rows. best = min over fitting (y << 16 | x).
__global__ void __launch_bounds__(128) k_atlasFind(const uint32_t* __restrict__ atlas, int32_t atlasSize, int32_t atlasRowWords,
                                                   const uint32_t* __restrict__ mask, ChartInfo c, int32_t y0, int32_t rangeX, int32_t rangeY,
                                                   uint32_t* best)
{
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const int32_t rows = rangeY - y0;
    if (i >= rows * rangeX) return;
    const int32_t x = i % rangeX, y = y0 + i / rangeX;
    const uint32_t key = ((uint32_t)y << 16) | (uint32_t)x;
    if (key >= *best) return;                       // a fit earlier in scan order is known already
    const int32_t shift = x & 31, w0 = x >> 5;
    for (int32_t r = 0; r < c.h; ++r)
    {
        const uint32_t* arow = atlas + (size_t)(y + r) * atlasRowWords;
        const uint32_t* mrow = mask + c.wordBase + r * c.rowWords;
        for (int32_t w = 0; w < c.rowWords; ++w)
        {
            const uint32_t m = mrow[w];
            if (m == 0u) continue;
            const int32_t aw = w0 + w;
            uint32_t a = arow[aw] >> shift;
            if (shift && aw + 1 < atlasRowWords) a |= arow[aw + 1] << (32 - shift);
            if (a & m) return;
        }
    }
    atomicMin(best, key);
}

// mask_write: the chart's set texels become set texels of the atlas
__global__ void __launch_bounds__(128) k_atlasWrite(uint32_t* __restrict__ atlas, int32_t atlasRowWords, const uint32_t* __restrict__ mask,
                                                    ChartInfo c, int32_t tx, int32_t ty)
{
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c.h * c.rowWords) return;
    const int32_t r = i / c.rowWords, w = i % c.rowWords;
    const uint32_t m = mask[c.wordBase + i];
    if (m == 0u) return;
    const int32_t shift = tx & 31, aw = (tx >> 5) + w;
    uint32_t* arow = atlas + (size_t)(ty + r) * atlasRowWords;
    atomicOr(&arow[aw], m << shift);
    if (shift && aw + 1 < atlasRowWords && (m >> (32 - shift))) atomicOr(&arow[aw + 1], m >> (32 - shift));
}

} // namespace

// Host part of ChartMaskFn (lightmap.c:537-556): the chart moves so that its lower corner is at (2, 2); its
// area is that of its bounds; its mask reaches 2 texels beyond its upper corner. Exact float operations
// (min, max, one subtraction each), so the same values as the reference's.
void chartPrepare(V2* tri, const int32_t* firstNode, int32_t chartCount, float* areaOut, int32_t* maskSizeOut)
{
    const float bigNum = (float)(1 << 23);
    for (int32_t c = 0; c < chartCount; ++c)
    {
        V2* t = tri + 3 * (size_t)firstNode[c];
        const size_t n = 3 * (size_t)(firstNode[c + 1] - firstNode[c]);
        float lox = bigNum, loy = bigNum, hix = -bigNum, hiy = -bigNum;
        for (size_t i = 0; i < n; ++i) { lox = std::min(lox, t[i].x); loy = std::min(loy, t[i].y); hix = std::max(hix, t[i].x); hiy = std::max(hiy, t[i].y); }
        const float sx = lox - 2.0f, sy = loy - 2.0f;
        for (size_t i = 0; i < n; ++i) { t[i].x = t[i].x - sx; t[i].y = t[i].y - sy; }
        lox = bigNum; loy = bigNum; hix = -bigNum; hiy = -bigNum;
        for (size_t i = 0; i < n; ++i) { lox = std::min(lox, t[i].x); loy = std::min(loy, t[i].y); hix = std::max(hix, t[i].x); hiy = std::max(hiy, t[i].y); }
        areaOut[c] = (hix - lox) * (hiy - loy);
        maskSizeOut[2 * c] = (int32_t)(hix + 2.0f);
        maskSizeOut[2 * c + 1] = (int32_t)(hiy + 2.0f);
    }
}

size_t chartMaskWords(const int32_t* maskSize, int32_t chartCount)
{
    size_t words = 0;
    for (int32_t c = 0; c < chartCount; ++c) words += (size_t)((maskSize[2 * c] + 31) / 32) * (size_t)std::max(maskSize[2 * c + 1], 0);
    return words;
}

// dCharts: chartCount ChartInfo records (device, 24 bytes each: filled here); dTri: translated triangle corners on the
// device; dBits: chartMaskWords() words
bool launchChartMasks(const int32_t* firstNode, const int32_t* maskSize, int32_t chartCount, void* dCharts, const V2* dTri,
                      uint32_t* dBits, std::vector<int32_t>& wordBaseOut, cudaStream_t stream)
{
    std::vector<ChartInfo> info((size_t)chartCount);
    wordBaseOut.assign((size_t)chartCount + 1, 0);
    int32_t words = 0;
    for (int32_t c = 0; c < chartCount; ++c)
    {
        ChartInfo& ci = info[(size_t)c];
        ci.firstNode = firstNode[c]; ci.nodeCount = firstNode[c + 1] - firstNode[c];
        ci.w = maskSize[2 * c]; ci.h = std::max(maskSize[2 * c + 1], 0);
        ci.rowWords = (ci.w + 31) / 32;
        ci.wordBase = words;
        wordBaseOut[(size_t)c] = words;
        words += ci.rowWords * ci.h;
    }
    wordBaseOut[(size_t)chartCount] = words;
    if (chartCount <= 0 || words <= 0) return true;
    if (!PT_CUDA(cudaMemcpyAsync(dCharts, info.data(), sizeof(ChartInfo) * info.size(), cudaMemcpyHostToDevice, stream))) return false;
    // the copy reads `info` before this function returns only if the stream has reached it
    if (!PT_CUDA(cudaStreamSynchronize(stream))) return false;
    k_chartMask<<<(unsigned)((words + 127) / 128), 128, 0, stream>>>((const ChartInfo*)dCharts, chartCount, dTri, dBits, words);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

size_t chartInfoBytes() { return sizeof(ChartInfo); }

bool launchMaskBytes(const void* dCharts, int32_t chart, int32_t w, int32_t h, const uint32_t* dBits, uint8_t* dBytes, cudaStream_t stream)
{
    const int32_t n = w * h;
    if (n <= 0) return true;
    k_maskBytes<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>((const ChartInfo*)dCharts, chart, dBits, dBytes);
    g_counters.kernelLaunches++;
    return PT_CUDA(cudaGetLastError());
}

// AtlasFn with one task chunk over all charts (lightmap.c:779-823): charts in the order given. dAtlas: atlasCount
// bit images of atlasSize rows x atlasSize/32 words, zeroed; dBest: one word. Returns false when a chart fits nowhere.
bool placeCharts(const int32_t* firstNode, const int32_t* maskSize, const float* area, const std::vector<int32_t>& wordBase,
                 int32_t chartCount, int32_t atlasSize, int32_t atlasCount, uint32_t* dAtlas, const uint32_t* dBits, uint32_t* dBest,
                 int32_t* atlasIndexOut, int32_t* translationOut, cudaStream_t stream)
{
    const int32_t kRowReset = -(1 << 20);
    const int32_t atlasRowWords = (atlasSize + 31) / 32;
    const size_t atlasWords = (size_t)atlasRowWords * (size_t)atlasSize;
    int32_t prevAtlas = 0, prevRow = kRowReset;
    float prevArea = (float)(1 << 23);
    for (int32_t c = 0; c < chartCount; ++c)
    {
        ChartInfo ci;
        ci.firstNode = firstNode[c]; ci.nodeCount = firstNode[c + 1] - firstNode[c];
        ci.w = maskSize[2 * c]; ci.h = std::max(maskSize[2 * c + 1], 0);
        ci.rowWords = (ci.w + 31) / 32; ci.wordBase = wordBase[(size_t)c];
        atlasIndexOut[c] = -1; translationOut[2 * c] = translationOut[2 * c + 1] = 0;
        if (area[c] < prevArea * 0.9f) { prevArea = area[c]; prevAtlas = 0; prevRow = kRowReset; }
        const int32_t rangeX = atlasSize - ci.w, rangeY = atlasSize - ci.h;
        bool placed = false;
        for (int attempt = 0; attempt < 2 && !placed; ++attempt)
        {
            for (int32_t a = prevAtlas; a < atlasCount && !placed; ++a)
            {
                const int32_t y0 = (prevRow != kRowReset) ? prevRow : 0;
                uint32_t best = 0xffffffffu;
                if (rangeX > 0 && rangeY > y0)
                {
                    const long long cand = (long long)(rangeY - y0) * rangeX;
                    if (!PT_CUDA(cudaMemcpyAsync(dBest, &best, 4, cudaMemcpyHostToDevice, stream))) return false;
                    k_atlasFind<<<(unsigned)((cand + 127) / 128), 128, 0, stream>>>(dAtlas + a * atlasWords, atlasSize, atlasRowWords, dBits, ci,
                                                                                   y0, rangeX, rangeY, dBest);
                    g_counters.kernelLaunches++;
                    if (!PT_CUDA(cudaMemcpyAsync(&best, dBest, 4, cudaMemcpyDeviceToHost, stream)) || !PT_CUDA(cudaStreamSynchronize(stream))) return false;
                }
                if (best != 0xffffffffu)
                {
                    const int32_t tx = (int32_t)(best & 0xffffu), ty = (int32_t)(best >> 16);
                    prevAtlas = a; prevRow = ty;
                    const int32_t n = ci.h * ci.rowWords;
                    if (n > 0)
                    {
                        k_atlasWrite<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(dAtlas + a * atlasWords, atlasRowWords, dBits, ci, tx, ty);
                        g_counters.kernelLaunches++;
                    }
                    atlasIndexOut[c] = a; translationOut[2 * c] = tx; translationOut[2 * c + 1] = ty;
                    placed = true;
                }
                else prevRow = kRowReset;
            }
            if (!placed) { prevArea = area[c]; prevAtlas = 0; prevRow = kRowReset; }     // AtlasFn's retry from the start
        }
        if (!placed)
        {
            // the reference would search forever here (a chart larger than an atlas, or every atlas full)
            setError("LmPack_AtlasDesc: chart %d (%d x %d texels) fits in none of the %d atlases", c, ci.w, ci.h, atlasCount);
            return false;
        }
    }
    return PT_CUDA(cudaGetLastError());
}

} // namespace pt
