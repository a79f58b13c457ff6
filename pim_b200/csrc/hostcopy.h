// hostcopy.h — large copies between the caller's memory and the device.
//
// The reference's buffers (ray arrays, PtTrace-less images, lightmap pages, screenshots) come from its own
// allocator: pageable memory. cudaMemcpy of pageable memory runs through the driver's bounce buffer at about
// 15 GB/s on this box (1080p tonemap call: 41.5 MB in 2.8 ms around a 20 us kernel). Copies of stagedMin() bytes or
// more therefore go through two page-locked slots per device: the resident host threads (hostpool.h) fill a slot
// with a parallel memcpy while the previous slot is on the bus, and the other way round for results. Memory that
// is already page-locked (cudaHostAlloc / cudaHostRegister, e.g. PtTrace images) is copied directly.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>

#include "arena.h"
#include "hostpool.h"

namespace pt {

constexpr size_t kRingSlotBytes = (size_t)8 << 20;
// smallest copy that goes through the ring; PIMPT_STAGED_COPY_MIN (bytes) overrides it, 0 switches the ring off
inline size_t stagedMin()
{
    static const size_t v = [] {
        const char* e = getenv("PIMPT_STAGED_COPY_MIN");
        if (!e || !*e) return (size_t)2 << 20;
        const unsigned long long b = strtoull(e, nullptr, 10);
        return b == 0 ? ~(size_t)0 : (size_t)b;
    }();
    return v;
}

struct CopyRing
{
    std::mutex mutex;
    char* slots[2] = { nullptr, nullptr };
    cudaEvent_t ev[2] = { nullptr, nullptr };
    bool pending[2] = { false, false };

    bool ensure()
    {
        if (slots[0]) return true;
        char* block = nullptr;
        if (!cudaOk(cudaMallocHost((void**)&block, 2 * kRingSlotBytes), "cudaMallocHost(copy ring)")) return false;
        slots[0] = block; slots[1] = block + kRingSlotBytes;
        return cudaOk(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming), "cudaEventCreate") &&
               cudaOk(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming), "cudaEventCreate");
    }
    bool drain(int slot)
    {
        if (!pending[slot]) return true;
        pending[slot] = false;
        return cudaOk(cudaEventSynchronize(ev[slot]), "cudaEventSynchronize(copy ring)");
    }
    void release()      // with the ring's device current
    {
        if (slots[0]) cudaFreeHost(slots[0]);
        for (int k = 0; k < 2; ++k) { if (ev[k]) cudaEventDestroy(ev[k]); ev[k] = nullptr; slots[k] = nullptr; pending[k] = false; }
    }
};

inline CopyRing& copyRing(int device)
{
    static CopyRing rings[kMaxDevices];
    return rings[device < 0 || device >= kMaxDevices ? 0 : device];
}

inline bool isPageLocked(const void* p)
{
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged;
}

constexpr size_t kRingFirstChunk = (size_t)1 << 20;

// hSrc is read completely before this returns; the device copy is ordered on `st` like a cudaMemcpyAsync
inline bool copyIn(void* dDst, const void* hSrc, size_t bytes, cudaStream_t st)
{
    if (bytes == 0) return true;
    if (bytes < stagedMin() || isPageLocked(hSrc))
        return cudaOk(cudaMemcpyAsync(dDst, hSrc, bytes, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(H2D)");
    CopyRing& R = copyRing(currentDevice());
    std::lock_guard<std::mutex> lock(R.mutex);
    if (!R.ensure()) return false;
    // chunks grow 1, 2, 4, 8, 8 ... MB: the first one, which nothing overlaps, is short
    int k = 0;
    size_t chunk = kRingFirstChunk;
    for (size_t off = 0, n = 0; off < bytes; off += n, ++k)
    {
        const int slot = k & 1;
        n = bytes - off < chunk ? bytes - off : chunk;
        chunk = chunk * 2 < kRingSlotBytes ? chunk * 2 : kRingSlotBytes;
        if (!R.drain(slot)) return false;
        hostCopyParallel(R.slots[slot], (const char*)hSrc + off, n);
        if (!cudaOk(cudaMemcpyAsync((char*)dDst + off, R.slots[slot], n, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(H2D ring)") ||
            !cudaOk(cudaEventRecord(R.ev[slot], st), "cudaEventRecord")) return false;
        R.pending[slot] = true;
    }
    return true;
}

// ordered on `st` after what produced dSrc. Staged copies have landed in hDst when this returns; direct ones when
// the caller has synchronised the stream (every caller does, right after its last copyOut)
inline bool copyOut(void* hDst, const void* dSrc, size_t bytes, cudaStream_t st)
{
    if (bytes == 0) return true;
    if (bytes < stagedMin() || isPageLocked(hDst))
        return cudaOk(cudaMemcpyAsync(hDst, dSrc, bytes, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync(D2H)");
    CopyRing& R = copyRing(currentDevice());
    std::lock_guard<std::mutex> lock(R.mutex);
    if (!R.ensure()) return false;
    // chunks shrink towards the end (half of what is left, at least 1 MB): the last one's host copy, which
    // nothing overlaps, is short
    size_t prevOff = 0, prevN = 0; int prevSlot = -1, k = 0;
    for (size_t off = 0, n = 0; off < bytes; off += n, ++k)
    {
        const int slot = k & 1;
        const size_t left = bytes - off;
        n = left / 2 > kRingSlotBytes ? kRingSlotBytes : left / 2;
        if (n < kRingFirstChunk) n = left < 2 * kRingFirstChunk ? left : kRingFirstChunk;
        n &= ~(size_t)63; if (n == 0 || left - n < 64) n = left;
        if (!R.drain(slot)) return false;
        if (!cudaOk(cudaMemcpyAsync(R.slots[slot], (const char*)dSrc + off, n, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync(D2H ring)") ||
            !cudaOk(cudaEventRecord(R.ev[slot], st), "cudaEventRecord")) return false;
        R.pending[slot] = true;
        if (prevSlot >= 0)
        {
            if (!R.drain(prevSlot)) return false;
            hostCopyParallel((char*)hDst + prevOff, R.slots[prevSlot], prevN);
        }
        prevSlot = slot; prevOff = off; prevN = n;
    }
    if (prevSlot >= 0)
    {
        if (!R.drain(prevSlot)) return false;
        hostCopyParallel((char*)hDst + prevOff, R.slots[prevSlot], prevN);
    }
    return true;
}

} // namespace pt
