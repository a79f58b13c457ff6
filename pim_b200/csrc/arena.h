// arena.h — the library's staging memory: per CUDA device one growable device block and one
// growable page-locked host block, handed out per call.
//
// Every entry point with host buffers (Pt_IntersectN, Pt_TraceRayN, Pt_RayGen, Cubemap_*,
// AmbCube_Bake, Pt_Tonemap, LmPack_BakeDesc ...) needs device staging for the length of the
// call. A cudaMalloc / cudaFree pair per call costs more than the kernels of the small calls
// (ambient cube: 1.65 ms per call against 1.23 ms on the CPU; 1080p tonemap: 4.1 ms around a
// 21 us kernel), so a call states everything it needs up front (Lease), the arena grows once
// if it is too small and the steady state allocates nothing. All these entry points block until
// their stream has drained, so a lease never outlives its call; the arena's mutex serialises
// calls from different host threads on the same device (threads on different devices — the
// in-library multi-GPU mode — use different arenas and do not meet).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <mutex>
#include <vector>

namespace pt {

bool cudaOk(cudaError_t e, const char* what);

constexpr int kMaxDevices = 64;

struct Arena
{
    std::mutex mutex;
    char* dev = nullptr; size_t devCap = 0;
    char* host = nullptr; size_t hostCap = 0;
    cudaEvent_t evA = nullptr, evB = nullptr;   // the pair KernelTimer records (created on first use)

    void release()      // with the arena's device current
    {
        if (dev) cudaFree(dev);
        if (host) cudaFreeHost(host);
        if (evA) cudaEventDestroy(evA);
        if (evB) cudaEventDestroy(evB);
        evA = evB = nullptr;
        dev = host = nullptr; devCap = hostCap = 0;
    }
};

Arena& stagingArena(int device);
int currentDevice();

// Collects the buffers of one call, then commit() sizes the arena and hands out the pointers.
// Holds the arena's lock from construction to destruction: one lease per call, never nested.
class Lease
{
public:
    Lease() : arena_(stagingArena(currentDevice())), lock_(arena_.mutex) {}
    template <class T> Lease& dev(T*& p, size_t count) { p = nullptr; reqs_.push_back({ (void**)&p, align(sizeof(T) * count), false }); return *this; }
    template <class T> Lease& host(T*& p, size_t count) { p = nullptr; reqs_.push_back({ (void**)&p, align(sizeof(T) * count), true }); return *this; }
    Arena& arena() { return arena_; }
    bool commit()
    {
        Arena& A = arena_;
        if (!A.evA && !(cudaOk(cudaEventCreate(&A.evA), "cudaEventCreate") && cudaOk(cudaEventCreate(&A.evB), "cudaEventCreate"))) return false;
        size_t needDev = 0, needHost = 0;
        for (const Req& r : reqs_) (r.host ? needHost : needDev) += r.bytes;
        if (needDev > A.devCap)
        {
            if (A.dev) cudaFree(A.dev);
            A.dev = nullptr; A.devCap = 0;
            const size_t cap = grow(needDev);
            if (!cudaOk(cudaMalloc((void**)&A.dev, cap), "cudaMalloc(staging arena)")) return false;
            A.devCap = cap;
        }
        if (needHost > A.hostCap)
        {
            if (A.host) cudaFreeHost(A.host);
            A.host = nullptr; A.hostCap = 0;
            const size_t cap = grow(needHost);
            if (!cudaOk(cudaMallocHost((void**)&A.host, cap), "cudaMallocHost(staging arena)")) return false;
            A.hostCap = cap;
        }
        size_t offDev = 0, offHost = 0;
        for (const Req& r : reqs_)
        {
            if (r.host) { *r.out = A.host + offHost; offHost += r.bytes; }
            else { *r.out = A.dev + offDev; offDev += r.bytes; }
        }
        reqs_.clear();
        return true;
    }

private:
    struct Req { void** out; size_t bytes; bool host; };
    static size_t align(size_t b) { return (b + 255) & ~(size_t)255; }
    static size_t grow(size_t need) { return (need + need / 4 + (1u << 20)) & ~(size_t)((1u << 20) - 1); }
    Arena& arena_;
    std::unique_lock<std::mutex> lock_;
    std::vector<Req> reqs_;
};

} // namespace pt
