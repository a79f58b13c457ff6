// hostpool.h — a few resident host threads for the byte shuffling the host-buffer entry points do before and
// after their kernels (LmPack_BakeDesc: find the covered texels of a page, pack their 108 bytes into page-locked
// staging, unpack the results). The work is memory bound and a millisecond long, so threads started per call
// would cost what they save; these wait on a condition variable between calls.
#pragma once

#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace pt {

class HostPool
{
public:
    static HostPool& get() { static HostPool pool; return pool; }

    int workers() const { return (int)threads_.size() + 1; }

    // fn(w, workers()) once per worker, w = 0 on the caller; returns when all are done. One run at a time.
    void run(const std::function<void(int, int)>& fn)
    {
        afterFork();
        Sync& S = *sync_;
        std::lock_guard<std::mutex> serial(S.runMutex);
        if (threads_.empty()) { fn(0, 1); return; }
        {
            std::lock_guard<std::mutex> lock(S.mutex);
            fn_ = &fn;
            pending_ = (int)threads_.size();
            ++generation_;
        }
        S.wake.notify_all();
        fn(0, workers());
        std::unique_lock<std::mutex> lock(S.mutex);
        S.done.wait(lock, [&] { return pending_ == 0; });
        fn_ = nullptr;
    }

    // joins the threads (PtSys_Shutdown); the next start() starts them again
    void stop()
    {
        afterFork();
        Sync& S = *sync_;
        std::lock_guard<std::mutex> serial(S.runMutex);
        {
            std::lock_guard<std::mutex> lock(S.mutex);
            quit_ = true;
        }
        S.wake.notify_all();
        for (std::thread& t : threads_) t.join();
        threads_.clear();
        quit_ = false;
        started_ = false;
    }

    void start()
    {
        afterFork();
        Sync& S = *sync_;
        std::lock_guard<std::mutex> serial(S.runMutex);
        if (started_) return;
        started_ = true;
        unsigned hw = std::thread::hardware_concurrency();
        int n = hw == 0 ? 1 : (int)(hw > 8 ? 8 : hw);
        if (const char* e = getenv("PIMPT_HOST_THREADS")) { const int v = atoi(e); if (v >= 1 && v <= 64) n = v; }
        unsigned long long seen;
        { std::lock_guard<std::mutex> lock(S.mutex); seen = generation_; }      // runs before this start are not theirs
        for (int w = 1; w < n; ++w) threads_.emplace_back([this, w, seen, &S] { loop(S, w, seen); });
    }

private:
    // the locks and condition variables live in a block of their own: a forked child inherits them in whatever
    // state the parent's threads left them (waiters that do not exist in the child), so it drops the block,
    // lets go of the thread handles and starts over
    struct Sync { std::mutex runMutex, mutex; std::condition_variable wake, done; };

    HostPool() : sync_(new Sync()), ownerPid_((long)getpid()) {}
    ~HostPool() { if (ownerPid_ == (long)getpid()) { stop(); delete sync_; } else abandon(); }

    void abandon()
    {
        for (std::thread& t : threads_) { try { t.detach(); } catch (...) {} }
        threads_.clear();
        sync_ = new Sync();         // the old block is left alone on purpose
        started_ = false; quit_ = false; pending_ = 0; fn_ = nullptr;
        ownerPid_ = (long)getpid();
    }
    void afterFork() { if (ownerPid_ != (long)getpid()) abandon(); }

    void loop(Sync& S, int w, unsigned long long seen)
    {
        for (;;)
        {
            const std::function<void(int, int)>* fn = nullptr;
            int total = 0;
            {
                std::unique_lock<std::mutex> lock(S.mutex);
                S.wake.wait(lock, [&] { return quit_ || generation_ != seen; });
                if (quit_) return;
                seen = generation_;
                fn = fn_;
                total = (int)threads_.size() + 1;
            }
            (*fn)(w, total);
            {
                std::lock_guard<std::mutex> lock(S.mutex);
                if (--pending_ == 0) S.done.notify_one();
            }
        }
    }

    Sync* sync_;
    std::vector<std::thread> threads_;
    const std::function<void(int, int)>* fn_ = nullptr;
    unsigned long long generation_ = 0;
    int pending_ = 0;
    bool quit_ = false, started_ = false;
    long ownerPid_ = 0;
};

// [begin, end) of worker w's share of n items
inline void hostShare(size_t n, int w, int workers, size_t& begin, size_t& end)
{
    const size_t base = n / (size_t)workers, rem = n % (size_t)workers;
    begin = (size_t)w * base + ((size_t)w < rem ? (size_t)w : rem);
    end = begin + base + ((size_t)w < rem ? 1 : 0);
}

// memcpy with non-temporal stores: the destination is read next by the DMA engine (or, for results, by nobody
// soon), so the lines need not be fetched for ownership nor kept in cache
inline void streamCopy(char* dst, const char* src, size_t bytes)
{
#if defined(__SSE2__)
    const size_t head = (16 - ((uintptr_t)dst & 15)) & 15;
    if (bytes < 256 + head) { memcpy(dst, src, bytes); return; }
    memcpy(dst, src, head);
    dst += head; src += head; bytes -= head;
    const size_t blocks = bytes >> 6;
    for (size_t i = 0; i < blocks; ++i, dst += 64, src += 64)
    {
        const __m128i a = _mm_loadu_si128((const __m128i*)src), b = _mm_loadu_si128((const __m128i*)(src + 16));
        const __m128i c = _mm_loadu_si128((const __m128i*)(src + 32)), d = _mm_loadu_si128((const __m128i*)(src + 48));
        _mm_stream_si128((__m128i*)dst, a); _mm_stream_si128((__m128i*)(dst + 16), b);
        _mm_stream_si128((__m128i*)(dst + 32), c); _mm_stream_si128((__m128i*)(dst + 48), d);
    }
    _mm_sfence();
    memcpy(dst, src, bytes & 63);
#else
    memcpy(dst, src, bytes);
#endif
}

// bytes a multiple of 16, both pointers 16-byte aligned
inline void streamStore16(void* dst, const void* src, size_t bytes)
{
#if defined(__SSE2__)
    for (size_t o = 0; o < bytes; o += 16) _mm_stream_si128((__m128i*)((char*)dst + o), _mm_load_si128((const __m128i*)((const char*)src + o)));
#else
    memcpy(dst, src, bytes);
#endif
}

inline void hostCopyParallel(void* dst, const void* src, size_t bytes)
{
    HostPool& pool = HostPool::get();
    pool.start();
    pool.run([&](int w, int workers)
    {
        size_t a, b; hostShare(bytes >> 6, w, workers, a, b);      // 64-byte lines
        a <<= 6; b = (w == workers - 1) ? bytes : b << 6;
        if (b > a) streamCopy((char*)dst + a, (const char*)src + a, b - a);
    });
}

} // namespace pt
