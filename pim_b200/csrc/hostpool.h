// hostpool.h — a few resident host threads for the byte shuffling the host-buffer entry points do before and
// after their kernels (LmPack_BakeDesc: find the covered texels of a page, pack their 108 bytes into page-locked
// staging, unpack the results). The work is memory bound and a millisecond long, so threads started per call
// would cost what they save; these wait on a condition variable between calls.
#pragma once

#include <stdint.h>
#include <stdlib.h>
#include <string.h>
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
        std::lock_guard<std::mutex> serial(runMutex_);
        if (threads_.empty()) { fn(0, 1); return; }
        {
            std::lock_guard<std::mutex> lock(mutex_);
            fn_ = &fn;
            pending_ = (int)threads_.size();
            ++generation_;
        }
        wake_.notify_all();
        fn(0, workers());
        std::unique_lock<std::mutex> lock(mutex_);
        done_.wait(lock, [&] { return pending_ == 0; });
        fn_ = nullptr;
    }

    // joins the threads (PtSys_Shutdown); the next run() starts them again
    void stop()
    {
        std::lock_guard<std::mutex> serial(runMutex_);
        {
            std::lock_guard<std::mutex> lock(mutex_);
            quit_ = true;
        }
        wake_.notify_all();
        for (std::thread& t : threads_) t.join();
        threads_.clear();
        quit_ = false;
        started_ = false;
    }

    void start()
    {
        std::lock_guard<std::mutex> serial(runMutex_);
        if (started_) return;
        started_ = true;
        unsigned hw = std::thread::hardware_concurrency();
        int n = hw == 0 ? 1 : (int)(hw > 8 ? 8 : hw);
        if (const char* e = getenv("PIMPT_HOST_THREADS")) { const int v = atoi(e); if (v >= 1 && v <= 64) n = v; }
        unsigned long long seen;
        { std::lock_guard<std::mutex> lock(mutex_); seen = generation_; }      // runs before this start are not theirs
        for (int w = 1; w < n; ++w) threads_.emplace_back([this, w, seen] { loop(w, seen); });
    }

private:
    HostPool() = default;
    ~HostPool() { stop(); }

    void loop(int w, unsigned long long seen)
    {
        for (;;)
        {
            const std::function<void(int, int)>* fn = nullptr;
            int total = 0;
            {
                std::unique_lock<std::mutex> lock(mutex_);
                wake_.wait(lock, [&] { return quit_ || generation_ != seen; });
                if (quit_) return;
                seen = generation_;
                fn = fn_;
                total = (int)threads_.size() + 1;
            }
            (*fn)(w, total);
            {
                std::lock_guard<std::mutex> lock(mutex_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }

    std::mutex runMutex_, mutex_;
    std::condition_variable wake_, done_;
    std::vector<std::thread> threads_;
    const std::function<void(int, int)>* fn_ = nullptr;
    unsigned long long generation_ = 0;
    int pending_ = 0;
    bool quit_ = false, started_ = false;
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
