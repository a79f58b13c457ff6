// Host test of pim_b200/csrc/hostpool.h: the covered-texel compaction LmPack_BakeDesc runs on the pool, against
// the plain loop. Built by tests/test_cpu_hostpool.py; not part of libpimpt.so.
#include "../../pim_b200/csrc/hostpool.h"
#include <stdint.h>

extern "C" int hp_workers(void) { pt::HostPool::get().start(); return pt::HostPool::get().workers(); }
extern "C" void hp_stop(void) { pt::HostPool::get().stop(); }

// indices of the non-zero entries of counts[0..n) into idx (two passes, as the library does it); returns how many
extern "C" int32_t hp_compact(const float* counts, int64_t n, int32_t* idx)
{
    pt::HostPool& pool = pt::HostPool::get();
    pool.start();
    std::vector<int32_t> slice((size_t)pool.workers() + 1, 0);
    pool.run([&](int w, int workers)
    {
        size_t t0, t1; pt::hostShare((size_t)n, w, workers, t0, t1);
        int32_t c = 0;
        for (size_t t = t0; t < t1; ++t) c += counts[t] != 0.0f ? 1 : 0;
        slice[(size_t)w + 1] = c;
    });
    for (size_t w = 1; w < slice.size(); ++w) slice[w] += slice[w - 1];
    pool.run([&](int w, int workers)
    {
        size_t t0, t1; pt::hostShare((size_t)n, w, workers, t0, t1);
        int32_t j = slice[(size_t)w];
        for (size_t t = t0; t < t1; ++t) if (counts[t] != 0.0f) idx[j++] = (int32_t)t;
    });
    return slice.back();
}

// every item of [0, n) visited exactly once by hostShare: visits[i] += 1
extern "C" void hp_touch(int32_t* visits, int64_t n)
{
    pt::HostPool::get().start();
    pt::HostPool::get().run([&](int w, int workers)
    {
        size_t a, b; pt::hostShare((size_t)n, w, workers, a, b);
        for (size_t i = a; i < b; ++i) visits[i] += 1;
    });
}

extern "C" void hp_copy(void* dst, const void* src, int64_t bytes) { pt::hostCopyParallel(dst, src, (size_t)bytes); }
extern "C" void hp_stream_copy(void* dst, const void* src, int64_t bytes) { pt::streamCopy((char*)dst, (const char*)src, (size_t)bytes); }

#include <sys/wait.h>
// fork after the pool has started: the child starts its own threads, runs, and exits cleanly (exit code 0)
extern "C" int hp_fork_child_runs(void)
{
    pt::HostPool::get().start();
    const pid_t pid = fork();
    if (pid == 0)
    {
        int32_t visits[64] = {};
        pt::HostPool::get().run([&](int w, int workers) { size_t a, b; pt::hostShare(64, w, workers, a, b); for (size_t i = a; i < b; ++i) visits[i] += 1; });
        int bad = 0;
        for (int i = 0; i < 64; ++i) bad += visits[i] != 1;
        pt::HostPool::get().stop();
        _exit(bad ? 3 : 0);
    }
    int status = -1;
    if (pid < 0 || waitpid(pid, &status, 0) != pid) return -1;
    return WIFEXITED(status) ? WEXITSTATUS(status) : -2;
}
