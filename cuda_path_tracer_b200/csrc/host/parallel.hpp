// parallel.hpp -- the host loader's fork/join helper.  Scene ingestion at 10 M triangles is parse- and hash-bound on
// one core; every phase of the loader that has no order dependence runs through this.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <exception>
#include <thread>
#include <vector>

namespace ptb {

// PTB_LOADER_THREADS overrides; default = hardware threads, capped at 32 (the phases are memory-bound beyond that).
inline int loader_threads() {
    if (const char* e = getenv("PTB_LOADER_THREADS")) {
        int n = atoi(e);
        if (n >= 1) return std::min(n, 256);
    }
    unsigned h = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min(h, 32u));
}

// fn(task) for task in [0, tasks), on up to `threads` threads, tasks handed out dynamically.  The first exception is
// rethrown on the caller's thread after every worker has stopped.
template <class F>
void parallel_tasks(size_t tasks, int threads, F&& fn) {
    if (tasks == 0) return;
    if (threads <= 1 || tasks == 1) {
        for (size_t i = 0; i < tasks; i++) fn(i);
        return;
    }
    std::atomic<size_t> next{0};
    std::exception_ptr first_error;
    std::atomic<bool> failed{false};
    auto worker = [&]() {
        for (;;) {
            size_t i = next.fetch_add(1);
            if (i >= tasks || failed.load()) return;
            try { fn(i); }
            catch (...) {
                if (!failed.exchange(true)) first_error = std::current_exception();
                return;
            }
        }
    };
    size_t n = std::min<size_t>((size_t)threads, tasks);
    std::vector<std::thread> pool;
    pool.reserve(n - 1);
    for (size_t t = 1; t < n; t++) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
    if (first_error) std::rethrow_exception(first_error);
}

// fn(begin, end) over [0, n) cut into contiguous ranges, one per thread.
template <class F>
void parallel_ranges(size_t n, int threads, F&& fn) {
    if (n == 0) return;
    size_t parts = std::min<size_t>((size_t)std::max(1, threads), (n + 4095) / 4096);
    parts = std::max<size_t>(parts, 1);
    parallel_tasks(parts, threads, [&](size_t p) { fn(n * p / parts, n * (p + 1) / parts); });
}

}  // namespace ptb
