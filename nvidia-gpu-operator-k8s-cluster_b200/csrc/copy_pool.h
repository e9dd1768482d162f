// copy_pool.h — a small pool of host threads that memcpy in parallel.
//
// Used by the host-buffer entry points when the caller's buffers are ordinary pageable memory
// (malloc, a Go slice, a numpy array): the DMA engines cannot read those directly, and the
// driver's own pageable path reaches ~13 GB/s on this box.  Copying pageable <-> pinned staging
// with several threads while the previous chunk is on the wire is ~3x faster.  One thread's
// memcpy is ~10 GB/s; the pool exists to get past that.
#pragma once

#include <algorithm>
#include <condition_variable>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace b200v {

class CopyPool {
public:
    // The process-wide pool.  Never destroyed: its detached workers simply end with the process,
    // which keeps static-destruction order out of the picture.
    static CopyPool &instance() {
        static CopyPool *p = new CopyPool();
        return *p;
    }

    int threads() const { return nthreads_; }

    // memcpy(dst, src, bytes) split over the pool; returns when every byte is copied.  Calls from
    // different host threads are serialised (each one uses the whole pool).
    void copy(void *dst, const void *src, size_t bytes) {
        if (bytes == 0) return;
        constexpr size_t kMinPart = 1 << 20;
        if (bytes < 2 * kMinPart || nthreads_ <= 1) {
            std::memcpy(dst, src, bytes);
            return;
        }
        std::lock_guard<std::mutex> serial(call_mu_);
        const size_t parts = std::min<size_t>((size_t)nthreads_ + 1, bytes / kMinPart);
        const size_t per = ((bytes / parts) + 4095) & ~(size_t)4095;
        {
            std::lock_guard<std::mutex> l(mu_);
            jobs_.clear();
            for (size_t off = per; off < bytes; off += per)   // part 0 is done by the caller below
                jobs_.push_back({(char *)dst + off, (const char *)src + off, std::min(per, bytes - off)});
            next_ = 0;
            pending_ = jobs_.size();
        }
        cv_.notify_all();
        std::memcpy(dst, src, std::min(per, bytes));
        std::unique_lock<std::mutex> l(mu_);
        done_cv_.wait(l, [&] { return pending_ == 0; });
    }

private:
    struct Job {
        char *d;
        const char *s;
        size_t n;
    };

    CopyPool() {
        unsigned hw = std::thread::hardware_concurrency();
        // measured on a 16-vCPU B200 host, 2^28-element call: 4 threads 35, 8 -> 48, 12 -> 56,
        // 16 -> 51 GB/s end to end
        nthreads_ = (int)std::min(12u, std::max(1u, hw * 3 / 4));
        if (const char *e = std::getenv("B200V_COPY_THREADS")) {
            const int v = std::atoi(e);
            if (v >= 1 && v <= 64) nthreads_ = v;
        }
        for (int i = 0; i < nthreads_; ++i) std::thread([this] { worker(); }).detach();
    }

    void worker() {
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> l(mu_);
                cv_.wait(l, [&] { return next_ < jobs_.size(); });
                j = jobs_[next_++];
            }
            std::memcpy(j.d, j.s, j.n);
            {
                std::lock_guard<std::mutex> l(mu_);
                if (--pending_ == 0) done_cv_.notify_all();
            }
        }
    }

    int nthreads_ = 1;
    std::mutex call_mu_, mu_;
    std::condition_variable cv_, done_cv_;
    std::vector<Job> jobs_;
    size_t next_ = 0, pending_ = 0;
};

}  // namespace b200v
