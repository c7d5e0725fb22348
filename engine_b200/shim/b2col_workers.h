// Helper threads of the drop-in systems.  The reference engine is single-threaded (src/application.cpp:390) and so is
// everything the shims do to the ECS; but three of their per-frame loops are pure memory traffic over component arrays
// of 144-byte TransformComponents — gathering the TRS of the entities that move, comparing collider boxes with the
// uploaded copy, scattering world matrices back — and at 1M entities each is milliseconds of cache misses on one core
// (bench.py update_e2e).  While the engine's thread is inside onUpdate nothing else touches the ECS, so these loops are
// cut into contiguous slices over a small persistent pool.  B2COL_SHIM_THREADS=1 (or a single-core machine) runs
// everything on the calling thread.
#pragma once

#include <condition_variable>
#include <cstddef>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace engine {

class B2colWorkers {
   public:
    static B2colWorkers& Get()
    {
        static B2colWorkers pool;
        return pool;
    }

    unsigned threads() const { return static_cast<unsigned>(workers_.size()) + 1u; }

    // fn(lo, hi) over [0, n) in contiguous slices, the first one on the calling thread; returns when all are done.
    // Fewer than `grain` items per slice are not worth a wake-up: small loops run inline.
    template <class F>
    void ParallelFor(std::size_t n, std::size_t grain, F&& fn)
    {
        std::size_t parts = grain ? n / grain : 1;
        if (parts > threads()) parts = threads();
        if (parts <= 1) {
            if (n) fn(std::size_t{0}, n);
            return;
        }
        const std::function<void(std::size_t, std::size_t)> f = std::ref(fn);
        {
            std::lock_guard<std::mutex> lock(m_);
            job_ = &f;
            n_ = n;
            parts_ = parts;
            next_part_ = 1;  // part 0 is the caller's
            pending_ = parts - 1;
            ++generation_;
        }
        wake_.notify_all();
        fn(std::size_t{0}, n / parts);
        std::unique_lock<std::mutex> lock(m_);
        done_.wait(lock, [&] { return pending_ == 0; });
        job_ = nullptr;
    }

   private:
    B2colWorkers()
    {
        unsigned want = std::thread::hardware_concurrency() / 2u;
        if (want > 8u) want = 8u;
        if (const char* e = std::getenv("B2COL_SHIM_THREADS")) want = static_cast<unsigned>(std::atoi(e));
        if (want < 1u) want = 1u;
        for (unsigned i = 1; i < want; ++i) workers_.emplace_back([this] { Loop(); });
    }
    ~B2colWorkers()
    {
        {
            std::lock_guard<std::mutex> lock(m_);
            stop_ = true;
        }
        wake_.notify_all();
        for (auto& t : workers_) t.join();
    }
    void Loop()
    {
        unsigned long seen = 0;
        std::unique_lock<std::mutex> lock(m_);
        for (;;) {
            wake_.wait(lock, [&] { return stop_ || (generation_ != seen && next_part_ < parts_); });
            if (stop_) return;
            while (next_part_ < parts_) {
                const std::size_t part = next_part_++;
                const std::size_t lo = n_ * part / parts_, hi = n_ * (part + 1) / parts_;
                const auto* job = job_;
                lock.unlock();
                (*job)(lo, hi);
                lock.lock();
                if (--pending_ == 0) done_.notify_all();
            }
            seen = generation_;
        }
    }

    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable wake_, done_;
    const std::function<void(std::size_t, std::size_t)>* job_ = nullptr;
    std::size_t n_ = 0, parts_ = 0, next_part_ = 0, pending_ = 0;
    unsigned long generation_ = 0;
    bool stop_ = false;
};

}  // namespace engine
