// TEST INFRASTRUCTURE ONLY -- stand-in for the legacy Intel TBB headers the reference includes (tbb_stddef.h, task_scheduler_init.h,
// blocked_range.h, parallel_for.h, parallel_reduce.h, parallel_for_each.h), implemented on a persistent std::thread pool so that
// the reference's own bodies run multi-threaded: chunks of the range are handed out from an atomic cursor; parallel_reduce gives
// every worker a split copy of the body (Body(body, tbb::split())) and joins them in order, as TBB does.
#pragma once
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstddef>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>
#define TBB_INTERFACE_VERSION 0
namespace tbb {
struct split {};
inline int TBB_runtime_interface_version() { return 0; }
namespace detail_shim {
inline int& nthreads() { static int n = int(std::max(1u, std::thread::hardware_concurrency())); return n; }
}
class task_scheduler_init {
public:
    static const int automatic = -1, deferred = -2;
    explicit task_scheduler_init(int n = automatic) { if (n != deferred) initialize(n); }
    void initialize(int n = automatic) { detail_shim::nthreads() = n > 0 ? n : default_num_threads(); m_active = true; }
    void terminate() { m_active = false; }
    bool is_active() const { return m_active; }
    static int default_num_threads() { return int(std::max(1u, std::thread::hardware_concurrency())); }
private:
    bool m_active = false;
};
template <class Value> class blocked_range {
public:
    using const_iterator = Value; using size_type = std::size_t;
    blocked_range() = default;
    blocked_range(Value b, Value e, size_type grain = 1) : m_b(b), m_e(e), m_grain(grain ? grain : 1) {}
    Value begin() const { return m_b; } Value end() const { return m_e; }
    size_type size() const { return size_type(m_e - m_b); } size_type grainsize() const { return m_grain; }
    bool empty() const { return !(m_b < m_e); } bool is_divisible() const { return m_grain < size(); }
private:
    Value m_b{}, m_e{}; size_type m_grain = 1;
};
namespace detail_shim {
// A persistent worker pool (threads are created once, not per parallel_for: the reference's ARACNE calls parallel_for for every
// block of 16384 edges). A job is a range of chunk indices handed out from an atomic cursor; the calling thread takes part.
// A parallel_* call made from inside a worker runs inline (no nested parallelism, like a saturated TBB arena).
class Pool {
public:
    static Pool& get() { static Pool p; return p; }
    template <class Fn> void run(std::size_t nchunks, int want_threads, Fn&& fn) {          // fn(worker_slot, chunk_index)
        if (nchunks == 0) return;
        if (t_inside || want_threads <= 1 || nchunks == 1) { for (std::size_t k = 0; k < nchunks; ++k) fn(0, k); return; }
        std::unique_lock<std::mutex> job_lock(m_job);                                       // one job at a time
        ensure(want_threads - 1);
        const int helpers = std::min<int>(want_threads - 1, (int)m_threads.size());
        std::function<void(int, std::size_t)> f = [&](int w, std::size_t k) { fn(w, k); };
        {
            std::lock_guard<std::mutex> lk(m_mx);
            m_fn = &f; m_n = nchunks; m_cursor.store(0); m_active = helpers; m_helpers = helpers; ++m_epoch;
        }
        m_cv.notify_all();
        t_inside = true;
        for (;;) { const std::size_t k = m_cursor.fetch_add(1); if (k >= nchunks) break; f(0, k); }
        t_inside = false;
        std::unique_lock<std::mutex> lk(m_mx);
        m_done.wait(lk, [&] { return m_active == 0; });
        m_fn = nullptr;
    }
    int max_slots(int want_threads) const { return std::max(1, want_threads); }
private:
    Pool() = default;
    ~Pool() { { std::lock_guard<std::mutex> lk(m_mx); m_stop = true; ++m_epoch; } m_cv.notify_all(); for (auto& t : m_threads) t.join(); }
    void ensure(int n) {
        while ((int)m_threads.size() < n) {
            const int slot = (int)m_threads.size() + 1;
            m_threads.emplace_back([this, slot] {
                t_inside = true;
                std::size_t seen = 0;
                for (;;) {
                    std::function<void(int, std::size_t)>* f; std::size_t n;
                    {
                        std::unique_lock<std::mutex> lk(m_mx);
                        m_cv.wait(lk, [&] { return m_stop || (m_epoch != seen && slot <= m_helpers && m_fn); });
                        if (m_stop) return;
                        seen = m_epoch; f = m_fn; n = m_n;
                    }
                    for (;;) { const std::size_t k = m_cursor.fetch_add(1); if (k >= n) break; (*f)(slot, k); }
                    { std::lock_guard<std::mutex> lk(m_mx); if (--m_active == 0) m_done.notify_all(); }
                }
            });
        }
    }
    static thread_local bool t_inside;
    std::mutex m_job, m_mx; std::condition_variable m_cv, m_done;
    std::vector<std::thread> m_threads;
    std::function<void(int, std::size_t)>* m_fn = nullptr; std::size_t m_n = 0, m_epoch = 0; std::atomic<std::size_t> m_cursor{0};
    int m_active = 0, m_helpers = 0; bool m_stop = false;
};
inline thread_local bool Pool::t_inside = false;

// chunk size: the range's grain size, but at most ~1/(8 * threads) of the range so that the load balances
template <class Value> std::size_t chunk_of(const blocked_range<Value>& r, int nt) {
    const std::size_t n = r.size(); std::size_t c = std::max<std::size_t>(1, n / (std::size_t(nt) * 8));
    return std::max<std::size_t>(1, std::min<std::size_t>(std::max<std::size_t>(r.grainsize(), 1), c));
}
// per_slot(slot) must return the callable body(sub_range) of that worker slot
template <class Value, class Fn> void run_chunks(const blocked_range<Value>& r, int nt, Fn&& per_chunk) {   // per_chunk(slot, sub_range)
    const std::size_t n = r.empty() ? 0 : r.size(); if (n == 0) return;
    const std::size_t chunk = chunk_of(r, nt), nchunks = (n + chunk - 1) / chunk;
    Pool::get().run(nchunks, nt, [&](int slot, std::size_t k) {
        const std::size_t b = k * chunk;
        per_chunk(slot, blocked_range<Value>(r.begin() + b, r.begin() + std::min(n, b + chunk), r.grainsize()));
    });
}
}
template <class Value, class Body> void parallel_for(const blocked_range<Value>& r, const Body& body) {
    const std::size_t n = r.empty() ? 0 : r.size(); if (n == 0) return;
    const int nt = int(std::min<std::size_t>(std::size_t(std::max(1, detail_shim::nthreads())), n));
    std::vector<std::unique_ptr<Body>> copies(nt);                      // one copy of the body per worker slot, made on first use
    detail_shim::run_chunks(r, nt, [&](int slot, blocked_range<Value> sub) {
        if (!copies[slot]) copies[slot].reset(new Body(body));
        (*copies[slot])(sub);
    });
}
template <class Value, class Body> void parallel_reduce(const blocked_range<Value>& r, Body& body) {
    const std::size_t n = r.empty() ? 0 : r.size(); if (n == 0) return;
    const int nt = int(std::min<std::size_t>(std::size_t(std::max(1, detail_shim::nthreads())), n));
    std::vector<Body*> copies(nt, nullptr); copies[0] = &body;
    for (int w = 1; w < nt; ++w) copies[w] = new Body(body, split());    // before any of them runs: a split never reads a body in use
    detail_shim::run_chunks(r, nt, [&](int slot, blocked_range<Value> sub) { (*copies[slot])(sub); });
    for (int w = 1; w < nt; ++w) { body.join(*copies[w]); delete copies[w]; }
}
template <class It, class Fn> void parallel_for_each(It first, It last, const Fn& fn) {
    std::vector<It> items; for (It it = first; it != last; ++it) items.push_back(it);
    blocked_range<std::size_t> r(0, items.size(), 1);
    const int nt = int(std::min<std::size_t>(std::size_t(std::max(1, detail_shim::nthreads())), std::max<std::size_t>(1, items.size())));
    detail_shim::run_chunks(r, nt, [&](int, blocked_range<std::size_t> sub) { for (std::size_t k = sub.begin(); k < sub.end(); ++k) fn(*items[k]); });
}
template <class Container, class Fn> void parallel_for_each(Container& c, const Fn& fn) { parallel_for_each(std::begin(c), std::end(c), fn); }
class mutex {
public:
    void lock() { m.lock(); } void unlock() { m.unlock(); } bool try_lock() { return m.try_lock(); }
    class scoped_lock { public: scoped_lock() = default; explicit scoped_lock(mutex& mm) : m_m(&mm) { mm.lock(); } ~scoped_lock() { if (m_m) m_m->unlock(); }
        void acquire(mutex& mm) { mm.lock(); m_m = &mm; } void release() { if (m_m) { m_m->unlock(); m_m = nullptr; } } private: mutex* m_m = nullptr; };
private:
    std::mutex m;
};
}
