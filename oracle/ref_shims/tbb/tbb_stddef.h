// TEST INFRASTRUCTURE ONLY -- stand-in for the legacy Intel TBB headers the reference includes (tbb_stddef.h, task_scheduler_init.h,
// blocked_range.h, parallel_for.h, parallel_reduce.h, parallel_for_each.h), implemented on std::thread so that the reference's
// own bodies run multi-threaded: chunks of the range are handed out from an atomic cursor; parallel_reduce gives every worker a
// split copy of the body (Body(body, tbb::split())) and joins them in order, as TBB does.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstddef>
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
// chunk size: the range's grain size, but at most ~1/(8 * threads) of the range so that the load balances
template <class Value> std::size_t chunk_of(const blocked_range<Value>& r, int nt) {
    const std::size_t n = r.size(); std::size_t c = std::max<std::size_t>(1, n / (std::size_t(nt) * 8));
    return std::max<std::size_t>(1, std::min<std::size_t>(std::max<std::size_t>(r.grainsize(), 1), c));
}
template <class Value, class Fn> void run_chunks(const blocked_range<Value>& r, Fn&& per_worker) {
    const std::size_t n = r.empty() ? 0 : r.size(); if (n == 0) return;
    const int nt = int(std::min<std::size_t>(std::size_t(std::max(1, nthreads())), n));
    const std::size_t chunk = chunk_of(r, nt);
    std::atomic<std::size_t> cursor{0};
    auto work = [&](int w) {
        per_worker(w, [&](auto&& body) {
            for (;;) { const std::size_t b = cursor.fetch_add(chunk); if (b >= n) break;
                       body(blocked_range<Value>(r.begin() + b, r.begin() + std::min(n, b + chunk), r.grainsize())); }
        });
    };
    if (nt == 1) { work(0); return; }
    std::vector<std::thread> th; th.reserve(nt);
    for (int w = 0; w < nt; ++w) th.emplace_back(work, w);
    for (auto& t : th) t.join();
}
}
template <class Value, class Body> void parallel_for(const blocked_range<Value>& r, const Body& body) {
    detail_shim::run_chunks(r, [&](int, auto&& drain) { Body mine(body); drain([&](blocked_range<Value> sub) { mine(sub); }); });
}
template <class Value, class Body> void parallel_reduce(const blocked_range<Value>& r, Body& body) {
    const std::size_t n = r.empty() ? 0 : r.size(); if (n == 0) return;
    const int nt = int(std::min<std::size_t>(std::size_t(std::max(1, detail_shim::nthreads())), n));
    std::vector<Body*> copies(nt, nullptr); copies[0] = &body;
    for (int w = 1; w < nt; ++w) copies[w] = new Body(body, split());
    detail_shim::run_chunks(r, [&](int w, auto&& drain) { Body& mine = *copies[w]; drain([&](blocked_range<Value> sub) { mine(sub); }); });
    for (int w = 1; w < nt; ++w) { body.join(*copies[w]); delete copies[w]; }
}
template <class It, class Fn> void parallel_for_each(It first, It last, const Fn& fn) {
    std::vector<It> items; for (It it = first; it != last; ++it) items.push_back(it);
    blocked_range<std::size_t> r(0, items.size(), 1);
    detail_shim::run_chunks(r, [&](int, auto&& drain) { drain([&](const blocked_range<std::size_t>& sub) { for (std::size_t k = sub.begin(); k < sub.end(); ++k) fn(*items[k]); }); });
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
