// ORACLE — TEST INFRASTRUCTURE ONLY.
// CPU restatement of AnkaChan/Gaia's VBD step (reference CPU/TBB path). Nothing under oracle/ is
// part of the product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it, and there only as the checker or as the reported CPU baseline.
//
// Parity status: "partially pinned". The reference ships no golden vectors; the
// restatement is checked against the reference's own __host__ __device__ math compiled from
// /root/reference into oracle/_ref/ (see oracle/Makefile, tests/test_oracle_vs_ref.py) and against
// fixtures generated from it (tests/golden/). Whole-trajectory outputs of the reference CPU app
// cannot be produced here (Eigen/oneTBB/Embree absent), so trajectory parity stays unpinned.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <array>
#include <string>
#include <algorithm>
#include <functional>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <atomic>

namespace orc {

struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
inline V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
inline V3& operator-=(V3& a, V3 b) { a = a - b; return a; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float sqnorm(V3 a) { return dot(a, a); }
inline float norm(V3 a) { return std::sqrt(dot(a, a)); }

// column-major 3x3 (Eigen default): m[r + 3*c]
struct M3 {
    float m[9];
    M3() { for (int i = 0; i < 9; i++) m[i] = 0.f; }
    float& operator()(int r, int c) { return m[r + 3 * c]; }
    float operator()(int r, int c) const { return m[r + 3 * c]; }
};

// a persistent pool standing in for tbb::parallel_for (M/Parallelization/CPUParallelization.h:20-28)
class Pool {
public:
    explicit Pool(int n) : nthreads_(n < 1 ? 1 : n) {
        for (int i = 1; i < nthreads_; i++) workers_.emplace_back([this, i] { loop(i); });
    }
    ~Pool() {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = true; gen_++; }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
    }
    int size() const { return nthreads_; }
    // chunks [0,n) statically; fn(begin,end)
    void parallel_for(int64_t n, const std::function<void(int64_t, int64_t)>& fn, int64_t minChunk = 256) {
        if (n <= 0) return;
        if (nthreads_ == 1 || n <= minChunk) { fn(0, n); return; }
        {
            std::lock_guard<std::mutex> lk(mu_);
            fn_ = &fn; n_ = n; pending_ = nthreads_ - 1; gen_++;
        }
        cv_.notify_all();
        run_chunk(0);
        std::unique_lock<std::mutex> lk(mu_);
        done_cv_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }
private:
    void run_chunk(int tid) {
        int64_t per = (n_ + nthreads_ - 1) / nthreads_;
        int64_t b = per * tid, e = std::min<int64_t>(n_, b + per);
        if (b < e) (*fn_)(b, e);
    }
    void loop(int tid) {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
            }
            run_chunk(tid);
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (--pending_ == 0) done_cv_.notify_one();
            }
        }
    }
    int nthreads_;
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_, done_cv_;
    const std::function<void(int64_t, int64_t)>* fn_ = nullptr;
    int64_t n_ = 0;
    int pending_ = 0;
    uint64_t gen_ = 0;
    bool stop_ = false;
};

}  // namespace orc
