// host_staging.cuh -- moving a PAGEABLE host slice (what a Rust Vec<T> is) to and
// from the device at PCIe speed.
//
// cudaMemcpyAsync on pageable memory is staged by the driver through one pinned
// buffer by one thread: 13-15 GB/s on the B200 boxes, a quarter of the link
// (r01: 1 B u64 keys end to end 1240 ms pageable vs 294 ms pinned).  Pinning the
// caller's slice for the call (cudaHostRegister) costs more than it saves.  So
// the library owns a few pinned bounce buffers and a small pool of copy threads:
// chunk k is copied slice -> bounce by the threads while the DMA of chunk k-1
// runs, and the other way round on the way out.
//
// Replaces nothing in the reference (its slices never leave the host); it is the
// host half of the drop-in boundary b200sort_keys_host.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <emmintrin.h>
#endif

namespace b200 {

// Copy whose stores bypass the cache (MOVNTDQ): the destination is the caller's
// multi-GB slice, which is not read again soon, so a write-allocating memcpy would
// read every destination line from DRAM first (glibc only switches to streaming
// stores for single copies far larger than the 1-2 MiB slices a copy thread gets).
inline void stream_copy(unsigned char* dst, const unsigned char* src, size_t bytes) {
#if defined(__x86_64__)
    const size_t head = std::min(bytes, (size_t)((16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15));
    if (head) std::memcpy(dst, src, head);
    dst += head; src += head; bytes -= head;
    const size_t body = bytes / 64 * 64;
    for (size_t i = 0; i < body; i += 64) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 16));
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 32));
        const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 48));
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i), a);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 16), b);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 32), c);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 48), d);
    }
    _mm_sfence();
    if (bytes > body) std::memcpy(dst + body, src + body, bytes - body);
#else
    std::memcpy(dst, src, bytes);
#endif
}

// ---- persistent copy threads: memcpy split into one slice per thread ---------------
class CopyPool {
public:
    static CopyPool& get() {
        static CopyPool p;
        return p;
    }
    // streaming: the destination is not read again soon (bounce -> caller's slice)
    void copy(void* dst, const void* src, size_t bytes, bool streaming = false) {
        streaming = streaming && nt_;
        if (bytes < ((size_t)1 << 20) || workers_.empty()) {
            if (streaming) stream_copy(static_cast<unsigned char*>(dst), static_cast<const unsigned char*>(src), bytes);
            else std::memcpy(dst, src, bytes);
            return;
        }
        std::lock_guard<std::mutex> one_job(submit_);
        const int parts = (int)workers_.size() + 1;
        const size_t per = ((bytes + parts - 1) / parts + 63) / 64 * 64;
        {
            std::lock_guard<std::mutex> lk(mu_);
            dst_ = static_cast<unsigned char*>(dst);
            src_ = static_cast<const unsigned char*>(src);
            bytes_ = bytes;
            per_ = per;
            streaming_ = streaming;
            pending_ = (int)workers_.size();
            ++generation_;
        }
        cv_.notify_all();
        slice(parts - 1);  // the caller copies the last slice itself
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [&] { return pending_ == 0; });
    }

private:
    CopyPool() {
        unsigned hc = std::thread::hardware_concurrency();
        // the call blocks its thread anyway: use the cores (r01: 16 threads 2.6 Gkeys/s, 8 threads 2.2)
        int n = hc ? (int)hc - 1 : 3;
        n = std::max(1, std::min(n, 15));
        // B200SORT_COPY_THREADS=k: k copy threads in all (the caller is one of them)
        if (const char* env = std::getenv("B200SORT_COPY_THREADS")) {
            const int want = std::atoi(env);
            if (want >= 1 && want <= 64) n = want - 1;
        }
        if (const char* env = std::getenv("B200SORT_COPY_NT")) nt_ = std::atoi(env) != 0;
        for (int i = 0; i < n; ++i) workers_.emplace_back([this, i] { run(i); });
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
    }
    void slice(int i) {
        const size_t off = (size_t)i * per_;
        if (off >= bytes_) return;
        const size_t len = std::min(per_, bytes_ - off);
        if (streaming_) stream_copy(dst_ + off, src_ + off, len);
        else std::memcpy(dst_ + off, src_ + off, len);
    }
    void run(int i) {
        unsigned long long seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || generation_ != seen; });
                if (stop_) return;
                seen = generation_;
            }
            slice(i);
            std::lock_guard<std::mutex> lk(mu_);
            if (--pending_ == 0) done_.notify_one();
        }
    }
    std::vector<std::thread> workers_;
    std::mutex submit_, mu_;
    std::condition_variable cv_, done_;
    unsigned char* dst_ = nullptr;
    const unsigned char* src_ = nullptr;
    size_t bytes_ = 0, per_ = 0;
    int pending_ = 0;
    unsigned long long generation_ = 0;
    bool stop_ = false;
    bool streaming_ = false;
    bool nt_ = true;
};

// ---- pinned bounce buffers, cached process-wide -----------------------------------
struct Bounce {
    static constexpr int NBUF = 4;
    static constexpr size_t BUF = (size_t)16 << 20;
    unsigned char* base = nullptr;  // NBUF * BUF bytes of pinned memory
    cudaEvent_t done[NBUF] = {};    // DMA that last used buffer i
    int dev = -1;                   // device the events belong to (an event is recorded on streams of its own device only)
    unsigned char* buf(int i) const { return base + (size_t)i * BUF; }
};

inline std::mutex& bounce_mu() {
    static std::mutex m;
    return m;
}
inline std::vector<Bounce>& bounce_free() {
    static std::vector<Bounce> v;
    return v;
}
inline cudaError_t bounce_acquire(Bounce* out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    {
        std::lock_guard<std::mutex> lk(bounce_mu());
        auto& free_list = bounce_free();
        for (size_t i = free_list.size(); i-- > 0;)
            if (free_list[i].dev == dev) {
                *out = free_list[i];
                free_list.erase(free_list.begin() + (long)i);
                return cudaSuccess;
            }
    }
    Bounce b;
    b.dev = dev;
    e = cudaHostAlloc(reinterpret_cast<void**>(&b.base), Bounce::NBUF * Bounce::BUF, cudaHostAllocPortable);
    if (e != cudaSuccess) return e;
    for (int i = 0; i < Bounce::NBUF; ++i) {
        e = cudaEventCreateWithFlags(&b.done[i], cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
    }
    *out = b;
    return cudaSuccess;
}
inline void bounce_release(const Bounce& b) {
    std::lock_guard<std::mutex> lk(bounce_mu());
    bounce_free().push_back(b);
}

// true when the driver does not know `p` as pinned / registered host memory
inline bool is_pageable(const void* p) {
    cudaPointerAttributes a{};
    const cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) {
        cudaGetLastError();  // older drivers report unregistered memory as an error: not one for us
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
}

// slice [host, host + bytes) -> device, through the bounce buffers, on stream `st`.
// Returns with every DMA enqueued (the last NBUF may still be running).
inline cudaError_t staged_copy_in(const Bounce& b, unsigned char* d_dst, const unsigned char* host, size_t bytes,
                                  cudaStream_t st, size_t* issued) {
    for (size_t off = 0; off < bytes; off += Bounce::BUF, ++*issued) {
        const int i = (int)(*issued % Bounce::NBUF);
        const size_t len = std::min(Bounce::BUF, bytes - off);
        if (*issued >= (size_t)Bounce::NBUF) {
            const cudaError_t e = cudaEventSynchronize(b.done[i]);
            if (e != cudaSuccess) return e;
        }
        CopyPool::get().copy(b.buf(i), host + off, len);
        cudaError_t e = cudaMemcpyAsync(d_dst + off, b.buf(i), len, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return e;
        e = cudaEventRecord(b.done[i], st);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// device -> slice through the bounce buffers: DMAs are issued up to NBUF ahead of the
// host copies that empty the buffers, in order.
class StagedCopyOut {
public:
    StagedCopyOut(const Bounce& b, cudaStream_t st) : b_(b), st_(st) {}
    cudaError_t add(unsigned char* host, const unsigned char* d_src, size_t bytes) {
        for (size_t off = 0; off < bytes; off += Bounce::BUF, ++issued_) {
            if (q_.size() == (size_t)Bounce::NBUF) {
                const cudaError_t e = drain_one();
                if (e != cudaSuccess) return e;
            }
            const int i = (int)(issued_ % Bounce::NBUF);
            const size_t len = std::min(Bounce::BUF, bytes - off);
            cudaError_t e = cudaMemcpyAsync(b_.buf(i), d_src + off, len, cudaMemcpyDeviceToHost, st_);
            if (e != cudaSuccess) return e;
            e = cudaEventRecord(b_.done[i], st_);
            if (e != cudaSuccess) return e;
            q_.push_back({i, host + off, len});
        }
        return cudaSuccess;
    }
    cudaError_t finish() {
        while (!q_.empty()) {
            const cudaError_t e = drain_one();
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    }

private:
    struct Pending { int buf; unsigned char* host; size_t len; };
    cudaError_t drain_one() {
        const Pending p = q_.front();
        q_.pop_front();
        const cudaError_t e = cudaEventSynchronize(b_.done[p.buf]);
        if (e != cudaSuccess) return e;
        CopyPool::get().copy(p.host, b_.buf(p.buf), p.len, /*streaming=*/true);
        return cudaSuccess;
    }
    const Bounce& b_;
    cudaStream_t st_;
    std::deque<Pending> q_;
    size_t issued_ = 0;
};

}  // namespace b200
