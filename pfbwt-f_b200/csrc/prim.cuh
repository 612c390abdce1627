// prim.cuh -- device primitive layer of pfbwt-b200.
//
// Everything the pipeline does on the GPU goes through this header:
//   * launch_for(n, f)       one logical thread per element, grid-stride kernel
//   * launch_blocks(...)     explicit CTA-shaped kernels (hand-written kernels in
//                            kernels.cuh use raw <<<>>> launches instead)
//   * scans / radix sorts / compaction
//   * DevBuf<T>              RAII device allocation
//
// Two builds exist.  The product build (nvcc, sm_100a) is the only thing that is
// ever linked into libpfbwt_b200.so.  Defining PFB_EMU gives a *developer logic
// check*: the same pipeline source is compiled by g++ with every primitive
// executed as a serial loop, so index arithmetic can be checked against the
// oracle on a machine without a GPU (tests/emu/, `-m "not gpu"`).  The emulator is
// test infrastructure; it is never shipped, never linked into the library and
// never used as a fallback -- the library fails loudly without a CUDA device.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#ifdef PFB_EMU
// ---------------------------------------------------------------- emulator --
#include <algorithm>
#include <numeric>
#define PFB_DEV
#define PFB_HD
#define PFB_LAMBDA [=]
typedef int pfbStream_t;
#else
// -------------------------------------------------------------------- CUDA --
#include <cuda_runtime.h>
#include <map>
#include <mutex>
#include <unordered_map>
#define PFB_DEV __device__
#define PFB_HD __host__ __device__
#define PFB_LAMBDA [=] __device__
typedef cudaStream_t pfbStream_t;
#endif

namespace pfb {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

#ifndef PFB_EMU
#define PFB_CUDA_CHECK(x)                                                                         \
    do {                                                                                          \
        cudaError_t e_ = (x);                                                                     \
        if (e_ != cudaSuccess)                                                                    \
            throw pfb::Error(-5, std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " +  \
                                     __FILE__ + ":" + std::to_string(__LINE__));                  \
    } while (0)
#endif

// launch counter: every kernel launched by this library (ours + CUB's) is counted
// so bench.py can report gpu_launches honestly.
struct LaunchStats {
    uint64_t own = 0;   // kernels written in this repo
    uint64_t lib = 0;   // CUB kernels (library code)
    // radix passes over >= 2^24 pairs, timed with CUDA events on the launching stream
    double radix_pass_ms = 0; uint64_t radix_passes = 0, radix_pairs = 0;
    float run_ms = 0;   // last pfb_run_resident, events on the pipeline stream
};
inline LaunchStats &launch_stats() {
    static LaunchStats s;
    return s;
}

// ------------------------------------------------------------------ DevBuf --
// Device memory is carved out of a few large slabs (cudaMalloc'ed on demand, kept until the
// last context goes away) by a best-fit free list with coalescing, so the many short-lived
// work arrays of a run cost no driver call and no device synchronisation.  Re-use of a
// freed range is ordered by the stream: all work of a context runs on one stream, and a
// call coming from a different stream first synchronises the device (merge_pfp, rare).
size_t &dev_bytes_live();
size_t &dev_bytes_peak();
inline pfbStream_t &alloc_stream() {
    static thread_local pfbStream_t s = 0;
    return s;
}
#ifndef PFB_EMU
struct SlabPool {
    struct Slab { char *base; size_t size; };
    std::mutex mu;
    std::vector<Slab> slabs;
    std::map<char *, size_t> free_;               // start -> bytes, coalesced, never spans slabs
    std::unordered_map<void *, size_t> live;      // ptr -> bytes
    pfbStream_t last_stream = 0;
    bool have_stream = false;
    static size_t round_up(size_t b) { return (b + 511) & ~(size_t)511; }

    void order(pfbStream_t st) {
        if (have_stream && st != last_stream) cudaDeviceSynchronize();
        last_stream = st; have_stream = true;
    }
    void *take(size_t bytes) {   // best fit
        auto best = free_.end();
        for (auto it = free_.begin(); it != free_.end(); ++it)
            if (it->second >= bytes && (best == free_.end() || it->second < best->second)) best = it;
        if (best == free_.end()) return nullptr;
        char *p = best->first;
        size_t sz = best->second;
        free_.erase(best);
        if (sz > bytes) free_[p + bytes] = sz - bytes;
        live[p] = bytes;
        return p;
    }
    bool grow(size_t bytes) {
        size_t fr = 0, tot = 0;
        cudaMemGetInfo(&fr, &tot);
        size_t want = (size_t)8 << 30;
        if (want > fr / 2) want = fr / 2;
        if (want < bytes) want = bytes;
        want = (want + ((size_t)2 << 20) - 1) & ~(((size_t)2 << 20) - 1);
        char *base = nullptr;
        if (cudaMalloc((void **)&base, want) != cudaSuccess) {
            cudaGetLastError();
            want = (bytes + ((size_t)2 << 20) - 1) & ~(((size_t)2 << 20) - 1);
            if (cudaMalloc((void **)&base, want) != cudaSuccess) { cudaGetLastError(); return false; }
        }
        slabs.push_back({base, want});
        free_[base] = want;
        return true;
    }
    void *alloc(size_t bytes, pfbStream_t st) {
        std::lock_guard<std::mutex> lk(mu);
        order(st);
        bytes = round_up(bytes);
        void *p = take(bytes);
        if (p) return p;
        if (!grow(bytes)) {
            cudaDeviceSynchronize();
            trim_locked();            // give empty slabs back and try once more
            if (!grow(bytes)) return nullptr;
        }
        return take(bytes);
    }
    void release(void *p, pfbStream_t st) {
        std::lock_guard<std::mutex> lk(mu);
        order(st);
        auto it = live.find(p);
        if (it == live.end()) return;
        char *b = (char *)p;
        size_t sz = it->second;
        live.erase(it);
        const Slab *sl = nullptr;
        for (auto &s : slabs) if (b >= s.base && b < s.base + s.size) sl = &s;
        auto nx = free_.lower_bound(b);
        if (nx != free_.end() && b + sz == nx->first && sl && nx->first < sl->base + sl->size) { sz += nx->second; nx = free_.erase(nx); }
        if (nx != free_.begin()) {
            auto pv = std::prev(nx);
            if (pv->first + pv->second == b && sl && pv->first >= sl->base) { pv->second += sz; return; }
        }
        free_[b] = sz;
    }
    void trim_locked() {   // cudaFree every slab that is entirely free
        for (size_t i = 0; i < slabs.size();) {
            auto it = free_.find(slabs[i].base);
            if (it != free_.end() && it->second == slabs[i].size) {
                free_.erase(it);
                cudaFree(slabs[i].base);
                slabs.erase(slabs.begin() + (long)i);
            } else ++i;
        }
    }
    void trim() { std::lock_guard<std::mutex> lk(mu); cudaDeviceSynchronize(); trim_locked(); }
    // the slab (one cudaMalloc allocation) a pointer lies in: what CUDA IPC exports to the peers of a multi-GPU job
    bool locate(const void *p, char *&base, size_t &off) {
        std::lock_guard<std::mutex> lk(mu);
        const char *c = (const char *)p;
        for (auto &s : slabs) if (c >= s.base && c < s.base + s.size) { base = s.base; off = (size_t)(c - s.base); return true; }
        return false;
    }
    size_t reserved() { std::lock_guard<std::mutex> lk(mu); size_t t = 0; for (auto &s : slabs) t += s.size; return t; }
    size_t idle() { std::lock_guard<std::mutex> lk(mu); size_t t = 0; for (auto &f : free_) t += f.second; return t; }   // bytes held but not handed out
};
SlabPool &slab_pool();   // one per process (one device per process: one rank per GPU)
#endif

template <class T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    pfbStream_t st_ = 0;
    DevBuf() {}
    explicit DevBuf(size_t count) { alloc(count); }
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    DevBuf(DevBuf &&o) noexcept : p(o.p), n(o.n), st_(o.st_) { o.p = nullptr; o.n = 0; }
    DevBuf &operator=(DevBuf &&o) noexcept {
        if (this != &o) { release(); p = o.p; n = o.n; st_ = o.st_; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    void alloc(size_t count) {
        release();
        n = count;
        size_t bytes = (count ? count : 1) * sizeof(T);
#ifdef PFB_EMU
        p = (T *)malloc(bytes);
        if (!p) throw Error(-4, "emu: out of memory");
#else
        st_ = alloc_stream();
        p = (T *)slab_pool().alloc(bytes, st_);
        if (!p) {
            n = 0;
            throw Error(-4, std::string("device out of memory allocating ") + std::to_string(bytes) + " bytes");
        }
#endif
        dev_bytes_live() += bytes;
        if (dev_bytes_live() > dev_bytes_peak()) dev_bytes_peak() = dev_bytes_live();
    }
    void release() {
        if (p) {
            dev_bytes_live() -= (n ? n : 1) * sizeof(T);
#ifdef PFB_EMU
            free(p);
#else
            slab_pool().release(p, st_);
#endif
        }
        p = nullptr; n = 0;
    }
    T *get() const { return p; }
    size_t size() const { return n; }
    void swap(DevBuf &o) { std::swap(p, o.p); std::swap(n, o.n); std::swap(st_, o.st_); }
};

// ----------------------------------------------------------------- memory ops
inline void dev_memset(void *p, int v, size_t bytes, pfbStream_t s) {
#ifdef PFB_EMU
    (void)s; memset(p, v, bytes);
#else
    PFB_CUDA_CHECK(cudaMemsetAsync(p, v, bytes, s));
#endif
}
inline void h2d(void *d, const void *h, size_t bytes, pfbStream_t s) {
#ifdef PFB_EMU
    (void)s; memcpy(d, h, bytes);
#else
    PFB_CUDA_CHECK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s));
#endif
}
inline void d2h(void *h, const void *d, size_t bytes, pfbStream_t s) {
#ifdef PFB_EMU
    (void)s; memcpy(h, d, bytes);
#else
    PFB_CUDA_CHECK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s));
#endif
}
inline void d2d(void *dst, const void *src, size_t bytes, pfbStream_t s) {
#ifdef PFB_EMU
    (void)s; memmove(dst, src, bytes);
#else
    PFB_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s));
#endif
}
inline void stream_sync(pfbStream_t s) {
#ifdef PFB_EMU
    (void)s;
#else
    PFB_CUDA_CHECK(cudaStreamSynchronize(s));
#endif
}
template <class T>
inline T read_scalar(const T *d, pfbStream_t s) {
    T v;
    d2h(&v, d, sizeof(T), s);
    stream_sync(s);
    return v;
}

// ------------------------------------------------------------------ atomics --
#ifdef PFB_EMU
template <class T> inline T atomic_add(T *a, T v) { T o = *a; *a = o + v; return o; }
template <class T> inline T atomic_min(T *a, T v) { T o = *a; if (v < o) *a = v; return o; }
template <class T> inline T atomic_max(T *a, T v) { T o = *a; if (v > o) *a = v; return o; }
template <class T> inline T atomic_or(T *a, T v) { T o = *a; *a = o | v; return o; }
template <class T> inline T atomic_cas(T *a, T cmp, T v) { T o = *a; if (o == cmp) *a = v; return o; }
inline int popc32(uint32_t x) { return __builtin_popcount(x); }
inline int ffs32(uint32_t x) { return __builtin_ffs((int)x); }
inline int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }
inline int popc64(uint64_t x) { return __builtin_popcountll(x); }
#else
__device__ __forceinline__ uint32_t atomic_add(uint32_t *a, uint32_t v) { return atomicAdd(a, v); }
__device__ __forceinline__ uint64_t atomic_add(uint64_t *a, uint64_t v) {
    return (uint64_t)atomicAdd((unsigned long long *)a, (unsigned long long)v);
}
__device__ __forceinline__ uint32_t atomic_min(uint32_t *a, uint32_t v) { return atomicMin(a, v); }
__device__ __forceinline__ uint64_t atomic_min(uint64_t *a, uint64_t v) {
    return (uint64_t)atomicMin((unsigned long long *)a, (unsigned long long)v);
}
__device__ __forceinline__ uint32_t atomic_max(uint32_t *a, uint32_t v) { return atomicMax(a, v); }
__device__ __forceinline__ uint64_t atomic_max(uint64_t *a, uint64_t v) {
    return (uint64_t)atomicMax((unsigned long long *)a, (unsigned long long)v);
}
__device__ __forceinline__ uint32_t atomic_or(uint32_t *a, uint32_t v) { return atomicOr(a, v); }
__device__ __forceinline__ uint64_t atomic_cas(uint64_t *a, uint64_t cmp, uint64_t v) {
    return (uint64_t)atomicCAS((unsigned long long *)a, (unsigned long long)cmp, (unsigned long long)v);
}
__device__ __forceinline__ int popc32(uint32_t x) { return __popc(x); }
__device__ __forceinline__ int ffs32(uint32_t x) { return __ffs((int)x); }
__host__ __device__ __forceinline__ int popc64(uint64_t x) {
#ifdef __CUDA_ARCH__
    return __popcll((unsigned long long)x);
#else
    return __builtin_popcountll(x);
#endif
}
__host__ __device__ __forceinline__ int clz64(uint64_t x) {
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
#endif

// -------------------------------------------------------------- launch_for --
#ifndef PFB_EMU
template <class F>
__global__ void __launch_bounds__(256) for_kernel(uint64_t n, F f) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) f(i);
}
#endif

// one logical thread per i in [0,n).  Grid is capped at a multiple of the SM
// count (148 SMs x 8 resident CTAs of 256 threads) and grid-strided beyond it.
template <class F>
inline void launch_for(uint64_t n, pfbStream_t s, F f) {
    if (n == 0) return;
#ifdef PFB_EMU
    (void)s;
    for (uint64_t i = 0; i < n; ++i) f(i);
#else
    uint64_t blocks = (n + 255) / 256;
    const uint64_t cap = 148ull * 8ull * 4ull;
    if (blocks > cap) blocks = cap;
    for_kernel<<<(unsigned)blocks, 256, 0, s>>>(n, f);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
#endif
}

typedef uint8_t u8;
typedef uint32_t u32;
typedef uint64_t u64;
}  // namespace pfb
#include "radix.cuh"
#ifndef PFB_EMU
#include "scan.cuh"
#endif
namespace pfb {

// ------------------------------------------------------------------- scans --
struct Scratch {
    DevBuf<uint8_t> buf;
    DevBuf<uint8_t> scanbuf;     // status words of the look-back scans
#ifndef PFB_EMU
    RadixWork rw;
#endif
    void *ensure(size_t bytes) {
        if (buf.size() < bytes) buf.alloc(bytes + bytes / 4 + 256);
        return buf.get();
    }
};

struct MaxOp {
    template <class T> PFB_HD T operator()(const T &a, const T &b) const { return a > b ? a : b; }
};
template <class A, class B>
struct CastOp {
    PFB_HD B operator()(const A &a) const { return (B)a; }
};

template <class TIn, class TOut>
inline void exclusive_sum(const TIn *in, TOut *out, uint64_t n, Scratch &sc, pfbStream_t s) {
    if (n == 0) return;
#ifdef PFB_EMU
    (void)sc; (void)s;
    TOut acc = 0;
    for (uint64_t i = 0; i < n; ++i) { TOut v = (TOut)in[i]; out[i] = acc; acc += v; }
#else
    device_scan<TIn, TOut, false, false>(in, out, n, sc.scanbuf, s);          // scan.cuh: single pass, decoupled look-back
#endif
}
template <class T>
inline void inclusive_max(const T *in, T *out, uint64_t n, Scratch &sc, pfbStream_t s) {
    if (n == 0) return;
#ifdef PFB_EMU
    (void)sc; (void)s;
    T acc = in[0];
    for (uint64_t i = 0; i < n; ++i) { acc = in[i] > acc ? in[i] : acc; out[i] = acc; }
#else
    device_scan<T, T, true, true>(in, out, n, sc.scanbuf, s);
#endif
}

template <class TIn, class TOut>
inline TOut reduce_sum(const TIn *in, uint64_t n, Scratch &sc, pfbStream_t s) {
    if (n == 0) return 0;
#ifdef PFB_EMU
    (void)sc; (void)s;
    TOut acc = 0;
    for (uint64_t i = 0; i < n; ++i) acc += (TOut)in[i];
    return acc;
#else
    (void)sc;
    DevBuf<unsigned long long> out(1);
    PFB_CUDA_CHECK(cudaMemsetAsync(out.get(), 0, 8, s));
    uint64_t blocks = (n + 256 * 16 - 1) / (256 * 16);
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_reduce_sum<TIn><<<(unsigned)blocks, 256, 0, s>>>(in, n, out.get());
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
    return (TOut)read_scalar(out.get(), s);
#endif
}

// ------------------------------------------------------------- radix sorts --
// Stable LSD radix sort of (key,value) pairs on key bits [begin_bit,end_bit).
// Result is left in keys_a/vals_a (buffers may be swapped internally).
// Pairs go through the hand-written onesweep sort of radix.cuh (the emulator uses std::stable_sort).  iota_vals: input values are 0..n-1 (vals_a need not be initialised);
// hist0: device histogram of the first 8-bit digit when the caller already has it.
template <class K, class V>
inline void sort_pairs(DevBuf<K> &keys_a, DevBuf<K> &keys_b, DevBuf<V> &vals_a, DevBuf<V> &vals_b, uint64_t n,
                       int begin_bit, int end_bit, Scratch &sc, pfbStream_t s, bool iota_vals = false, const uint64_t *hist0 = nullptr) {
#ifndef PFB_EMU
    // the product build sorts (u64 key, u32 / u64 value) pairs with the hand-written onesweep sort of radix.cuh
    static_assert(sizeof(K) == 8 && (sizeof(V) == 4 || sizeof(V) == 8), "sort_pairs: (u64, u32) and (u64, u64) pairs only");
    radix_sort_pairs<typename std::conditional<sizeof(V) == 4, RxDefault, RxWide>::type>(keys_a, keys_b, vals_a, vals_b, n, begin_bit, end_bit, iota_vals, hist0, sc.rw, s);
    return;
#else
    if (iota_vals && n) { V *v = vals_a.get(); launch_for(n, s, PFB_LAMBDA(uint64_t i) { v[i] = (V)i; }); }
    if (n <= 1) return;
    if (end_bit <= begin_bit) return;
    (void)sc; (void)s; (void)keys_b; (void)vals_b;
    std::vector<uint64_t> perm(n);
    std::iota(perm.begin(), perm.end(), 0);
    const K mask = (end_bit - begin_bit >= (int)(8 * sizeof(K))) ? (K)~(K)0 : (K)((((K)1) << (end_bit - begin_bit)) - 1);
    K *kp = keys_a.get(); V *vp = vals_a.get();
    std::stable_sort(perm.begin(), perm.end(), [&](uint64_t a, uint64_t b) {
        return ((kp[a] >> begin_bit) & mask) < ((kp[b] >> begin_bit) & mask);
    });
    std::vector<K> k2(n); std::vector<V> v2(n);
    for (uint64_t i = 0; i < n; ++i) { k2[i] = kp[perm[i]]; v2[i] = vp[perm[i]]; }
    memcpy(kp, k2.data(), n * sizeof(K)); memcpy(vp, v2.data(), n * sizeof(V));
#endif
}

inline int bits_for(uint64_t max_value) {   // number of bits needed to represent max_value
    int b = 1;
    while (b < 64 && (max_value >> b)) ++b;
    return b;
}

}  // namespace pfb
