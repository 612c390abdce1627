// scan.cuh -- hand-written single-pass prefix scans and reductions (replace cub::DeviceScan / cub::DeviceReduce on the
// product path).  "Decoupled look-back": one launch, every element read once and written once.
//
//   k_scan<TIn, TOut, MAX, INCL>   CTA = 256 threads x 16 consecutive items (4096-item tile, tile ids from an atomic counter
//                                  so that a tile only waits for tiles already running).  A thread reduces its 16 items, the
//                                  CTA scans the 256 partials (warp shuffles), thread 0 publishes the tile aggregate
//                                  (flag 1), walks back over the predecessors' status words until it meets an inclusive
//                                  prefix (flag 2), publishes its own inclusive prefix, and the tile is written out.
//                                  Status word: flag in the top 2 bits, 62-bit value (sums and maxima here stay below 2^62).
//   k_reduce_sum<TIn>              grid-stride sum, one atomicAdd per CTA.
// (included by prim.cuh inside the CUDA build)
#pragma once

namespace pfb {

static const int SCAN_THREADS = 256, SCAN_ITEMS = 16, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

template <class TIn, class TOut, bool MAX, bool INCL>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan(const TIn *__restrict__ in, TOut *__restrict__ out, unsigned long long n,
                                                       unsigned long long *__restrict__ status, unsigned *__restrict__ counter) {
    __shared__ unsigned s_tile;
    __shared__ TOut s_warp[SCAN_THREADS / 32];
    __shared__ TOut s_base;
    __shared__ TOut s_item[SCAN_TILE + SCAN_TILE / SCAN_ITEMS];      // the tile, one pad slot per thread: global accesses are striped (unit stride across the warp), a thread's 16 consecutive items come from here
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
    __syncthreads();
    const unsigned long long tile = s_tile;
    auto op = [](TOut a, TOut b) -> TOut { return MAX ? (a > b ? a : b) : (TOut)(a + b); };
    const unsigned long long tbase = tile * SCAN_TILE;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const int e = k * SCAN_THREADS + tid;
        s_item[e + e / SCAN_ITEMS] = tbase + e < n ? (TOut)in[tbase + e] : (TOut)0;
    }
    __syncthreads();
    TOut v[SCAN_ITEMS];
    TOut acc = 0;                                          // identity of both operations on unsigned values
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        v[k] = s_item[tid * (SCAN_ITEMS + 1) + k];
        acc = op(acc, v[k]);
    }
    // inclusive scan of the thread partials over the CTA
    TOut inc = acc;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        TOut o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc = op(inc, o);
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    TOut wbase = 0, total = 0;
#pragma unroll
    for (int k = 0; k < SCAN_THREADS / 32; ++k) { if (k < wid) wbase = op(wbase, s_warp[k]); total = op(total, s_warp[k]); }
    if (tid == 0) {
        const unsigned long long VMASK = (1ull << 62) - 1;
        unsigned long long excl = 0;
        if (tile == 0) {
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(status), "l"((2ull << 62) | ((unsigned long long)total & VMASK)) : "memory");
        } else {
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(status + tile), "l"((1ull << 62) | ((unsigned long long)total & VMASK)) : "memory");
            for (long long t = (long long)tile - 1; t >= 0; --t) {
                unsigned long long s;
                do { asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(s) : "l"(status + t) : "memory"); } while ((s >> 62) == 0);
                excl = MAX ? (excl > (s & VMASK) ? excl : (s & VMASK)) : excl + (s & VMASK);
                if ((s >> 62) == 2) break;
            }
            const unsigned long long incl = MAX ? (excl > (unsigned long long)total ? excl : (unsigned long long)total) : excl + (unsigned long long)total;
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(status + tile), "l"((2ull << 62) | (incl & VMASK)) : "memory");
        }
        s_base = (TOut)excl;
    }
    __syncthreads();
    TOut run = op(s_base, wbase);
    // exclusive prefix of this thread inside the CTA = partials of the lanes before it
    {
        TOut before = __shfl_up_sync(0xffffffffu, inc, 1);
        if (lane == 0) before = 0;
        run = op(run, before);
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (INCL) { run = op(run, v[k]); s_item[tid * (SCAN_ITEMS + 1) + k] = run; }
        else { s_item[tid * (SCAN_ITEMS + 1) + k] = run; run = op(run, v[k]); }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const int e = k * SCAN_THREADS + tid;
        if (tbase + e < n) out[tbase + e] = s_item[e + e / SCAN_ITEMS];
    }
}

template <class TIn>
__global__ void __launch_bounds__(256) k_reduce_sum(const TIn *__restrict__ in, unsigned long long n, unsigned long long *__restrict__ out) {
    __shared__ unsigned long long ws[8];
    unsigned long long acc = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * 256) acc += (unsigned long long)in[i];
#pragma unroll
    for (int d = 16; d; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int k = 0; k < 8; ++k) t += ws[k];
        if (t) atomicAdd(out, t);
    }
}

// scratch: status words (one per tile) + the tile counter
template <class TIn, class TOut, bool MAX, bool INCL>
inline void device_scan(const TIn *in, TOut *out, uint64_t n, DevBuf<uint8_t> &scratch, cudaStream_t st) {
    if (n == 0) return;
    const uint64_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    const size_t need = (ntiles + 2) * 8;
    if (scratch.size() < need) scratch.alloc(need + need / 4 + 256);
    PFB_CUDA_CHECK(cudaMemsetAsync(scratch.get(), 0, need, st));
    unsigned long long *status = reinterpret_cast<unsigned long long *>(scratch.get()) + 1;
    unsigned *counter = reinterpret_cast<unsigned *>(scratch.get());
    k_scan<TIn, TOut, MAX, INCL><<<(unsigned)ntiles, SCAN_THREADS, 0, st>>>(in, out, n, status, counter);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
}

}  // namespace pfb
