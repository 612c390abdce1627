// radix.cuh -- hand-written stable LSD radix sort of (u64 key, u32 or u64 value) pairs, one kernel
// launch per digit ("onesweep": per-tile digit counts are chained through a decoupled
// look-back array, so every pass reads and writes each pair exactly once).
//
//   k_radix_hist_all  digit histograms of ALL passes in one read of the keys (shared-memory counters per CTA)
//   k_radix_pass    one CTA per tile of RX_TILE pairs, tile ids handed out by an atomic
//                   counter (a tile only ever waits for tiles that are already running):
//                     warp-striped loads -> per-warp digit counters in shared memory with
//                     match.any ranking (stable) -> column scan -> look-back per digit ->
//                     pairs reordered through shared memory -> unit-stride runs to HBM.
//   values can be "iota" in the first pass (value = position): the identity permutation is never
//   materialised.
//
// HBM traffic per pass: 12 B read + 12 B written per pair (first pass with iota values: 8 B read)
// + 3 x 8 B x 256 look-back words per 5120-pair tile (5 %); the histogram kernel reads the keys once more (8 B per pair
// for the whole sort).  Fusing the counts into the key kernel was measured slower on B200.
#pragma once

namespace pfb {

#ifndef PFB_EMU
static const int RX_BITS = 8, RX_R = 1 << RX_BITS;
// tile shape: NT threads x IPT pairs, MINB resident CTAs per SM asked of ptxas, MATCH 0 = ballots, 1 = match.any
template <int NT_, int IPT_, int MINB_, int MATCH_, class V_ = unsigned>
struct RxCfg {
    typedef V_ V;                           // value type: u32, or u64 (positions in a dictionary of 2^32 bytes or more)
    static const int NT = NT_, IPT = IPT_, MINB = MINB_, MATCH = MATCH_, NW = NT_ / 32, TILE = NT_ * IPT_;
    static const size_t SMEM = (size_t)NW * RX_R * 4 + (size_t)TILE * 8 + (size_t)TILE * sizeof(V_) + (size_t)RX_R * 8 + (size_t)RX_R * 4 + 16 * 8;
};
typedef RxCfg<256, 20, 3, 0> RxDefault;   // best of the tools/bench_radix.cu sweep on B200 with full-tile specialisation (5120-pair tiles, 3 CTAs/SM)
typedef RxCfg<256, 15, 3, 0, u64> RxWide;   // 64-bit values: 3840-pair tiles keep 3 CTAs/SM within shared memory
static const int RX_MAX_PASSES = 8;

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// digit histograms of ALL passes in one read of the keys (per-CTA shared-memory counters, then one
// global add per non-empty counter)
__global__ void __launch_bounds__(256) k_radix_hist_all(const unsigned long long *__restrict__ keys, unsigned long long n, int begin_bit,
                                                        int end_bit, int npass, unsigned long long *__restrict__ hist) {
    __shared__ unsigned sh[RX_MAX_PASSES * RX_R];
    for (int i = threadIdx.x; i < npass * RX_R; i += 256) sh[i] = 0;
    __syncthreads();
    unsigned long long i = ((unsigned long long)blockIdx.x * 256 + threadIdx.x) * 2;
    const unsigned long long stride = (unsigned long long)gridDim.x * 512;
    for (; i < n; i += stride) {
        unsigned long long k0, k1 = 0;
        const bool two = i + 1 < n;
        if (two) { ulonglong2 kk = *reinterpret_cast<const ulonglong2 *>(keys + i); k0 = kk.x; k1 = kk.y; }
        else k0 = keys[i];
        for (int p = 0; p < npass; ++p) {
            int sh_ = begin_bit + p * RX_BITS, b = end_bit - sh_;
            unsigned m = b >= RX_BITS ? (unsigned)(RX_R - 1) : (1u << b) - 1u;
            atomicAdd(&sh[p * RX_R + ((unsigned)(k0 >> sh_) & m)], 1u);
            if (two) atomicAdd(&sh[p * RX_R + ((unsigned)(k1 >> sh_) & m)], 1u);
        }
    }
    __syncthreads();
    for (int i2 = threadIdx.x; i2 < npass * RX_R; i2 += 256)
        if (sh[i2]) atomicAdd(&hist[i2], (unsigned long long)sh[i2]);
}

template <class V>
struct RadixPassArgs {
    const unsigned long long *kin; const V *vin; unsigned long long *kout; V *vout;
    unsigned long long n;
    int shift; unsigned mask;               // digit of this pass
    const unsigned long long *hist;         // RX_R counts of this digit over all keys
    unsigned long long *lookback;           // ntiles * RX_R words, tag in the top byte
    unsigned *tile_counter;
    unsigned tag;                           // 2*pass+1 = aggregate published, 2*pass+2 = inclusive prefix published
};

// lanes of the warp holding the same 8-bit digit (8 ballots; cheaper than match.any, whose result
// latency showed up as the top stall in the first ncu capture of this kernel)
__device__ __forceinline__ unsigned rx_peers(unsigned d) {
    unsigned peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < RX_BITS; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

// FULL: the tile holds RX_TILE pairs (all but the last tile of a pass): no per-item bounds checks
template <bool IOTA, class C, bool FULL>
__device__ void rx_pass_body(const RadixPassArgs<typename C::V> &a, const unsigned long long tile, const int tile_n_rt) {
    constexpr int RX_NT = C::NT, RX_IPT = C::IPT, RX_NW = C::NW, RX_TILE = C::TILE;
    typedef typename C::V V;
    extern __shared__ __align__(16) unsigned char rx_smem[];
    unsigned *whist = reinterpret_cast<unsigned *>(rx_smem);                                   // [RX_NW][RX_R]
    unsigned long long *skeys = reinterpret_cast<unsigned long long *>(whist + RX_NW * RX_R);  // [RX_TILE]
    V *svals = reinterpret_cast<V *>(skeys + RX_TILE);                                         // [RX_TILE]
    unsigned long long *gbase = reinterpret_cast<unsigned long long *>(svals + RX_TILE);       // [RX_R]
    unsigned *dstart = reinterpret_cast<unsigned *>(gbase + RX_R);                             // [RX_R]
    unsigned long long *wsum = reinterpret_cast<unsigned long long *>(dstart + RX_R);          // [RX_NW]
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const unsigned long long base = tile * RX_TILE;
    const int tile_n = FULL ? RX_TILE : tile_n_rt;

    unsigned long long key[RX_IPT];
    V val[RX_IPT];
    unsigned rank[RX_IPT];
#pragma unroll
    for (int j = 0; j < RX_IPT; ++j) {
        const int li = w * 32 * RX_IPT + j * 32 + lane;
        if (FULL || li < tile_n) {
            key[j] = a.kin[base + li];
            val[j] = IOTA ? (V)(base + li) : a.vin[base + li];
        } else { key[j] = ~0ull; val[j] = 0; }
    }
    // stable ranks inside the warp: items are ordered (j, lane); invalid items of the last tile carry
    // digit mask+1.. (never equal to a real digit because they are excluded from the counters)
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned *wh = whist + w * RX_R;
#pragma unroll
    for (int j = 0; j < RX_IPT; ++j) {
        const int li = w * 32 * RX_IPT + j * 32 + lane;
        const bool valid = FULL || li < tile_n;
        const unsigned d = (unsigned)(key[j] >> a.shift) & a.mask;
        unsigned peers = C::MATCH ? __match_any_sync(0xffffffffu, valid ? d : (unsigned)(RX_R + lane)) : (FULL ? rx_peers(d) : (rx_peers(d) & __ballot_sync(0xffffffffu, valid)));
        const int leader = __ffs((int)peers) - 1;
        unsigned old = 0;
        if (valid && lane == leader) { old = wh[d]; wh[d] = old + (unsigned)__popc(peers); }
        old = __shfl_sync(0xffffffffu, old, leader & 31);
        rank[j] = old + (unsigned)__popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    // digit thread: counts of this tile, exclusive prefix over warps; publish the aggregate early
    unsigned tcount = 0;
    unsigned long long gcount = 0;
    unsigned long long *lb = a.lookback + tile * RX_R + tid;
    if (tid < RX_R) {
        unsigned run = 0;
#pragma unroll
        for (int k = 0; k < RX_NW; ++k) { unsigned c = whist[k * RX_R + tid]; whist[k * RX_R + tid] = run; run += c; }
        tcount = run;
        st_volatile_u64(lb, ((unsigned long long)(tile == 0 ? a.tag + 1 : a.tag) << 56) | tcount);
        gcount = a.hist[tid];
    }
    // exclusive scans over the digits (tile-local starts and global bin bases)
    unsigned inc32 = tcount;
    unsigned long long inc64 = gcount;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
        unsigned o32 = __shfl_up_sync(0xffffffffu, inc32, dd);
        unsigned long long o64 = __shfl_up_sync(0xffffffffu, inc64, dd);
        if (lane >= dd) { inc32 += o32; inc64 += o64; }
    }
    if (tid < RX_R && lane == 31) { wsum[w] = ((unsigned long long)inc32 << 40) | inc64; }   // n < 2^40
    __syncthreads();
    unsigned ds = 0;
    unsigned long long bin_base = 0;
    if (tid < RX_R) {
        unsigned b32 = 0;
        unsigned long long b64 = 0;
        for (int k = 0; k < w; ++k) { b32 += (unsigned)(wsum[k] >> 40); b64 += wsum[k] & ((1ull << 40) - 1); }
        ds = b32 + inc32 - tcount;
        bin_base = b64 + inc64 - gcount;
        dstart[tid] = ds;
    }
    __syncthreads();
    // reorder the tile in shared memory (needs only tile-local offsets) while the preceding tiles
    // publish their counts
#pragma unroll
    for (int j = 0; j < RX_IPT; ++j) {
        const int li = w * 32 * RX_IPT + j * 32 + lane;
        if (FULL || li < tile_n) {
            const unsigned d = (unsigned)(key[j] >> a.shift) & a.mask;
            const unsigned p = dstart[d] + wh[d] + rank[j];
            skeys[p] = key[j];
            svals[p] = val[j];
        }
    }
    if (tid < RX_R) {
        // decoupled look-back over the preceding tiles for digit `tid`
        const unsigned long long TAGV = 1ull << 56;
        unsigned long long excl = 0;
        if (tile != 0) {
            for (long long t = (long long)tile - 1; t >= 0; --t) {
                const unsigned long long *pp = a.lookback + (unsigned long long)t * RX_R + tid;
                unsigned long long v;
                do { v = ld_volatile_u64(pp); } while ((v >> 56) < a.tag);
                excl += v & (TAGV - 1);
                if ((v >> 56) == a.tag + 1) break;
            }
            st_volatile_u64(lb, ((unsigned long long)(a.tag + 1) << 56) | (excl + tcount));
        }
        gbase[tid] = bin_base + excl - ds;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RX_IPT; ++j) {
        const int i = j * RX_NT + tid;
        if (FULL || i < tile_n) {
            const unsigned long long k = skeys[i];
            const unsigned long long g = gbase[(unsigned)(k >> a.shift) & a.mask] + (unsigned long long)i;
            a.kout[g] = k;
            a.vout[g] = svals[i];
        }
    }
}

// the one partial tile of a pass goes through an out-of-line copy so that it does not cost the full tiles registers
template <bool IOTA, class C>
__device__ __noinline__ void rx_pass_tail(const RadixPassArgs<typename C::V> a, const unsigned long long tile, const int tile_n) {
    rx_pass_body<IOTA, C, false>(a, tile, tile_n);
}

template <bool IOTA, class C>
__global__ void __launch_bounds__(C::NT, C::MINB) k_radix_pass(RadixPassArgs<typename C::V> a) {
    extern __shared__ __align__(16) unsigned char rx_smem0[];
    unsigned *whist = reinterpret_cast<unsigned *>(rx_smem0);
    __shared__ unsigned s_tile;
    if (threadIdx.x == 0) s_tile = atomicAdd(a.tile_counter, 1u);
    for (int i = threadIdx.x; i < C::NW * RX_R; i += C::NT) whist[i] = 0;
    __syncthreads();
    const unsigned long long tile = s_tile;
    const unsigned long long left = a.n - tile * C::TILE;
    if (left >= (unsigned long long)C::TILE) rx_pass_body<IOTA, C, true>(a, tile, C::TILE);
    else rx_pass_tail<IOTA, C>(a, tile, (int)left);
}

// ---- host driver ------------------------------------------------------------------------------
struct RadixWork {                 // reusable scratch of the sort (look-back array, histograms, tile counters)
    DevBuf<unsigned long long> lookback, hist;
    DevBuf<unsigned> counters;
    struct Timed { cudaEvent_t a, b; u64 n; };
    std::vector<Timed> pending;        // big passes whose events have not been read yet
    std::vector<cudaEvent_t> pool;
    cudaEvent_t get_event() {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e; cudaEventCreate(&e); return e;
    }
    void collect() {                   // call when the stream is idle
        for (auto &t : pending) {
            float ms = 0;
            if (cudaEventSynchronize(t.b) == cudaSuccess && cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) {
                launch_stats().radix_pass_ms += ms; launch_stats().radix_passes++; launch_stats().radix_pairs += t.n;
            }
            pool.push_back(t.a); pool.push_back(t.b);
        }
        pending.clear();
    }
};

// Stable sort of n pairs on key bits [begin_bit, end_bit).  Result is left in keys_a/vals_a (buffers
// may be swapped).  iota_vals: the input values are 0..n-1 and vals_a need not be initialised.
// hist0: optional device histogram (RX_R u64) of the first digit, already computed by the caller.
template <class C = RxDefault>
inline void radix_sort_pairs(DevBuf<u64> &keys_a, DevBuf<u64> &keys_b, DevBuf<typename C::V> &vals_a, DevBuf<typename C::V> &vals_b, u64 n, int begin_bit,
                             int end_bit, bool iota_vals, const u64 *hist0, RadixWork &rw, pfbStream_t st) {
    typedef typename C::V V;
    if (n == 0 || end_bit <= begin_bit) {
        if (iota_vals && n) { V *v = vals_a.get(); launch_for(n, st, PFB_LAMBDA(u64 i) { v[i] = (V)i; }); }
        return;
    }
    const int npass = (end_bit - begin_bit + RX_BITS - 1) / RX_BITS;
    if (npass > RX_MAX_PASSES) throw Error(-8, "radix sort: too many passes");
    const u64 ntiles = (n + C::TILE - 1) / C::TILE;
    static bool attr_set = false;
    if (!attr_set) {
        PFB_CUDA_CHECK(cudaFuncSetAttribute(k_radix_pass<true, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
        PFB_CUDA_CHECK(cudaFuncSetAttribute(k_radix_pass<false, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
        attr_set = true;
    }
    if (rw.lookback.size() < ntiles * RX_R) rw.lookback.alloc(ntiles * RX_R + ntiles * RX_R / 8);
    if (rw.hist.size() < (size_t)RX_MAX_PASSES * RX_R) rw.hist.alloc((size_t)RX_MAX_PASSES * RX_R);
    if (rw.counters.size() < (size_t)RX_MAX_PASSES) rw.counters.alloc(RX_MAX_PASSES);
    dev_memset(rw.lookback.get(), 0, ntiles * RX_R * 8, st);
    dev_memset(rw.hist.get(), 0, (size_t)RX_MAX_PASSES * RX_R * 8, st);
    dev_memset(rw.counters.get(), 0, RX_MAX_PASSES * 4, st);
    auto dmask = [&](int p) { int b = end_bit - (begin_bit + p * RX_BITS); if (b > RX_BITS) b = RX_BITS; return (unsigned)((1u << b) - 1u); };
    (void)hist0;
    {
        u64 blocks = (n + 512 * 8 - 1) / (512 * 8);
        if (blocks > 148 * 8) blocks = 148 * 8;
        k_radix_hist_all<<<(unsigned)blocks, 256, 0, st>>>((const unsigned long long *)keys_a.get(), n, begin_bit, end_bit, npass,
                                                           (unsigned long long *)rw.hist.get());
        PFB_CUDA_CHECK(cudaGetLastError());
        launch_stats().own++;
    }
    for (int p = 0; p < npass; ++p) {
        RadixPassArgs<V> a;
        a.kin = (const unsigned long long *)keys_a.get(); a.vin = vals_a.get();
        a.kout = (unsigned long long *)keys_b.get(); a.vout = vals_b.get();
        a.n = n; a.shift = begin_bit + p * RX_BITS; a.mask = dmask(p);
        a.hist = (const unsigned long long *)rw.hist.get() + (size_t)p * RX_R;
        a.lookback = (unsigned long long *)rw.lookback.get();
        a.tile_counter = rw.counters.get() + p;
        a.tag = (unsigned)(2 * p + 1);
        const bool timed = n >= (1ull << 24);
        RadixWork::Timed tv;
        if (timed) { tv.a = rw.get_event(); tv.b = rw.get_event(); tv.n = n; cudaEventRecord(tv.a, st); }
        if (p == 0 && iota_vals) k_radix_pass<true, C><<<(unsigned)ntiles, C::NT, C::SMEM, st>>>(a);
        else k_radix_pass<false, C><<<(unsigned)ntiles, C::NT, C::SMEM, st>>>(a);
        if (timed) { cudaEventRecord(tv.b, st); rw.pending.push_back(tv); }
        PFB_CUDA_CHECK(cudaGetLastError());
        launch_stats().own++;
        keys_a.swap(keys_b);
        vals_a.swap(vals_b);
    }
}
#endif

}  // namespace pfb
