// parsekernels.cuh -- hand-written sm_100a kernels of the parse step (include/hash.hpp:12-43,
// include/pfparser.hpp:345-350): the trigger scan K1.
//
// The trigger test `wang_hash(2-bit packed last w bases) % p == 0` is a pure function of the w-mer, and a w-mer over
// {A,C,G,T} is one of only 4^w values: for the default w = 10 the whole predicate is a 2^20-bit table = 128 KB, which fits
// the 227 KB of shared memory of a B200 SM.  So the scan never hashes in its inner loop:
//   * k_trigger_lut     builds the table once per (w, p): bit k = [p | wang_hash(k)], division-free (DivTest)
//   * k_trigger_tiles   one persistent CTA per SM keeps the table in shared memory (w <= 10; for 11 <= w <= 14 it stays in
//                       L2) and streams the text through two 16 KB tiles filled by the TMA engine (cp.async.bulk +
//                       mbarrier: the copy of tile t+1 runs while tile t is scanned).  A tile is first packed to 2 bits per
//                       base, 8 bytes at a time with word-wide arithmetic, into shared memory; every thread then owns one
//                       32-position bitmap word: it assembles its 63-base window from the packed tile with shifts, and each
//                       position costs one funnel shift, one shared-memory lookup and one bit insert.
// The reference hashes every position (hash.hpp:29-35); identical trigger sets are what the parity tests check.
// w >= 15 (tables past L2 size) keeps the hashing kernel of pipeline.cuh.
// (included by pipeline.cuh once wang_hash / DivTest / launch_for are defined)
#pragma once

namespace pfb {

#ifndef PFB_EMU
static const int TRG_TILE = 16384;          // text positions per tile = 512 threads x one 32-position word
static const int TRG_THREADS = 512;
static const int TRG_HALO = 32;             // bases kept before a tile (>= w - 1)
static const int TRG_BUF = TRG_TILE + 64;   // bytes of one tile buffer: halo + tile + 16-byte alignment slack on both ends
static const int TRG_LUT_SMEM_W = 10;       // largest w whose table (4^w bits) is kept in shared memory
static const int TRG_LUT_MAX_W = 14;        // largest w served by a table at all (2^28 bits = 32 MB, L2 resident)

__device__ __forceinline__ u32 smem_addr(const void *p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64 *bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// TMA bulk copy global -> shared, completion counted in bytes on the mbarrier (16-byte aligned addresses and size)
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, u32 bytes, u64 *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64 *bar, u32 parity) {
    asm volatile("{\n"
                 ".reg .pred P1;\n"
                 "LAB_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
                 "@P1 bra DONE;\n"
                 "bra LAB_WAIT;\n"
                 "DONE:\n"
                 "}" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}

// 8 text bytes (little endian: base 0 in the low byte) -> 16 bits, base 0 in the top two: the hasher's codes
// A,N -> 0  C -> 1  G -> 2  T,'-' -> 3 (src/utils.c:139-161), word-wide: ((c >> 1) & 3) ^ its own bit 1, N (bits 3 and 6 set) -> 0
__device__ __forceinline__ u32 trg_pack8(u64 x) {
    u64 z = ((x & 0x00ff00ff00ff00ffull) << 8) | ((x >> 8) & 0x00ff00ff00ff00ffull);
    z = ((z & 0x0000ffff0000ffffull) << 16) | ((z >> 16) & 0x0000ffff0000ffffull);
    z = (z << 32) | (z >> 32);
    const u64 ONE = 0x0101010101010101ull;
    u64 y = (z >> 1) & (3 * ONE);
    y ^= (y >> 1) & ONE;
    y &= ~(((z >> 3) & (z >> 6) & ONE) * 3);
    u64 a = (y | (y >> 6)) & 0x000F000F000F000Full;
    a = (a | (a >> 12)) & 0x000000FF000000FFull;
    a = (a | (a >> 24)) & 0xFFFFull;
    return (u32)a;
}

struct TrigArgs {
    const u8 *Tp;               // T'[0]
    u64 n, qmin;                // positions qmin..n may trigger
    int w;
    const u32 *lut; u32 lut_words; int lut_in_smem;
    u32 *trig, *cnt; u64 nwords;
    const u8 *mem_lo, *mem_hi;  // 16-byte aligned bounds of the memory that may be read
    u64 ntiles;
};

__global__ void __launch_bounds__(TRG_THREADS, 1) k_trigger_tiles(TrigArgs a) {
    extern __shared__ __align__(128) u8 trg_smem[];
    u8 *buf0 = trg_smem;                                                  // two tile buffers
    unsigned short *pk = reinterpret_cast<unsigned short *>(trg_smem + 2 * TRG_BUF);   // packed tile: one u16 per 8 bases
    u32 *slut = reinterpret_cast<u32 *>(trg_smem + 2 * TRG_BUF + (TRG_BUF / 8 + 8) * 2);
    __shared__ __align__(8) u64 bar[2];
    __shared__ __align__(8) u64 lbar;
    const int tid = threadIdx.x;
    const u64 kmask = (1ull << (2 * a.w)) - 1;
    // address range of tile j: bases [jT - HALO, (j+1)T) of T', widened to 16-byte alignment and clamped to the buffer
    auto tile_range = [&](u64 j, const u8 *&lo, u32 &bytes) {
        const u8 *s = a.Tp + j * TRG_TILE - TRG_HALO;
        const u8 *e = a.Tp + (j + 1) * TRG_TILE;
        lo = reinterpret_cast<const u8 *>(reinterpret_cast<uintptr_t>(s) & ~(uintptr_t)15);
        const u8 *hi = reinterpret_cast<const u8 *>((reinterpret_cast<uintptr_t>(e) + 15) & ~(uintptr_t)15);
        if (lo < a.mem_lo) lo = a.mem_lo;
        if (hi > a.mem_hi) hi = a.mem_hi;
        bytes = hi > lo ? (u32)(hi - lo) : 0u;
    };
    if (tid == 0) {
        mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_init(&lbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    u64 j = blockIdx.x;
    if (tid == 0) {
        if (a.lut_in_smem) {
            // the whole trigger table -> shared memory, in chunks the TMA engine takes (<= 64 KB each)
            const u32 total = (a.lut_words * 4 + 15u) & ~15u;       // the table's allocation has the slack
            mbar_expect_tx(&lbar, total);
            for (u32 o = 0; o < total; o += 65536u) bulk_g2s(reinterpret_cast<u8 *>(slut) + o, reinterpret_cast<const u8 *>(a.lut) + o, total - o < 65536u ? total - o : 65536u, &lbar);
        }
        if (j < a.ntiles) {
            const u8 *lo; u32 bytes;
            tile_range(j, lo, bytes);
            mbar_expect_tx(&bar[0], bytes);
            if (bytes) bulk_g2s(buf0, lo, bytes, &bar[0]);
        }
    }
    if (a.lut_in_smem) mbar_wait(&lbar, 0);
    const u32 *lut = a.lut_in_smem ? slut : a.lut;
    u32 it = 0;
    for (; j < a.ntiles; j += gridDim.x, ++it) {
        const u32 b = it & 1;
        u8 *buf = buf0 + b * TRG_BUF;
        // the other buffer was last read in phase A of the previous tile, which every thread has left: refill it now, the
        // copy runs while this tile is packed and scanned
        if (tid == 0 && j + gridDim.x < a.ntiles) {
            const u8 *nlo; u32 nbytes;
            tile_range(j + gridDim.x, nlo, nbytes);
            mbar_expect_tx(&bar[b ^ 1], nbytes);
            if (nbytes) bulk_g2s(buf0 + (b ^ 1) * TRG_BUF, nlo, nbytes, &bar[b ^ 1]);
        }
        mbar_wait(&bar[b], (it >> 1) & 1);
        const u8 *lo; u32 bytes;
        tile_range(j, lo, bytes);
        // phase A: 8 bases -> 16 bits for every aligned 8-byte group of the tile
        for (u32 g = tid; g < bytes / 8; g += TRG_THREADS) pk[g] = (unsigned short)trg_pack8(*reinterpret_cast<const u64 *>(buf + 8 * g));
        __syncthreads();
        // phase B: thread = one bitmap word (32 positions q0..q0+31); window = the 32 bases before q0 and the 32 from q0 on
        const u64 t = j * (TRG_TILE / 32) + (u64)tid;
        if (t < a.nwords) {
            const u64 q0 = t * 32;
            const long long o = (long long)((a.Tp + q0 - 32) - lo);          // buffer offset of the first window base (may be < 0 in the very first word)
            const long long g0 = o >> 3;                                      // floor
            const u32 r = (u32)(o & 7);
            const long long ng = (long long)(bytes / 8);
            u64 hi = 0, mid = 0, low = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) { const long long g = g0 + k; hi = (hi << 16) | ((g >= 0 && g < ng) ? (u64)pk[g] : 0ull); }
#pragma unroll
            for (int k = 4; k < 8; ++k) { const long long g = g0 + k; mid = (mid << 16) | ((g >= 0 && g < ng) ? (u64)pk[g] : 0ull); }
            { const long long g = g0 + 8; low = ((g >= 0 && g < ng) ? (u64)pk[g] : 0ull) << 48; }
            const u64 H = r ? (hi << (2 * r)) | (mid >> (64 - 2 * r)) : hi;   // bases q0-32 .. q0-1, first in the top bits
            const u64 M = r ? (mid << (2 * r)) | (low >> (64 - 2 * r)) : mid; // bases q0 .. q0+31
            u32 bits = 0;
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                const int sh = 62 - 2 * k;                                    // the w-mer ending at base q0 + k
                const u64 win = sh ? (M >> sh) | (H << (64 - sh)) : M;
                const u32 km = (u32)(win & kmask);
                bits |= ((lut[km >> 5] >> (km & 31)) & 1u) << k;
            }
            // positions outside [qmin, n] never trigger
            if (q0 < a.qmin) bits &= a.qmin - q0 >= 32 ? 0u : ~0u << (u32)(a.qmin - q0);
            if (q0 + 31 > a.n) bits &= a.n < q0 ? 0u : ~0u >> (u32)(31 - (a.n - q0));
            a.trig[t] = bits;
            a.cnt[t] = (u32)__popc(bits);
        }
        __syncthreads();       // pk is rewritten by the next tile's phase A
    }
}
#endif

// table of the trigger predicate over all 4^w w-mers, cached per (w, p)
struct TriggerLut {
    int w = 0; u64 p = 0;
    DevBuf<u32> bits;
    u32 words = 0;
};

#ifndef PFB_EMU
inline void build_trigger_lut(TriggerLut &L, int w, u64 p, pfbStream_t st) {
    if (L.w == w && L.p == p && L.words) return;
    const u64 nk = 1ull << (2 * w);
    const u64 words = nk < 32 ? 1 : nk / 32;
    L.bits.alloc(words + 4);
    u32 *out = L.bits.get();
    const DivTest dt = DivTest::make(p);
    launch_for(words, st, PFB_LAMBDA(u64 i) {
        u32 b = 0;
        for (u32 k = 0; k < 32; ++k) {
            const u64 km = i * 32 + k;
            if (km < nk && dt.divides(wang_hash(km))) b |= 1u << k;
        }
        out[i] = b;
    });
    L.w = w; L.p = p; L.words = (u32)words;
}

// K1 with the table; false when this w is not served by a table (caller uses k_trigger_words)
inline bool k_trigger_tiled(const u8 *Tp, u64 n, int w, u64 p, u32 *trig, u32 *cnt, u64 nwords, u64 qmin, const u8 *mem_lo, const u8 *mem_hi,
                            TriggerLut &L, pfbStream_t st) {
    static const bool off = getenv("PFB_TRIGGER") && !strcmp(getenv("PFB_TRIGGER"), "hash");      // A/B knob: the hashing kernel
    if (off || w > TRG_LUT_MAX_W || w > TRG_HALO) return false;
    build_trigger_lut(L, w, p, st);
    TrigArgs a;
    a.Tp = Tp; a.n = n; a.qmin = qmin ? qmin : (u64)w + 1; a.w = w;
    a.lut = L.bits.get(); a.lut_words = L.words; a.lut_in_smem = w <= TRG_LUT_SMEM_W ? 1 : 0;
    a.trig = trig; a.cnt = cnt; a.nwords = nwords;
    a.mem_lo = reinterpret_cast<const u8 *>((reinterpret_cast<uintptr_t>(mem_lo) + 15) & ~(uintptr_t)15);
    a.mem_hi = reinterpret_cast<const u8 *>(reinterpret_cast<uintptr_t>(mem_hi) & ~(uintptr_t)15);
    a.ntiles = (nwords * 32 + TRG_TILE - 1) / TRG_TILE;
    size_t smem = 2 * (size_t)TRG_BUF + ((size_t)TRG_BUF / 8 + 8) * 2 + (a.lut_in_smem ? (((size_t)L.words * 4 + 15) & ~(size_t)15) : 0) + 64;
    static size_t smem_set = 0;
    if (smem > smem_set) {
        PFB_CUDA_CHECK(cudaFuncSetAttribute(k_trigger_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    u64 grid = a.ntiles < 148 ? a.ntiles : 148;            // persistent: one CTA per SM
    if (w < TRG_LUT_SMEM_W - 1 && a.ntiles >= 296) grid = 296;          // small tables leave room for two CTAs per SM
    if (grid == 0) return true;
    k_trigger_tiles<<<(unsigned)grid, TRG_THREADS, smem, st>>>(a);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
    return true;
}
#endif

}  // namespace pfb
