// emit.cuh -- hand-written kernels of the BWT assembly (K7 classify / groups, K8 emission).
// Every kernel works on TILES of 2048 sorted dictionary suffixes (256 threads x 8 consecutive
// elements) so that per-element arrays are read and written with 16/32-byte vector accesses, and a
// tile's output range is contiguous in the BWT: offsets are a block scan on top of one 64-bit
// base per tile (no per-element offset array), output bytes / SA entries are staged in shared
// memory and leave as aligned 16-byte stores.
//
// meta[t] (u8) of the t-th smallest dictionary suffix:
//   bits 0..3  code of the preceding byte (dictkeys.cuh: prev_code / prev_byte)
//   bit  4     head: first member of a group of identical suffix strings
//   bit  5     full: the suffix is a whole dictionary word (include/pfbwt.hpp:116)
//   bit  6     valid: suffix length > w (include/pfbwt.hpp:114)
//   bit  7     hard: member of a group that is merged by ilist position (include/pfbwt.hpp:146-178);
//              during the suffix sort the bit marks "tied and still open" instead
// Under PFB_EMU (tests/emu, developer logic check) each kernel has a serial twin with the same
// contract.
#pragma once

namespace pfb {

static const int TILE = 2048;
static const u8 M_HEAD = 0x10, M_FULL = 0x20, M_VALID = 0x40, M_HARD = 0x80;

inline u64 tiles_of(u64 n) { return (n + TILE - 1) / TILE; }
inline u64 padded(u64 n) { return tiles_of(n) * TILE; }

template <class DP>
struct FlagsArgsT {
    const u64 *key; const DP *sa; u64 D; u8 *meta; u32 *cnt; u32 *tilecnt; KeyExtra kx; WordLookupT<DP> wl; u64 w;
};
// valid / full bits and occurrence count of dictionary suffix idx by position -> word lookup
// (include/pfbwt.hpp:83-85,111-119)
template <class DP>
PFB_HD inline u32 suffix_class(const WordLookupT<DP> &wl, DP idx, u64 w, u32 &cnt, u32 &ibase, u32 &slen) {
    u32 d = wl.word(idx);
    u64 len = wl.suflen(d, idx);
    cnt = 0; ibase = 0; slen = (u32)len;
    if (len <= w) return 0;
    u32 c0 = wl.cum(d);
    cnt = wl.cum(d + 1) - c0;
    ibase = 1 + c0;
    return (u32)M_VALID | (idx == wl.start(d) ? (u32)M_FULL : 0u);
}
template <class DP>
PFB_HD inline void flags_one(const FlagsArgsT<DP> &a, u64 t, u64 kprev, u64 k, u64 knext, u32 &m, u32 &c, bool &act) {
    u64 code = a.kx.code(k);
    bool head = t == 0 || code != a.kx.code(kprev);
    bool last = t + 1 == a.D || code != a.kx.code(knext);
    act = !(head && last) && !key_closed(k);
    u64 cf = a.kx.count(k);
    m = (u32)(k & 15) | (head ? M_HEAD : 0) | (act ? M_HARD : 0);
    if (a.kx.cmax && cf != a.kx.cmax) { c = (u32)cf; m |= (c ? M_VALID : 0) | (a.kx.full(k) ? M_FULL : 0); }
    else { u32 ib, sl; m |= suffix_class<DP>(a.wl, a.sa[t], a.w, c, ib, sl); }      // no room in the key / count overflow
}


#ifndef PFB_EMU
// exclusive prefix of one value per thread over a 256-thread block; returns the block total
template <class T>
__device__ __forceinline__ T block_excl_scan(T v, T &total, T *warp_sums /* 8 */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    T inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) warp_sums[wid] = inc;
    __syncthreads();
    T base = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { T s = warp_sums[k]; if (k < wid) base += s; tot += s; }
    __syncthreads();
    total = tot;
    return base + inc - v;
}

// ---- after the main radix sort: head flags, "tied and open" flags, per-tile open counts; the
//      valid / full bits and the occurrence count come out of the key's upper bits (dictkeys.cuh) ----
template <class DP>
__global__ void __launch_bounds__(256) k_sorted_flags_kernel(FlagsArgsT<DP> a) {
    __shared__ u32 ws[8];
    const u64 t0 = (u64)blockIdx.x * TILE + (u64)threadIdx.x * 8;
    u64 k[10];
#pragma unroll
    for (int e = 0; e < 10; ++e) {
        u64 t = t0 + (u64)e - 1;
        k[e] = (t0 + (u64)e >= 1 && t < a.D) ? a.key[t] : ~0ull;
    }
    u64 m8 = 0;
    u32 nact = 0, c[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        u64 t = t0 + (u64)e;
        u32 m = M_HEAD;
        c[e] = 0;
        if (t < a.D) { bool act; flags_one(a, t, k[e], k[e + 1], k[e + 2], m, c[e], act); nact += act; }
        m8 |= (u64)m << (8 * e);
    }
    *reinterpret_cast<u64 *>(a.meta + t0) = m8;
    reinterpret_cast<uint4 *>(a.cnt + t0)[0] = make_uint4(c[0], c[1], c[2], c[3]);
    reinterpret_cast<uint4 *>(a.cnt + t0)[1] = make_uint4(c[4], c[5], c[6], c[7]);
    u32 tot;
    block_excl_scan<u32>(nact, tot, ws);
    if (threadIdx.x == 0) a.tilecnt[blockIdx.x] = tot;
}

// per-tile sums of cnt[] (BWT positions emitted by a tile)
__global__ void __launch_bounds__(256) k_tile_sums_kernel(const u32 *__restrict__ cnt, u64 *__restrict__ tilesum) {
    __shared__ u64 ws[8];
    const u64 t0 = (u64)blockIdx.x * TILE + (u64)threadIdx.x * 8;
    uint4 x = reinterpret_cast<const uint4 *>(cnt + t0)[0], y = reinterpret_cast<const uint4 *>(cnt + t0)[1];
    u64 sum = (u64)x.x + x.y + x.z + x.w + y.x + y.y + y.z + y.w, tot;
    block_excl_scan<u64>(sum, tot, ws);
    if (threadIdx.x == 0) tilesum[blockIdx.x] = tot;
}

// ---- ordered compaction of the positions whose meta byte has `bit` set ----
__global__ void __launch_bounds__(256) k_compact_bit_kernel(const u8 *__restrict__ meta, u8 bit, const u32 *__restrict__ tilebase,
                                                            u32 *__restrict__ list) {
    __shared__ u32 ws[8];
    const u64 t0 = (u64)blockIdx.x * TILE + (u64)threadIdx.x * 8;
    u64 m8 = *reinterpret_cast<const u64 *>(meta + t0);
    u32 c = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) c += ((m8 >> (8 * e)) & bit) ? 1u : 0u;
    u32 tot;
    u32 o = tilebase[blockIdx.x] + block_excl_scan<u32>(c, tot, ws);
#pragma unroll
    for (int e = 0; e < 8; ++e)
        if ((m8 >> (8 * e)) & bit) list[o++] = (u32)(t0 + (u64)e);
}
#endif


// meta (pc | head | full | valid | open bit) and cnt from sorted keys; returns the ordered list of open positions
template <class DP>
inline u64 sorted_flags_and_open_list(FlagsArgsT<DP> fa, DevBuf<u32> &list, Workspace &ws) {
    pfbStream_t st = ws.st;
    const u64 D = fa.D, nt = tiles_of(D);
#ifdef PFB_EMU
    std::vector<u32> v;
    for (u64 t = 0; t < padded(D); ++t) {
        if (t >= D) { fa.meta[t] = M_HEAD; fa.cnt[t] = 0; continue; }
        u32 m, c; bool act;
        flags_one(fa, t, t ? fa.key[t - 1] : 0, fa.key[t], t + 1 < D ? fa.key[t + 1] : 0, m, c, act);
        fa.meta[t] = (u8)m; fa.cnt[t] = c;
        if (act) v.push_back((u32)t);
    }
    list.alloc(v.size());
    if (!v.empty()) memcpy(list.get(), v.data(), v.size() * 4);
    (void)st; (void)nt;
    return v.size();
#else
    DevBuf<u32> tcnt(nt + 1), tbase(nt + 1);
    fa.tilecnt = tcnt.get();
    k_sorted_flags_kernel<DP><<<(unsigned)nt, 256, 0, st>>>(fa);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
    dev_memset(tcnt.get() + nt, 0, 4, st);
    exclusive_sum<u32, u32>(tcnt.get(), tbase.get(), nt + 1, ws.sc, st);
    u64 m = read_scalar(tbase.get() + nt, st);
    list.alloc(m);
    if (m) {
        k_compact_bit_kernel<<<(unsigned)nt, 256, 0, st>>>(fa.meta, M_HARD, tbase.get(), list.get());
        PFB_CUDA_CHECK(cudaGetLastError());
        launch_stats().own++;
    }
    return m;
#endif
}

inline void k_tile_sums(const u32 *cnt, u64 D, u64 *tilesum, pfbStream_t st) {
#ifdef PFB_EMU
    (void)st;
    for (u64 tile = 0; tile < tiles_of(D); ++tile) {
        u64 s = 0;
        for (u64 t = tile * TILE; t < (tile + 1) * TILE; ++t) s += cnt[t];
        tilesum[tile] = s;
    }
#else
    k_tile_sums_kernel<<<(unsigned)tiles_of(D), 256, 0, st>>>(cnt, tilesum);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
#endif
}

// ---------------------------------------------------------------------------
// K7a classify: position -> word through the sampled table; valid / full bits, number of
// occurrences the suffix emits (occ of its word, 0 when suffix length <= w), per-tile sums.
// With SA output also the first ilist slot of the word and the suffix length, so that the
// emission kernels need no second lookup.        (include/pfbwt.hpp:83-85,111-119)
// ---------------------------------------------------------------------------
template <class DP>
struct ClassifyArgsT {
    const DP *sa; u8 *meta; WordLookupT<DP> wl; u64 D; u64 w;
    u32 *cnt; u64 *tilesum; u32 *ibase; u32 *sl;   // ibase/sl only when want_sa
};

template <class DP>
PFB_HD inline void classify_one(const ClassifyArgsT<DP> &a, u64 t, u32 &m, u32 &c, u32 &ib, u32 &s) {
    m = (m & (u32)(0x0f | M_HEAD)) | suffix_class<DP>(a.wl, a.sa[t], a.w, c, ib, s);
}

#ifndef PFB_EMU
template <bool SA, class DP>
__global__ void __launch_bounds__(256) k_classify_kernel(ClassifyArgsT<DP> a) {
    __shared__ u64 ws[8];
    const u64 t0 = (u64)blockIdx.x * TILE + (u64)threadIdx.x * 8;
    u64 m8 = *reinterpret_cast<const u64 *>(a.meta + t0);
    u32 c[8], ib[8], s[8];
    u64 sum = 0, o8 = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        u64 t = t0 + (u64)e;
        u32 m = (u32)(m8 >> (8 * e)) & 0xffu;
        c[e] = 0; ib[e] = 0; s[e] = 0;
        if (t < a.D) classify_one(a, t, m, c[e], ib[e], s[e]);
        else m = M_HEAD;
        o8 |= (u64)m << (8 * e);
        sum += c[e];
    }
    *reinterpret_cast<u64 *>(a.meta + t0) = o8;
    reinterpret_cast<uint4 *>(a.cnt + t0)[0] = make_uint4(c[0], c[1], c[2], c[3]);
    reinterpret_cast<uint4 *>(a.cnt + t0)[1] = make_uint4(c[4], c[5], c[6], c[7]);
    if (SA) {
        reinterpret_cast<uint4 *>(a.ibase + t0)[0] = make_uint4(ib[0], ib[1], ib[2], ib[3]);
        reinterpret_cast<uint4 *>(a.ibase + t0)[1] = make_uint4(ib[4], ib[5], ib[6], ib[7]);
        reinterpret_cast<uint4 *>(a.sl + t0)[0] = make_uint4(s[0], s[1], s[2], s[3]);
        reinterpret_cast<uint4 *>(a.sl + t0)[1] = make_uint4(s[4], s[5], s[6], s[7]);
    }
    u64 tot;
    block_excl_scan<u64>(sum, tot, ws);
    if (threadIdx.x == 0) a.tilesum[blockIdx.x] = tot;
}
#endif

template <class DP>
inline void k_classify(const ClassifyArgsT<DP> &a, bool want_sa, pfbStream_t st) {
    const u64 nt = tiles_of(a.D);
#ifdef PFB_EMU
    (void)st;
    for (u64 tile = 0; tile < nt; ++tile) {
        u64 sum = 0;
        for (u64 t = tile * TILE; t < (tile + 1) * TILE; ++t) {
            u32 m = a.meta[t], c = 0, ib = 0, s = 0;
            if (t < a.D) classify_one(a, t, m, c, ib, s); else m = M_HEAD;
            a.meta[t] = (u8)m; a.cnt[t] = c;
            if (want_sa) { a.ibase[t] = ib; a.sl[t] = s; }
            sum += c;
        }
        a.tilesum[tile] = sum;
    }
#else
    if (want_sa) k_classify_kernel<true, DP><<<(unsigned)nt, 256, 0, st>>>(a);
    else k_classify_kernel<false, DP><<<(unsigned)nt, 256, 0, st>>>(a);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
#endif
}

// ---------------------------------------------------------------------------
// K7b hard groups without SA output: a group of >1 identical suffixes is merged by ilist position
// only when its members are preceded by different bytes (include/pfbwt.hpp:146); all equal = the
// members emit independently.  The head of each valid group walks its members; groups longer
// than WALK_CAP go to a list handled by one CTA each.  With SA output every multi-member group
// is hard (include/pfbwt.hpp:148) and no walk is needed.
// ---------------------------------------------------------------------------
static const u32 WALK_CAP = 1024;

#ifndef PFB_EMU
// 8 meta bytes per thread: nearly every suffix is a singleton group, found with a few word-wide bit operations;
// only the heads of real multi-member groups walk their members
__global__ void __launch_bounds__(256) k_group_walk_kernel(u8 *meta, u64 D, u32 *longlist, u32 *nlong, u32 *nhard) {
    const u64 nwords = (D + 7) / 8;
    const u64 stride = (u64)gridDim.x * 256;
    const u64 ones = 0x0101010101010101ull;
    for (u64 wd = (u64)blockIdx.x * 256 + threadIdx.x; wd < nwords; wd += stride) {
        const u64 t0 = wd * 8;
        const u64 m8 = *reinterpret_cast<const u64 *>(meta + t0);                 // meta is padded with head bytes past D
        const u64 nx8 = (m8 >> 8) | ((u64)meta[t0 + 8] << 56);                     // the following element of each
        u64 cand = (m8 >> 4) & (m8 >> 6) & ~(nx8 >> 4) & ones;                     // head & valid & next is not a head
        while (cand) {
            const int el = (__ffsll((long long)cand) - 1) >> 3;
            cand &= cand - 1;
            const u64 t = t0 + (u64)el;
            if (t >= D) break;
            const u32 m = (u32)(m8 >> (8 * el)) & 0xffu;
            bool diff = false;
            u64 e = t + 1;
            for (; e < D && e - t <= WALK_CAP; ++e) {
                u32 x = meta[e];
                if (x & M_HEAD) break;
                diff |= (x & 15) != (m & 15);
            }
            if (e < D && e - t > WALK_CAP) { longlist[atomicAdd(nlong, 1u)] = (u32)t; continue; }
            if (diff) { for (u64 i = t; i < e; ++i) meta[i] |= M_HARD; atomicAdd(nhard, 1u); }
        }
    }
}
__global__ void __launch_bounds__(256) k_group_long_kernel(u8 *meta, u64 D, const u32 *longlist, u32 *nhard) {
    const u64 t = longlist[blockIdx.x];
    const u32 pc = meta[t] & 15;
    __shared__ u64 s_end;
    if (threadIdx.x == 0) s_end = D;
    __syncthreads();
    int any = 0;
    // find the end of the group and whether any preceding byte differs
    for (u64 base = t + 1; base < D; base += 256) {
        u64 i = base + threadIdx.x;
        u32 x = i < D ? meta[i] : (u32)M_HEAD;
        if (x & M_HEAD) atomicMin((unsigned long long *)&s_end, (unsigned long long)i);
        __syncthreads();
        u64 end = s_end;
        __syncthreads();
        if (i < end && (x & 15) != pc) any = 1;
        if (end < D || base + 256 >= D) break;
    }
    any = __syncthreads_or(any);
    if (!any) return;
    if (threadIdx.x == 0) atomicAdd(nhard, 1u);
    const u64 end = s_end;
    for (u64 i = t + threadIdx.x; i < end; i += 256) meta[i] |= M_HARD;
}
#endif

// returns the number of hard groups
inline u64 k_group_classes(u8 *meta, u64 D, bool any_sa, Workspace &ws) {
    pfbStream_t st = ws.st;
    DevBuf<u32> nhard(1);
    dev_memset(nhard.get(), 0, 4, st);
    u32 *nh = nhard.get();
    if (any_sa) {
        launch_for(D, st, PFB_LAMBDA(u64 t) {
            u32 m = meta[t];
            if (!(m & M_VALID)) return;
            bool single = (m & M_HEAD) && (meta[t + 1] & M_HEAD);
            if (!single) { meta[t] = (u8)(m | M_HARD); if (m & M_HEAD) atomic_add(nh, 1u); }
        });
        return read_scalar(nh, st);
    }
#ifdef PFB_EMU
    for (u64 t = 0; t < D; ++t) {
        u32 m = meta[t];
        if ((m & (M_HEAD | M_VALID)) != (M_HEAD | M_VALID)) continue;
        u64 e = t + 1;
        bool diff = false;
        for (; e < D && !(meta[e] & M_HEAD); ++e) diff |= (meta[e] & 15) != (m & 15);
        if (diff) { for (u64 i = t; i < e; ++i) meta[i] |= M_HARD; ++*nh; }
    }
    return *nh;
#else
    DevBuf<u32> longlist(D / WALK_CAP + 2), nlong(1);
    dev_memset(nlong.get(), 0, 4, st);
    k_group_walk_kernel<<<148 * 8, 256, 0, st>>>(meta, D, longlist.get(), nlong.get(), nh);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
    u32 nl = read_scalar(nlong.get(), st);
    if (nl) {
        k_group_long_kernel<<<nl, 256, 0, st>>>(meta, D, longlist.get(), nh);
        PFB_CUDA_CHECK(cudaGetLastError());
        launch_stats().own++;
    }
    return read_scalar(nh, st);
#endif
}

// ---------------------------------------------------------------------------
// K8 emission (include/pfbwt.hpp:111-178 + the writer of src/pfbwt-f.cpp:298-303).
// ---------------------------------------------------------------------------
template <int U> struct UIntOf { typedef u32 type; };
template <> struct UIntOf<8> { typedef u64 type; };

struct EmitArgs {
    const u8 *meta; const u32 *cnt; const u32 *sa; const u32 *ibase; const u32 *sl;   // ibase/sl only with SA output
    WordLookup wl; const u64 *tilebase; u64 D;
    const u32 *ilist; const u8 *bwlast; const u64 *bwsai;
    u8 *bwt; u8 *sa_out;
    u32 *hard_t; u64 *hard_x;          // heads of hard groups and their first output position
    u32 *n_hard_groups; u64 *n_hard_occ;
    u32 small_lim, mid_lim;            // <= HARD_SMALL / HARD_MID (test knob PFB_HARD_LIMITS lowers them)
};

// first ilist slot of the word that dictionary suffix t belongs to.  a.sa == nullptr: the positions are 64-bit (dictionary of
// 2^32 bytes or more); the classify pass has then filled ibase[] also without SA output and no lookup happens here
PFB_HD inline u32 ilist_base(const EmitArgs &a, u64 t, bool sa_mode) {
    return (sa_mode || !a.sa) ? a.ibase[t] : 1u + a.wl.cum(a.wl.word(a.sa[t]));
}

#ifndef PFB_EMU
static const int EMIT_BIGCAP = 128;   // elements of a tile with > 32 occurrences are expanded by the whole CTA

template <int U, bool SA>
__device__ __forceinline__ void emit_slot(const EmitArgs &a, u32 m, u32 ib, u32 sl, u64 j, int pos, u8 *obwt,
                                          typename UIntOf<U>::type *osa) {
    u8 byte = prev_byte(m);
    if (SA || (m & M_FULL)) {
        u32 ii = a.ilist[(u64)ib + j];
        if (m & M_FULL) byte = a.bwlast[ii];                      // include/pfbwt.hpp:119
        if (SA) osa[pos] = (typename UIntOf<U>::type)(a.bwsai[ii] - (u64)sl);
    }
    obwt[pos] = byte;
}

template <int U, bool SA>
__global__ void __launch_bounds__(256) k_emit_kernel(EmitArgs a) {
    typedef typename UIntOf<U>::type UT;
    constexpr int WIN = SA ? 4096 : 8192;
    __shared__ __align__(16) u8 obwt[WIN];
    __shared__ __align__(16) UT osa[SA ? WIN : 2];
    __shared__ u64 ws64[8];
    __shared__ u32 ws32[8];
    __shared__ u32 s_hbase, s_nbig;
    __shared__ u64 big_o[EMIT_BIGCAP];
    __shared__ u32 big_c[EMIT_BIGCAP], big_ib[EMIT_BIGCAP], big_sl[EMIT_BIGCAP], big_m[EMIT_BIGCAP];
    const int tid = threadIdx.x;
    const u64 t0 = (u64)blockIdx.x * TILE + (u64)tid * 8;
    if (tid == 0) s_nbig = 0;
    const u64 m8 = *reinterpret_cast<const u64 *>(a.meta + t0);
    u32 c[8];
    {
        uint4 x = reinterpret_cast<const uint4 *>(a.cnt + t0)[0], y = reinterpret_cast<const uint4 *>(a.cnt + t0)[1];
        c[0] = x.x; c[1] = x.y; c[2] = x.z; c[3] = x.w; c[4] = y.x; c[5] = y.y; c[6] = y.z; c[7] = y.w;
    }
    u64 sum = 0, hocc = 0;
    u32 nh = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        u32 m = (u32)(m8 >> (8 * e)) & 0xffu;
        sum += c[e];
        if (m & M_HARD) { hocc += c[e]; nh += (m & M_HEAD) ? 1u : 0u; }
    }
    u64 T, Hocc;
    const u64 toff = block_excl_scan<u64>(sum, T, ws64);
    block_excl_scan<u64>(hocc, Hocc, ws64);
    u32 htot;
    u32 hoff = block_excl_scan<u32>(nh, htot, ws32);
    if (tid == 0) {
        s_hbase = htot ? atomicAdd(a.n_hard_groups, htot) : 0u;
        if (Hocc) atomicAdd((unsigned long long *)a.n_hard_occ, (unsigned long long)Hocc);
    }
    __syncthreads();
    const u64 X0 = a.tilebase[blockIdx.x];
    // per-element ilist base / suffix length, hard heads, big elements
    u32 ib[8], sl[8];
    u32 bigmask = 0;
    {
        u64 o = toff;
        u32 hi = s_hbase + hoff;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            u32 m = (u32)(m8 >> (8 * e)) & 0xffu;
            ib[e] = 0; sl[e] = 0;
            if (c[e] && !(m & M_HARD)) {
                if (SA) { ib[e] = a.ibase[t0 + e]; sl[e] = a.sl[t0 + e]; }
                else if (m & M_FULL) ib[e] = ilist_base(a, t0 + e, false);
                if (c[e] > 32) {
                    u32 b = atomicAdd(&s_nbig, 1u);
                    if (b < EMIT_BIGCAP) { big_o[b] = o; big_c[b] = c[e]; big_ib[b] = ib[e]; big_sl[b] = sl[e]; big_m[b] = m; bigmask |= 1u << e; }
                }
            }
            if ((m & (M_HARD | M_HEAD)) == (M_HARD | M_HEAD)) { a.hard_t[hi] = (u32)(t0 + e); a.hard_x[hi] = X0 + o; ++hi; }
            o += c[e];
        }
    }
    __syncthreads();
    const int nbig = s_nbig < (u32)EMIT_BIGCAP ? (int)s_nbig : EMIT_BIGCAP;
    const long long A = (long long)(X0 & 15);          // windows are 16-byte aligned in the output
    for (long long wlo = -A; wlo < (long long)T; wlo += WIN) {
        const long long whi = wlo + WIN;
        u64 o = toff;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            u32 m = (u32)(m8 >> (8 * e)) & 0xffu;
            if (c[e] && !(m & M_HARD) && !((bigmask >> e) & 1)) {
                long long lo = (long long)o > wlo ? (long long)o : wlo;
                long long hi = (long long)(o + c[e]) < whi ? (long long)(o + c[e]) : whi;
                for (long long s = lo; s < hi; ++s) emit_slot<U, SA>(a, m, ib[e], sl[e], (u64)s - o, (int)(s - wlo), obwt, osa);
            }
            o += c[e];
        }
        for (int b = 0; b < nbig; ++b) {
            long long lo = (long long)big_o[b] > wlo ? (long long)big_o[b] : wlo;
            long long hi = (long long)(big_o[b] + big_c[b]) < whi ? (long long)(big_o[b] + big_c[b]) : whi;
            for (long long s = lo + tid; s < hi; s += 256)
                emit_slot<U, SA>(a, big_m[b], big_ib[b], big_sl[b], (u64)s - big_o[b], (int)(s - wlo), obwt, osa);
        }
        __syncthreads();
        const long long vlo = wlo > 0 ? wlo : 0, vhi = whi < (long long)T ? whi : (long long)T;
        for (int ch = tid; ch < WIN / 16; ch += 256) {
            long long s0 = wlo + (long long)ch * 16;
            if (s0 + 16 <= vlo || s0 >= vhi) continue;
            u8 *g = a.bwt + (X0 + s0);      // X0 + s0 is a multiple of 16 (may lie before X0 only when skipped above)
            if (s0 >= vlo && s0 + 16 <= vhi) *reinterpret_cast<uint4 *>(g) = *reinterpret_cast<const uint4 *>(obwt + ch * 16);
            else for (int b = 0; b < 16; ++b) if (s0 + b >= vlo && s0 + b < vhi) g[b] = obwt[ch * 16 + b];
        }
        if (SA) {
            constexpr int PER = 16 / U;
            UT *gs = reinterpret_cast<UT *>(a.sa_out);
            for (int ch = tid; ch < WIN / PER; ch += 256) {
                long long s0 = wlo + (long long)ch * PER;
                if (s0 + PER <= vlo || s0 >= vhi) continue;
                if (s0 >= vlo && s0 + PER <= vhi) *reinterpret_cast<uint4 *>(gs + (X0 + s0)) = *reinterpret_cast<const uint4 *>(osa + ch * PER);
                else for (int b = 0; b < PER; ++b) if (s0 + b >= vlo && s0 + b < vhi) gs[X0 + s0 + b] = osa[ch * PER + b];
            }
        }
        __syncthreads();
    }
}
#endif

#ifndef PFB_EMU
// ---- emission with SA output: load-balanced expansion ---------------------------------------------------
// With SA output every BWT position costs two dependent gathers (ilist -> bwsai), and a suffix emits occ(word)
// positions - tens to hundreds in a pangenome.  One thread walking "its" 8 suffixes leaves the CTA waiting at the
// barrier for the thread with the longest walk (ncu on the first version: 18 barrier-stall cycles per issue, 15 % of
// DRAM peak).  Here the tile's per-suffix output offsets go to shared memory and every OUTPUT SLOT of a window is
// handled by its own thread (binary search slot -> suffix), four slots per thread in flight.
static const int EMIT_SA_WIN = 4096;
template <int U> struct EmitSaSmem {
    static const size_t BYTES = (size_t)EMIT_SA_WIN * U + (size_t)(TILE + 2) * 8 + (size_t)TILE * 8 + EMIT_SA_WIN + TILE;
};

template <int U>
__global__ void __launch_bounds__(256) k_emit_sa_kernel(EmitArgs a) {
    typedef typename UIntOf<U>::type UT;
    constexpr int WIN = EMIT_SA_WIN;
    extern __shared__ __align__(16) u8 esm[];
    UT *osa = reinterpret_cast<UT *>(esm);                              // [WIN]
    u64 *eoff = reinterpret_cast<u64 *>(esm + (size_t)WIN * U);         // [TILE + 1] first output slot of every suffix of the tile
    u32 *s_ib = reinterpret_cast<u32 *>(eoff + TILE + 2);               // [TILE]
    u32 *s_sl = s_ib + TILE;                                            // [TILE]
    u8 *obwt = reinterpret_cast<u8 *>(s_sl + TILE);                     // [WIN]
    u8 *s_m = obwt + WIN;                                               // [TILE]
    __shared__ u64 ws64[8];
    __shared__ u32 ws32[8];
    __shared__ u32 s_hbase;
    const int tid = threadIdx.x;
    const u64 t0 = (u64)blockIdx.x * TILE + (u64)tid * 8;
    const u64 m8 = *reinterpret_cast<const u64 *>(a.meta + t0);
    u32 c[8];
    {
        uint4 x = reinterpret_cast<const uint4 *>(a.cnt + t0)[0], y = reinterpret_cast<const uint4 *>(a.cnt + t0)[1];
        c[0] = x.x; c[1] = x.y; c[2] = x.z; c[3] = x.w; c[4] = y.x; c[5] = y.y; c[6] = y.z; c[7] = y.w;
    }
    u64 sum = 0, hocc = 0;
    u32 nh = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        u32 m = (u32)(m8 >> (8 * e)) & 0xffu;
        sum += c[e];
        if (m & M_HARD) { hocc += c[e]; nh += (m & M_HEAD) ? 1u : 0u; }
    }
    u64 T, Hocc;
    const u64 toff = block_excl_scan<u64>(sum, T, ws64);
    block_excl_scan<u64>(hocc, Hocc, ws64);
    u32 htot;
    u32 hoff = block_excl_scan<u32>(nh, htot, ws32);
    if (tid == 0) {
        s_hbase = htot ? atomicAdd(a.n_hard_groups, htot) : 0u;
        if (Hocc) atomicAdd((unsigned long long *)a.n_hard_occ, (unsigned long long)Hocc);
    }
    __syncthreads();
    const u64 X0 = a.tilebase[blockIdx.x];
    {
        u64 o = toff;
        u32 hi = s_hbase + hoff;
        uint4 ib0 = reinterpret_cast<const uint4 *>(a.ibase + t0)[0], ib1 = reinterpret_cast<const uint4 *>(a.ibase + t0)[1];
        uint4 sl0 = reinterpret_cast<const uint4 *>(a.sl + t0)[0], sl1 = reinterpret_cast<const uint4 *>(a.sl + t0)[1];
        const u32 ibv[8] = {ib0.x, ib0.y, ib0.z, ib0.w, ib1.x, ib1.y, ib1.z, ib1.w};
        const u32 slv[8] = {sl0.x, sl0.y, sl0.z, sl0.w, sl1.x, sl1.y, sl1.z, sl1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const u32 m = (u32)(m8 >> (8 * e)) & 0xffu;
            const int le = tid * 8 + e;
            eoff[le] = o; s_ib[le] = ibv[e]; s_sl[le] = slv[e]; s_m[le] = (u8)m;
            if ((m & (M_HARD | M_HEAD)) == (M_HARD | M_HEAD)) { a.hard_t[hi] = (u32)(t0 + e); a.hard_x[hi] = X0 + o; ++hi; }
            o += c[e];
        }
        if (tid == 255) eoff[TILE] = o;
    }
    __syncthreads();
    const long long A = (long long)(X0 & 15);          // windows are 16-byte aligned in the output
    for (long long wlo = -A; wlo < (long long)T; wlo += WIN) {
        const long long whi = wlo + WIN;
        const long long vlo = wlo > 0 ? wlo : 0, vhi = whi < (long long)T ? whi : (long long)T;
        for (long long s0 = vlo + tid; s0 < vhi; s0 += 4 * 256) {
            u32 ii[4], me[4], sle[4];
            bool ok[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const long long sk = s0 + 256 * k;
                ok[k] = sk < vhi;
                ii[k] = 0; me[k] = 0; sle[k] = 0;
                if (ok[k]) {
                    int lo = 0, hi = TILE;                              // largest element with eoff <= sk
                    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (eoff[mid] <= (u64)sk) lo = mid; else hi = mid; }
                    me[k] = s_m[lo];
                    ok[k] = !(me[k] & M_HARD);
                    if (ok[k]) { sle[k] = s_sl[lo]; ii[k] = a.ilist[(u64)s_ib[lo] + ((u64)sk - eoff[lo])]; }
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (!ok[k]) continue;
                const int pos = (int)(s0 + 256 * k - wlo);
                osa[pos] = (UT)(a.bwsai[ii[k]] - (u64)sle[k]);
                obwt[pos] = (me[k] & M_FULL) ? a.bwlast[ii[k]] : prev_byte(me[k]);      // include/pfbwt.hpp:119
            }
        }
        __syncthreads();
        for (int ch = tid; ch < WIN / 16; ch += 256) {
            long long s0 = wlo + (long long)ch * 16;
            if (s0 + 16 <= vlo || s0 >= vhi) continue;
            u8 *g = a.bwt + (X0 + s0);
            if (s0 >= vlo && s0 + 16 <= vhi) *reinterpret_cast<uint4 *>(g) = *reinterpret_cast<const uint4 *>(obwt + ch * 16);
            else for (int b = 0; b < 16; ++b) if (s0 + b >= vlo && s0 + b < vhi) g[b] = obwt[ch * 16 + b];
        }
        {
            constexpr int PER = 16 / U;
            UT *gs = reinterpret_cast<UT *>(a.sa_out);
            for (int ch = tid; ch < WIN / PER; ch += 256) {
                long long s0 = wlo + (long long)ch * PER;
                if (s0 + PER <= vlo || s0 >= vhi) continue;
                if (s0 >= vlo && s0 + PER <= vhi) *reinterpret_cast<uint4 *>(gs + (X0 + s0)) = *reinterpret_cast<const uint4 *>(osa + ch * PER);
                else for (int b = 0; b < PER; ++b) if (s0 + b >= vlo && s0 + b < vhi) gs[X0 + s0 + b] = osa[ch * PER + b];
            }
        }
        __syncthreads();
    }
}
#endif

// hard groups of up to HARD_SMALL members: 8 lanes per group rank every occurrence among the
// (ascending) ilist ranges of the other members by binary search; larger groups are listed for the
// radix-sort path (emit_big_groups).
static const u32 HARD_SMALL = 16;

PFB_HD inline u32 lower_bound_u32(const u32 *p, u32 n, u32 v) {
    u32 lo = 0, hi = n;
    while (lo < hi) { u32 mid = (lo + hi) >> 1; if (p[mid] < v) lo = mid + 1; else hi = mid; }
    return lo;
}

#ifndef PFB_EMU
template <int U, bool SA>
__global__ void __launch_bounds__(256) k_emit_hard_kernel(EmitArgs a, u32 nh, u32 *biglist, u32 *nbig) {
    typedef typename UIntOf<U>::type UT;
    __shared__ u32 s_ib[32][HARD_SMALL], s_c[32][HARD_SMALL];
    const int sub = threadIdx.x >> 3, l8 = threadIdx.x & 7;
    const u64 gi = (u64)blockIdx.x * 32 + (u64)sub;
    if (gi >= nh) return;
    const int sh = 8 * ((threadIdx.x >> 3) & 3);
    const u32 mask = 0xffu << sh;
    const u64 t0 = a.hard_t[gi], X = a.hard_x[gi];
    u32 g = 0;
    for (u32 base = 1; base <= HARD_SMALL; base += 8) {            // meta is padded with head bytes past D
        bool stop = (a.meta[t0 + base + l8] & M_HEAD) != 0;
        u32 bal = (__ballot_sync(mask, stop) >> sh) & 0xffu;
        if (bal) { g = base + (u32)__ffs((int)bal) - 1; break; }
    }
    if (g == 0 || g > a.small_lim) { if (l8 == 0) biglist[atomicAdd(nbig, 1u)] = (u32)gi; return; }
    for (u32 k = l8; k < g; k += 8) { s_ib[sub][k] = ilist_base(a, t0 + k, SA); s_c[sub][k] = a.cnt[t0 + k]; }
    __syncwarp(mask);
    u64 T = 0;
    for (u32 k = 0; k < g; ++k) T += s_c[sub][k];
    for (u64 q = l8; q < T; q += 8) {
        u32 k = 0;
        u64 pre = 0;
        while (pre + s_c[sub][k] <= q) { pre += s_c[sub][k]; ++k; }
        u32 j = (u32)(q - pre);
        u32 v = a.ilist[(u64)s_ib[sub][k] + j];
        u64 rank = j;
        for (u32 k2 = 0; k2 < g; ++k2)
            if (k2 != k) rank += lower_bound_u32(a.ilist + s_ib[sub][k2], s_c[sub][k2], v);
        u64 x = X + rank;
        a.bwt[x] = prev_byte(a.meta[t0 + k]);                     // include/pfbwt.hpp:173-178
        if (SA) reinterpret_cast<UT *>(a.sa_out)[x] = (UT)(a.bwsai[v] - (u64)a.sl[t0 + k]);
    }
}
#endif

// hard groups with more members but at most HARD_MID occurrences: one warp per group loads every
// occurrence's ilist position into shared memory and ranks it by counting the smaller ones (the
// positions are distinct); all lanes read the same shared word at a time (broadcast).
static const u32 HARD_MID = 1024;

#ifndef PFB_EMU
template <int U, bool SA>
__global__ void __launch_bounds__(256) k_emit_hard_mid_kernel(EmitArgs a, const u32 *midlist, u32 nmid, u32 *biglist, u32 *nbig) {
    typedef typename UIntOf<U>::type UT;
    extern __shared__ __align__(16) unsigned char mid_smem[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    u32 *s_val = reinterpret_cast<u32 *>(mid_smem) + wid * HARD_MID;
    unsigned short *s_mem = reinterpret_cast<unsigned short *>(mid_smem + 8 * HARD_MID * 4) + wid * HARD_MID;
    const u64 wg = (u64)blockIdx.x * 8 + wid;
    if (wg >= nmid) return;
    const u32 gi = midlist[wg];
    const u64 t0 = a.hard_t[gi], X = a.hard_x[gi];
    // members and occurrences (every member of a valid group has at least one occurrence)
    u32 g = 0;
    u64 T = 0;
    for (u32 base = 0; base <= HARD_MID; base += 32) {
        const u64 t = t0 + base + lane;
        const bool stop = (base + lane) != 0 && (t >= a.D || (a.meta[t] & M_HEAD));
        const u32 bal = __ballot_sync(0xffffffffu, stop);
        const u32 nm = bal ? (u32)__ffs((int)bal) - 1 : 32u;      // members in this chunk
        u32 c = lane < nm ? a.cnt[t] : 0u;
#pragma unroll
        for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
        T += c; g += nm;
        if (bal || T > HARD_MID) break;
        if (base + 32 > HARD_MID) { T = HARD_MID + 1; break; }
    }
    if (T > a.mid_lim) { if (lane == 0) biglist[atomicAdd(nbig, 1u)] = gi; return; }
    u32 run = 0;
    for (u32 base = 0; base < g; base += 32) {
        const u32 k = base + lane;
        u32 c = 0, ib = 0;
        if (k < g) { c = a.cnt[t0 + k]; ib = ilist_base(a, t0 + k, SA); }
        u32 inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { u32 o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
        u32 off = run + inc - c;
        for (u32 j = 0; j < c; ++j) { s_val[off + j] = a.ilist[(u64)ib + j]; s_mem[off + j] = (unsigned short)k; }
        run += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncwarp();
    const u32 Tn = (u32)T;
    for (u32 q = lane; q < Tn; q += 32) {
        const u32 v = s_val[q];
        u32 rank = 0;
        for (u32 i = 0; i < Tn; ++i) rank += s_val[i] < v ? 1u : 0u;
        const u32 k = s_mem[q];
        const u64 x = X + rank;
        a.bwt[x] = prev_byte(a.meta[t0 + k]);                     // include/pfbwt.hpp:173-178
        if (SA) reinterpret_cast<UT *>(a.sa_out)[x] = (UT)(a.bwsai[v] - (u64)a.sl[t0 + k]);
    }
}
static const size_t MID_SMEM = 8 * HARD_MID * 4 + 8 * HARD_MID * 2;
#endif

// serial twins (emulator) ------------------------------------------------------------------
#ifdef PFB_EMU
template <int U>
inline void emu_store_sa(u8 *sa_out, u64 x, u64 v) {
    if (U == 4) reinterpret_cast<u32 *>(sa_out)[x] = (u32)v; else reinterpret_cast<u64 *>(sa_out)[x] = v;
}
template <int U>
inline void emu_emit(const EmitArgs &a, bool SA) {
    for (u64 tile = 0; tile < tiles_of(a.D); ++tile) {
        u64 x = a.tilebase[tile];
        for (u64 t = tile * TILE; t < (tile + 1) * TILE && t < a.D; ++t) {
            u32 m = a.meta[t], c = a.cnt[t];
            if (m & M_HARD) {
                *a.n_hard_occ += c;
                if (m & M_HEAD) { u32 h = (*a.n_hard_groups)++; a.hard_t[h] = (u32)t; a.hard_x[h] = x; }
            } else if (c) {
                u32 ib = (SA || (m & M_FULL)) ? ilist_base(a, t, SA) : 0;
                for (u32 j = 0; j < c; ++j) {
                    u8 byte = prev_byte(m);
                    if (SA || (m & M_FULL)) {
                        u32 ii = a.ilist[(u64)ib + j];
                        if (m & M_FULL) byte = a.bwlast[ii];
                        if (SA) emu_store_sa<U>(a.sa_out, x + j, a.bwsai[ii] - (u64)a.sl[t]);
                    }
                    a.bwt[x + j] = byte;
                }
            }
            x += c;
        }
    }
}
template <int U>
inline void emu_emit_hard(const EmitArgs &a, bool SA, u32 nh, u32 *biglist, u32 *nbig) {
    for (u32 gi = 0; gi < nh; ++gi) {
        u64 t0 = a.hard_t[gi], X = a.hard_x[gi];
        u32 g = 1;
        while (!(a.meta[t0 + g] & M_HEAD)) ++g;
        if (g > a.small_lim) { biglist[(*nbig)++] = gi; continue; }
        for (u32 k = 0; k < g; ++k) {
            u32 ib = ilist_base(a, t0 + k, SA), c = a.cnt[t0 + k];
            for (u32 j = 0; j < c; ++j) {
                u32 v = a.ilist[(u64)ib + j];
                u64 rank = j;
                for (u32 k2 = 0; k2 < g; ++k2)
                    if (k2 != k) rank += lower_bound_u32(a.ilist + ilist_base(a, t0 + k2, SA), a.cnt[t0 + k2], v);
                a.bwt[X + rank] = prev_byte(a.meta[t0 + k]);
                if (SA) emu_store_sa<U>(a.sa_out, X + rank, a.bwsai[v] - (u64)a.sl[t0 + k]);
            }
        }
    }
}
template <int U>
inline void emu_emit_hard_mid(const EmitArgs &a, bool SA, const u32 *midlist, u32 nmid, u32 *biglist, u32 *nbig) {
    for (u32 w = 0; w < nmid; ++w) {
        u32 gi = midlist[w];
        u64 t0 = a.hard_t[gi], X = a.hard_x[gi];
        u32 g = 1;
        while (t0 + g < a.D && !(a.meta[t0 + g] & M_HEAD)) ++g;
        u64 T = 0;
        for (u32 k = 0; k < g; ++k) T += a.cnt[t0 + k];
        if (T > a.mid_lim) { biglist[(*nbig)++] = gi; continue; }
        std::vector<std::pair<u32, u32>> occ;
        for (u32 k = 0; k < g; ++k) {
            u32 ib = ilist_base(a, t0 + k, SA);
            for (u32 j = 0; j < a.cnt[t0 + k]; ++j) occ.push_back({a.ilist[(u64)ib + j], k});
        }
        std::sort(occ.begin(), occ.end());
        for (size_t r = 0; r < occ.size(); ++r) {
            a.bwt[X + r] = prev_byte(a.meta[t0 + occ[r].second]);
            if (SA) emu_store_sa<U>(a.sa_out, X + r, a.bwsai[occ[r].first] - (u64)a.sl[t0 + occ[r].second]);
        }
    }
}
#endif

// ---------------------------------------------------------------------------
// hard groups with more than HARD_MID occurrences: all occurrences of a chunk of groups get a
// (dense group id, ilist position) key, one radix sort orders them, a scatter writes them.  Chunks
// bound the work memory (BIG_CHUNK occurrences at a time).
// ---------------------------------------------------------------------------
static const u64 BIG_CHUNK = 1ull << 28;

template <int U>
inline int emit_big_groups(const EmitArgs &a, bool SA, const u32 *biglist, u64 nbig, u64 P, Workspace &ws, std::string &err) {
    pfbStream_t st = ws.st;
    const u64 D = a.D;
    const u8 *meta = a.meta;
    // group sizes: 32 logical lanes per group look for the next head
    DevBuf<u32> gend(nbig + 1);
    {
        u32 *ge = gend.get(); const u32 *ht = a.hard_t;
        launch_for(nbig + 1, st, PFB_LAMBDA(u64 b) { ge[b] = b < nbig ? (u32)D : 0u; });
        launch_for(nbig * 32, st, PFB_LAMBDA(u64 i) {
            u64 b = i >> 5, lane = i & 31;
            u64 t0 = ht[biglist[b]];
            for (u64 t = t0 + 1 + lane; t < D; t += 32) {
                if (t >= ge[b]) break;
                if (meta[t] & M_HEAD) { atomic_min(&ge[b], (u32)t); break; }
            }
        });
    }
    DevBuf<u32> gsz(nbig + 1);
    DevBuf<u64> moff(nbig + 1);
    {
        const u32 *ge = gend.get(); const u32 *ht = a.hard_t; u32 *gs = gsz.get();
        launch_for(nbig + 1, st, PFB_LAMBDA(u64 b) { gs[b] = b < nbig ? ge[b] - ht[biglist[b]] : 0u; });
    }
    exclusive_sum<u32, u64>(gsz.get(), moff.get(), nbig + 1, ws.sc, st);
    const u64 mh = read_scalar(moff.get() + nbig, st);
    if (mh > 0xFFFFFFFFull) { err = "internal: too many members in large hard groups"; return PFB_ERR_INTERNAL; }
    DevBuf<u32> hl(mh), mb(mh), hc(mh + 1), mib(mh);
    DevBuf<u64> hoff(mh + 1), goff(nbig + 1);
    {
        const u32 *ht = a.hard_t; const u32 *gs = gsz.get(); const u64 *mo = moff.get();
        u32 *hp = hl.get(); u32 *mbp = mb.get(); u32 *c_ = hc.get(); u32 *ib = mib.get();
        const EmitArgs aa = a;
        launch_for(nbig * 32, st, PFB_LAMBDA(u64 i) {
            u64 b = i >> 5, lane = i & 31;
            u64 t0 = ht[biglist[b]];
            for (u64 k = lane; k < gs[b]; k += 32) {
                u64 m = mo[b] + k;
                hp[m] = (u32)(t0 + k); mbp[m] = (u32)b; c_[m] = aa.cnt[t0 + k]; ib[m] = ilist_base(aa, t0 + k, SA);
            }
            if (i == 0) c_[mh] = 0;
        });
    }
    exclusive_sum<u32, u64>(hc.get(), hoff.get(), mh + 1, ws.sc, st);
    {
        const u64 *ho = hoff.get(); const u64 *mo = moff.get(); u64 *go = goff.get();
        launch_for(nbig + 1, st, PFB_LAMBDA(u64 b) { go[b] = ho[mo[b]]; });
    }
    std::vector<u64> h_goff(nbig + 1), h_moff(nbig + 1);
    d2h(h_goff.data(), goff.get(), (nbig + 1) * 8, st);
    d2h(h_moff.data(), moff.get(), (nbig + 1) * 8, st);
    stream_sync(st);
    const int ibits = bits_for(P);           // ilist values are in [0, P]
    const u64 imask = (1ull << ibits) - 1;
    for (u64 b0 = 0; b0 < nbig;) {
        u64 b1 = b0 + 1;
        while (b1 < nbig && h_goff[b1 + 1] - h_goff[b0] <= BIG_CHUNK) ++b1;
        const u64 E0 = h_goff[b0], ne = h_goff[b1] - E0, m0 = h_moff[b0], nm = h_moff[b1] - m0;
        const int gbits = bits_for(b1 - b0 - 1);
        if (ibits + gbits > 64) { err = "internal: hard key overflow"; return PFB_ERR_INTERNAL; }
        if (ne) {
            DevBuf<u64> hk(ne), hkb(ne);
            DevBuf<u32> hv(ne), hvb(ne);
            {
                const u32 *mbp = mb.get(); const u32 *c_ = hc.get(); const u32 *ib = mib.get(); const u64 *ho = hoff.get();
                const u32 *ilist = a.ilist; u64 *kp = hk.get(); u32 *vp = hv.get();
                launch_for(nm * 32, st, PFB_LAMBDA(u64 i) {
                    u64 m = m0 + (i >> 5), lane = i & 31;
                    u64 gg = (u64)mbp[m] - b0, e = ho[m] - E0;
                    for (u32 j = (u32)lane; j < c_[m]; j += 32) {
                        kp[e + j] = (gg << ibits) | (u64)ilist[(u64)ib[m] + j];
                        vp[e + j] = (u32)m;
                    }
                });
            }
            sort_pairs(hk, hkb, hv, hvb, ne, 0, ibits + gbits, ws.sc, st);
            {
                const u64 *kp = hk.get(); const u32 *vp = hv.get(); const u64 *go = goff.get(); const u32 *hp = hl.get();
                const u64 *hx = a.hard_x; u8 *bwt = a.bwt; u8 *sa_out = a.sa_out; const u64 *bwsai = a.bwsai; const u32 *slp = a.sl;
                launch_for(ne, st, PFB_LAMBDA(u64 e) {
                    u64 b = b0 + (kp[e] >> ibits);
                    u64 x = hx[biglist[b]] + (E0 + e - go[b]);
                    u32 t = hp[vp[e]];
                    bwt[x] = prev_byte(meta[t]);                      // include/pfbwt.hpp:173-178
                    if (SA) {
                        u64 v = bwsai[kp[e] & imask] - (u64)slp[t];
                        if (U == 4) reinterpret_cast<u32 *>(sa_out)[x] = (u32)v; else reinterpret_cast<u64 *>(sa_out)[x] = v;
                    }
                });
            }
        }
        b0 = b1;
    }
    return PFB_OK;
}

}  // namespace pfb
