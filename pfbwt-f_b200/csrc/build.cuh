// build.cuh -- BWT build from a device-resident parse: dictionary suffix sort,
// group detection and segmented emission of .bwt/.sa/.ssa/.esa.
// Replaces PrefixFreeBWT::sort_dict_suffixes + generate_bwt_lcp
// (include/pfbwt.hpp:96-223) and the writer lambda of run_pfbwt
// (src/pfbwt-f.cpp:298-328).
//
// Data layout in HBM (D = dict bytes, Dw = words, P = phrases, n = |S|; per-suffix arrays are
// padded to a multiple of the 2048-element tile):
//   key[D]   u64  enumerative code of the first L symbols of the suffix, symbols consumed,
//                 closed flag, code of the preceding byte (dictkeys.cuh); dropped after the sort
//   sa[D]    DP   dictionary position of the j-th smallest suffix (DP = u32; u64 for a dictionary of 2^32 bytes or more, which
//                 is always sorted in key ranges of fewer than 2^32 suffixes: indices INTO the sorted arrays stay u32)
//   meta[D]  u8   preceding-byte code | head | full | valid | hard (emit.cuh)
//   wc[Dw+2] u64  (word start | ilist base << 32): one 16-byte read gives both ends
//   wsamp[D/128] u32  word id at every 128th dictionary byte (position -> word)
//   cnt[D]   u32  occurrences emitted by the j-th suffix (0 when suffix length <= w)
//   tilebase[D/2048] u64  BWT position of the first output of a tile (offsets inside a tile are
//                 recomputed by a block scan in the emission kernel)
//   ibase[D], sl[D] u32  first ilist slot of the word / suffix length (only with SA output)
#pragma once
#include "dictkeys.cuh"

#include "emit.cuh"

namespace pfb {

// ordered compaction of the indices j in [0,n) with flags[j] != 0
inline u64 compact_indices(const u8 *flags, u64 n, DevBuf<u32> &scratch_idx, DevBuf<u32> &list, Workspace &ws) {
    if (scratch_idx.size() < n) scratch_idx.alloc(n);
    u64 m = scan_flags(flags, scratch_idx.get(), n, ws);
    list.alloc(m);
    if (m) {
        const u32 *ix = scratch_idx.get(); u32 *lp = list.get();
        launch_for(n, ws.st, PFB_LAMBDA(u64 j) { if (flags[j]) lp[ix[j]] = (u32)j; });
    }
    return m;
}

// ---------------------------------------------------------------------------
// K6  dictionary suffix sort.  One radix sort on the enumerative code of the first
//     L symbols (dictkeys.cuh), then only the tie groups that are still open are
//     refined: rounds that append the code of further symbols read from the
//     dictionary (each open group continues at its own depth, because a stopper
//     symbol ends a window early), and - for long repeats - prefix doubling over
//     rank[] (Larsson-Sadakane) from the smallest depth still open.  Replaces
//     gsacak + the LCP array (include/pfbwt.hpp:206-223): a group of equal strings
//     is exactly a run with glcp >= suff_len (:136-145).
//     Result: sa[] and meta[] (preceding-byte code + head bit), both padded(D) long.
// ---------------------------------------------------------------------------
//     cnt[] (occurrences per suffix) and the valid/full bits of meta come out of the keys' upper
//     bits; *reclassify is set when they have to be recomputed by lookup (prefix doubling ran).
//     Multi-GPU (sh != nullptr): the rank sorts only the valid suffixes whose code falls in its
//     range of the key space (splitters from a weighted histogram of the leading code bits, computed
//     identically on every rank from the replicated dictionary); *n_sorted = how many those are.
struct ShardSpec { int rank = 0, world = 1; Comm *comm = nullptr; };   // comm == nullptr: key range `rank` of `world` on ONE GPU

// ---- pivot rounds: refinement of LONG ties (repetitive dictionaries: variant phrases of a pangenome
// share suffixes of hundreds of symbols, which rounds of <= 26 symbols close far too slowly) ----------
// Every member of an open tie group is compared byte by byte with the group's first member (the
// pivot), both read from the group's depth on.  With l = common length and c = the member's byte at
// the mismatch, the group is ordered in ONE sort by
//     less than pivot:  (0, l, c)         -- a shorter common prefix with the pivot sorts first
//     identical:        (1, 0, 0) closed  -- equal up to and including EndOfWord
//     greater:          (2, LMAX - l, c)  -- a longer common prefix sorts first
// and members with equal keys are tied on l+1 more symbols (a new, smaller group).  Comparisons
// stop after PIVOT_CAP symbols ("clamped": stays tied with the pivot, continues next round).
// 8 dictionary bytes starting at any position (little endian), from two aligned words
PFB_HD inline u64 load8_unaligned(const u8 *dict, u64 pos) {
    const u64 *w = reinterpret_cast<const u64 *>(dict + (pos & ~7ull));
    const u32 sh = (u32)(pos & 7) * 8;
    const u64 lo = w[0];
    return sh ? (lo >> sh) | (w[1] << (64 - sh)) : lo;
}
// 32 dictionary bytes starting at any position as four little-endian words, from three aligned 16-byte loads: a divergent
// warp pays per load instruction (one L1 wavefront per distinct line), so the comparison loop wants few, wide loads
PFB_HD inline void load32_unaligned(const u8 *dict, u64 pos, u64 *o) {
    const u64 A = pos & ~15ull;
    u64 t[6];
#ifdef __CUDA_ARCH__
    const uint4 *q = reinterpret_cast<const uint4 *>(dict + A);
    const uint4 v0 = q[0], v1 = q[1], v2 = q[2];
    t[0] = (u64)v0.x | ((u64)v0.y << 32); t[1] = (u64)v0.z | ((u64)v0.w << 32);
    t[2] = (u64)v1.x | ((u64)v1.y << 32); t[3] = (u64)v1.z | ((u64)v1.w << 32);
    t[4] = (u64)v2.x | ((u64)v2.y << 32); t[5] = (u64)v2.z | ((u64)v2.w << 32);
#else
    const u64 *q = reinterpret_cast<const u64 *>(dict + A);
    for (int i = 0; i < 6; ++i) t[i] = q[i];
#endif
    const u32 wi = (u32)(pos & 15) >> 3, bs = (u32)(pos & 7) * 8;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const u64 lo = wi ? t[k + 1] : t[k], hi2 = wi ? t[k + 2] : t[k + 1];
        o[k] = bs ? (lo >> bs) | (hi2 << (64 - bs)) : lo;
    }
}
// Long single-symbol runs of the dictionary (poly-A phrases left by --non-acgt-to-a on Mbp N runs, poly-N without it):
// two suffixes inside such runs share min(remaining run lengths) symbols, which the comparison skips in one step
// instead of reading them - otherwise ties of millions of symbols would need prefix doubling over the whole dictionary.
template <class DP>
struct RunTableT {
    const DP *start = nullptr, *end = nullptr;       // maximal runs [start, end) of one byte value, length >= RUN_MIN, ascending
    u32 R = 0;
    PFB_HD inline u64 remaining(u64 pos) const {     // symbols of the run containing pos from pos on (0 if none)
        if (!R || pos < start[0]) return 0;
        u32 lo = 0, hi = R;
        while (hi - lo > 1) { u32 mid = (lo + hi) >> 1; if ((u64)start[mid] <= pos) lo = mid; else hi = mid; }
        return pos < (u64)end[lo] ? (u64)end[lo] - pos : 0;
    }
};
static const u64 RUN_BLOCK = 256, RUN_MIN = 3 * RUN_BLOCK;
// All positions inside a single-symbol run are tied with each other after the main sort, so an open list shorter than this at
// the first pivot round means that no run with open ties is longer than it: every comparison ends within the default cap
// (2^PIVOT_LBITS symbols) and the scan of the whole dictionary for runs (0.6 ms per GB) is skipped.
static const u64 RUN_TABLE_MIN_OPEN = 2048;

template <class RT>
PFB_HD inline u64 pivot_key(const u8 *dict, u64 D, u64 a, u64 p, int lb, const RT &rt) {
    const u64 cap = (1ull << lb) - 1;
    u64 l = 0;
    // 8 symbols per step while both windows (and the second aligned word behind them) lie inside the dictionary
    const u64 hi = a > p ? a : p;
    // four 8-symbol steps at a time while nothing happens in them (all loads of a step are independent: the comparison is
    // bound by memory latency, variants of a phrase agree for tens of symbols); the exact step loop below takes over at the
    // first difference, terminator or candidate run
    while (l + 32 <= cap && hi + l + 48 <= D) {
        u64 x[4], y[4];
        load32_unaligned(dict, a + l, x);
        load32_unaligned(dict, p + l, y);
        u64 acc = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) acc |= (x[c] ^ y[c]) | ((x[c] - 0x0202020202020202ull) & ~x[c] & 0x8080808080808080ull);
        if (acc || (rt.R && x[0] == (x[0] & 0xffull) * 0x0101010101010101ull)) break;
        l += 32;
    }
    while (l + 8 <= cap && hi + l + 16 <= D) {
        const u64 x = load8_unaligned(dict, a + l), y = load8_unaligned(dict, p + l);
        const u64 d = x ^ y;
        const u64 e = (x - 0x0202020202020202ull) & ~x & 0x8080808080808080ull;      // lowest set bit: first byte < 2 (EndOfWord / EndOfDict)
        if (d | e) {
            const u64 ev = (d ? d & (0 - d) : ~0ull), ee = (e ? e & (0 - e) : ~0ull);   // lowest set bits
            // the first event decides: a differing byte (also when that byte is EndOfWord on one side), else a common terminator
            const u64 first = ev < ee ? ev : ee;
            const u32 b = (u32)(63 - clz64(first)) >> 3;
            return pivot_key_result((u32)l + b, (u32)(x >> (8 * b)) & 0xffu, (u32)(y >> (8 * b)) & 0xffu, lb);
        }
        u64 step = 8;
        if (rt.R && x == (x & 0xffull) * 0x0101010101010101ull) {                    // 8 equal symbols on both sides: inside runs?
            const u64 ra = rt.remaining(a + l), rp = rt.remaining(p + l);
            const u64 both = ra < rp ? ra : rp;
            if (both > 8) step = both;
        }
        l += step;
    }
    if (l > cap) l = cap;
    for (; l < cap; ++l) {
        const u32 x = a + l < D ? dict[a + l] : 0u, y = p + l < D ? dict[p + l] : 0u;
        if (x != y || x <= 1u) return pivot_key_result((u32)l, x, y, lb);
    }
    return pivot_key_clamped(lb);
}

// run table of the dictionary: 256-byte blocks of one byte value are found on the device, merged into candidate runs on
// the host (a handful even for a human genome) and trimmed to their exact ends
template <class DP>
inline void build_run_table(const u8 *dict, u64 D, DevBuf<DP> &rs, DevBuf<DP> &re, u32 *R_out, u64 *longest, Workspace &ws) {
    pfbStream_t st = ws.st;
    *R_out = 0; *longest = 0;
    const u64 nb = D / RUN_BLOCK;
    if (nb == 0) return;
    DevBuf<u8> uni(nb);
    DevBuf<u32> uidx(nb);
    {
        u8 *up = uni.get();
        launch_for(nb, st, PFB_LAMBDA(u64 b) {
            const u64 *w = reinterpret_cast<const u64 *>(dict + b * RUN_BLOCK);
            const u64 first = (w[0] & 0xffull) * 0x0101010101010101ull;
            u64 acc = 0;
            for (int k = 0; k < (int)(RUN_BLOCK / 8); ++k) acc |= w[k] ^ first;
            up[b] = acc ? 0 : 1;
        });
    }
    const u64 nu = scan_flags(uni.get(), uidx.get(), nb, ws);
    if (nu < RUN_MIN / RUN_BLOCK) return;
    DevBuf<u32> ub(nu);
    DevBuf<u8> us(nu);
    {
        const u8 *up = uni.get(); const u32 *ui = uidx.get(); u32 *o = ub.get(); u8 *osym = us.get();
        launch_for(nb, st, PFB_LAMBDA(u64 b) { if (up[b]) { o[ui[b]] = (u32)b; osym[ui[b]] = dict[b * RUN_BLOCK]; } });
    }
    std::vector<u32> hb(nu);
    std::vector<u8> hs(nu);
    d2h(hb.data(), ub.get(), nu * 4, st);
    d2h(hs.data(), us.get(), nu, st);
    stream_sync(st);
    std::vector<DP> cs, ce;
    std::vector<u8> csym;
    for (u64 i = 0; i < nu;) {
        u64 j = i + 1;
        while (j < nu && hb[j] == hb[j - 1] + 1 && hs[j] == hs[i]) ++j;
        if (j - i >= RUN_MIN / RUN_BLOCK) { cs.push_back((DP)((u64)hb[i] * RUN_BLOCK)); ce.push_back((DP)(((u64)hb[j - 1] + 1) * RUN_BLOCK)); csym.push_back(hs[i]); }
        i = j;
    }
    const u32 R = (u32)cs.size();
    if (!R) return;
    rs.alloc(R); re.alloc(R);
    DevBuf<u8> rsym(R);
    h2d(rs.get(), cs.data(), (size_t)R * sizeof(DP), st);
    h2d(re.get(), ce.data(), (size_t)R * sizeof(DP), st);
    h2d(rsym.get(), csym.data(), R, st);
    DevBuf<u64> lmax(1);
    dev_memset(lmax.get(), 0, 8, st);
    {
        DP *sp = rs.get(); DP *ep = re.get(); const u8 *sy = rsym.get(); u64 *lm = lmax.get();
        launch_for(R, st, PFB_LAMBDA(u64 r) {          // exact ends: at most RUN_BLOCK - 1 steps each way
            u64 s = sp[r], e = ep[r];
            const u8 c = sy[r];
            while (s > 0 && dict[s - 1] == c) --s;
            while (e < D && dict[e] == c) ++e;
            sp[r] = (DP)s; ep[r] = (DP)e;
            atomic_max(lm, e - s);
        });
    }
    *longest = read_scalar(lmax.get(), st);
    stream_sync(st);            // the host vectors above fed asynchronous copies
    *R_out = R;
}

// positions as u32 for prefix doubling (identity for DP = u32)
inline DevBuf<u32> &sa_as_u32(DevBuf<u32> &sa, DevBuf<u32> &, u64, pfbStream_t) { return sa; }
inline DevBuf<u32> &sa_as_u32(DevBuf<u64> &sa, DevBuf<u32> &tmp, u64 n, pfbStream_t st) {
    tmp.alloc(n);
    const u64 *s = sa.get(); u32 *d = tmp.get();
    launch_for(n, st, PFB_LAMBDA(u64 j) { d[j] = (u32)s[j]; });
    return tmp;
}
inline void sa_from_u32(DevBuf<u32> &, DevBuf<u32> &, u64, pfbStream_t) {}
inline void sa_from_u32(DevBuf<u64> &sa, DevBuf<u32> &tmp, u64 n, pfbStream_t st) {
    const u32 *s = tmp.get(); u64 *d = sa.get();
    launch_for(n, st, PFB_LAMBDA(u64 j) { d[j] = s[j]; });
}

template <class DP>
inline int sort_dict_suffixes(const DeviceParse &pp, int w, DevBuf<DP> &sa, DevBuf<u8> &meta, DevBuf<u32> &cnt, const WordLookupT<DP> &wl,
                              bool *reclassify, Workspace &ws, StageTimes &tm, Timer &timer, std::string &err,
                              const ShardSpec *sh = nullptr, u64 *n_sorted = nullptr) {
    pfbStream_t st = ws.st;
    const u64 D = pp.dsize;
    const u8 *dict = pp.dict.get();
    *reclassify = false;
    int L = 1;
    while (L < KEY_MAX_SYMS && (1ull << (2 * L)) < 128ull * D) ++L;     // ties become rare once 4^L >> D
    if (const char *e = getenv("PFB_KEY_SYMS")) { L = atoi(e); if (L < 1) L = 1; if (L > KEY_MAX_SYMS) L = KEY_MAX_SYMS; }   // test knob
    DevBuf<u32> apos, adep;
    u64 m;
    u64 N = D;                       // entries in the sorted array
    {
        DevBuf<u64> key, key_b;
        DevBuf<DP> val_b;
        const KeyExtra kx = key_extra_for(L);
        if (!sh && D > 0xFFFFFFF0ull) { err = "internal: a dictionary of 2^32 bytes or more is sorted in key ranges"; return PFB_ERR_INTERNAL; }
        if (!sh) {
            key.alloc(D); key_b.alloc(D); val_b.alloc(padded(D)); sa.alloc(padded(D));
            k_suffix_keys(dict, D, L, key.get(), wl, pp.occ.get(), w, kx, st);
            timer.lap("K6 keys");
            sort_pairs(key, key_b, sa, val_b, D, KEY_CODE_SHIFT, KEY_CODE_SHIFT + enum_bits(L), ws.sc, st, /*iota_vals=*/true);
        } else {
            const int cb = enum_bits(L), hb = cb < 14 ? cb : 14, hshift = cb - hb;
            const u64 nbins = 1ull << hb;
            std::vector<u64> h(nbins);
            {
                DevBuf<u64> hist(nbins);
                dev_memset(hist.get(), 0, nbins * 8, st);
                k_suffix_key_hist(dict, D, L, wl, pp.occ.get(), w, kx, hist.get(), hshift, st);
                d2h(h.data(), hist.get(), nbins * 8, st);
                stream_sync(st);
            }
            u64 total = 0;
            for (u64 b = 0; b < nbins; ++b) total += h[b];
            auto bound = [&](int r) -> u64 {      // first bin of rank r: smallest b with weight(bins < b) >= total*r/world
                if (r <= 0) return 0;
                if (r >= sh->world) return nbins;
                const u64 want = (u64)(((unsigned __int128)total * (unsigned)r + (unsigned)sh->world - 1) / (unsigned)sh->world);
                u64 cum = 0, b = 0;
                while (b < nbins && cum < want) cum += h[b++];
                return b;
            };
            timer.lap("K6 splitter histogram");
            if (!sh->comm) {
                // time-multiplexed form (one GPU sorting one key range after the other): select from all positions
                const u64 lo = bound(sh->rank) << hshift;
                const u64 hi = sh->rank + 1 >= sh->world ? ~0ull : bound(sh->rank + 1) << hshift;
                N = k_suffix_keys_select<DP>(dict, D, L, wl, pp.occ.get(), w, kx, lo, hi, key, sa, padded, ws);
                if (N > 0xFFFFFFF0ull) { err = "more than 2^32 dictionary suffixes in one key range (more buckets needed)"; return PFB_ERR_LIMIT; }
                timer.lap("K6 keys (selected)");
            } else {
                // position-sliced: this rank computes the keys of ITS 1/world of the dictionary positions and routes every
                // (key, position) pair to the owner of the key's range - one all-to-all over NVLink instead of every rank
                // scanning all D positions.  Pairs arrive ordered by source rank = by dictionary position.
                const int G = sh->world, g = sh->rank;
                if (G > KEY_MAX_DEST) { err = "more ranks than the key router supports"; return PFB_ERR_LIMIT; }
                std::vector<u64> bnd((size_t)G + 1);
                for (int r = 0; r <= G; ++r) bnd[(size_t)r] = r >= G ? ~0ull : bound(r) << hshift;
                const u64 ntiles = (D + KEYS_TILE - 1) / KEYS_TILE;
                const u64 t0 = ntiles * (u64)g / (u64)G, t1 = ntiles * (u64)(g + 1) / (u64)G;
                KeyRouterT<DP> kr;
                kr.dict = dict; kr.D = D; kr.L = L; kr.wl = wl; kr.occ = pp.occ.get(); kr.w = w; kr.kx = kx;
                kr.bound = bnd; kr.tile0 = t0; kr.ntl = t1 - t0;
                kr.count(ws);
                timer.lap("K6 keys (count per owner)");
                std::vector<u64> mat = sh->comm->gather_u64(kr.counts.data(), G, st);     // mat[s * G + d] = pairs s -> d
                N = 0;
                for (int s = 0; s < G; ++s) N += mat[(size_t)s * G + g];
                if (N > 0xFFFFFFF0ull) { err = "more than 2^32 dictionary suffixes in one rank's key range (use more GPUs)"; return PFB_ERR_LIMIT; }
                key.alloc(N); sa.alloc(padded(N));
                std::vector<u64 *> kouts((size_t)G);
                std::vector<DP *> iouts((size_t)G);
                void *mine[2] = {key.get(), sa.get()};
                std::vector<void *> peers;
                if (sh->comm->peer_pointers(mine, 2, peers, st)) {
                    // FUSED: the key kernel stores every pair straight into its owner's receive buffer over NVLink (peer
                    // memory): the all-to-all is the kernel's own unit-stride stores, overlapped with the key arithmetic
                    for (int d = 0; d < G; ++d) {
                        u64 before = 0;                                   // pairs the ranks before this one send to d
                        for (int s = 0; s < g; ++s) before += mat[(size_t)s * G + d];
                        kouts[(size_t)d] = (u64 *)peers[(size_t)d * 2] + before;
                        iouts[(size_t)d] = (DP *)peers[(size_t)d * 2 + 1] + before;
                    }
                    kr.scatter(kouts.data(), iouts.data(), ws);
                    stream_sync(st);
                    sh->comm->barrier(st);                                // every rank's stores into this rank's buffers have landed
                    u64 got = 0;
                    for (int s = 0; s < G; ++s) if (s != g) got += mat[(size_t)s * G + g];
                    sh->comm->bytes_moved += got * (8 + sizeof(DP));
                    timer.lap("K6 keys -> owners (peer stores over NVLink)");
                } else {
                    const u64 m = kr.destbase[(size_t)G];
                    DevBuf<u64> skey(m);
                    DevBuf<DP> sidx(m);
                    for (int d = 0; d < G; ++d) { kouts[(size_t)d] = skey.get() + kr.destbase[(size_t)d]; iouts[(size_t)d] = sidx.get() + kr.destbase[(size_t)d]; }
                    kr.scatter(kouts.data(), iouts.data(), ws);
                    timer.lap("K6 keys (grouped by owner)");
                    std::vector<u64> m8(mat), m4(mat);
                    for (auto &x : m8) x *= 8;
                    for (auto &x : m4) x *= sizeof(DP);
                    sh->comm->all_to_all_v(skey.get(), key.get(), m8.data(), st);
                    sh->comm->all_to_all_v(sidx.get(), sa.get(), m4.data(), st);
                    timer.lap("K6 key exchange (all-to-all)");
                }
            }
            key_b.alloc(N); val_b.alloc(padded(N));
            sort_pairs(key, key_b, sa, val_b, N, KEY_CODE_SHIFT, KEY_CODE_SHIFT + enum_bits(L), ws.sc, st, /*iota_vals=*/false);
        }
        if (n_sorted) *n_sorted = N;
        meta.alloc(padded(N) + 64); cnt.alloc(padded(N));
        dev_memset(meta.get() + padded(N), M_HEAD, 64, st);       // the flags pass below writes every slot of the tiles, padding included
        if (N == 0) { tm.dict_sort_rounds = 0; return PFB_OK; }
        timer.lap("K6 radix sort");
        FlagsArgsT<DP> fa;
        fa.key = key.get(); fa.sa = sa.get(); fa.D = N; fa.meta = meta.get(); fa.cnt = cnt.get(); fa.tilecnt = nullptr;
        fa.kx = kx; fa.wl = wl; fa.w = (u64)w;
        m = sorted_flags_and_open_list(fa, apos, ws);
        if (m) {
            adep.alloc(m);
            const u32 *ap = apos.get(); const u64 *ks = key.get(); u32 *dp = adep.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) { dp[a] = key_consumed(ks[ap[a]]); });
        }
    }
    timer.lap("K6 flags+open ties");
    u32 rounds = 0;
    // Extension rounds go on while they make progress: a big open set after a few rounds means long
    // repeats (prefix doubling, O(log) rounds); a handful of stragglers (e.g. the Dollar^w tail, one
    // stopper per round) is cheaper to finish here than to hand all D suffixes to doubling.
    u32 max_ext_rounds = 4, max_small_rounds = 256;
    bool knob = false;
    if (const char *e = getenv("PFB_EXT_ROUNDS")) { max_ext_rounds = (u32)atoi(e); knob = true; }   // test knob: 0 = doubling only
    const u64 small_set = D / 2048 > 4096 ? D / 2048 : 4096;
    u8 *mt = meta.get();
    // a rank that holds only its key range cannot switch to prefix doubling (rank[] of all suffixes):
    // it keeps extending; phrases of ordinary length close within a few dozen rounds
    const u32 shard_rounds = 1u << 16;
    // pivot rounds take over when a large set is still open after two window rounds (long ties); they stop
    // when many comparisons hit the cap (self-overlapping repeats such as poly-A: prefix doubling is the tool)
    int pivot_knob = -1;
    if (const char *e = getenv("PFB_PIVOT")) pivot_knob = atoi(e);         // test knob: 1 = pivot rounds from the start, 0 = never
    bool pivot_ok = pivot_knob != 0;
    u32 pivot_rounds = 0;
    const u32 max_pivot_rounds = 64;
    // ... right away when a large part of the dictionary is tied after the main sort (repetitive dictionary), else
    // after two window rounds (short random ties are cheaper to close with windows)
    RunTableT<DP> rt;                  // built before the first pivot round
    DevBuf<DP> run_s, run_e;
    bool have_rt = false;
    int lb = PIVOT_LBITS;
    auto want_pivot = [&]() { return pivot_ok && pivot_rounds < max_pivot_rounds && (pivot_knob == 1 || rounds >= 2 || m > N / 16); };
    while (m && (want_pivot() || rounds < max_ext_rounds || (!knob && m <= small_set && rounds < max_small_rounds) || (sh && rounds < shard_rounds))) {
        const bool pivot = want_pivot();
        // dense group ids inside the active list (whole groups are active, so the
        // first listed member of a group is its head)
        DevBuf<u8> ah(m);
        DevBuf<u32> gidx(m);
        {
            const u32 *ap = apos.get(); u8 *a_ = ah.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) { a_[a] = (mt[ap[a]] & M_HEAD) ? 1 : 0; });
        }
        u64 G = scan_flags(ah.get(), gidx.get(), m, ws);
        timer.lap("K6 round: group ids");
        int gbits = bits_for(G ? G - 1 : 0);
        int L2 = KEY_MAX_SYMS;
        while (L2 > 0 && 6 + enum_bits(L2) + gbits > 64) --L2;
        if (L2 < 1) break;   // cannot happen with D < 2^32; doubling below takes over
        const int ebits = enum_bits(L2);
        DevBuf<u64> k2(m), k2b(m);
        DevBuf<DP> v(m), vb(m);
        // what travels with a suffix (occurrence count, preceding-byte code, valid / full bits) is gathered once and follows
        // the permutation of the small groups instead of being looked up again (position -> word) after every round
        DevBuf<u32> tcnt(m), srcslot(m);
        DevBuf<u8> tmeta(m);
        {
            const u32 *ap = apos.get(); const u32 *cp = cnt.get(); u32 *tc = tcnt.get(); u8 *tm_ = tmeta.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) { const u32 j = ap[a]; tc[a] = cp[j]; tm_[a] = mt[j]; });
        }
        if (pivot && !have_rt && m < RUN_TABLE_MIN_OPEN) have_rt = true;     // see RUN_TABLE_MIN_OPEN
        if (pivot && !have_rt) {
            u32 R = 0; u64 longest = 0;
            build_run_table<DP>(dict, D, run_s, run_e, &R, &longest, ws);
            if (R) { rt.start = run_s.get(); rt.end = run_e.get(); rt.R = R; lb = bits_for(longest + 65536); if (lb < PIVOT_LBITS) lb = PIVOT_LBITS; }
            have_rt = true;
            timer.lap("K6 run table");
        }
        if (pivot && gbits + lb + 6 > 64) break;               // cannot happen with D < 2^32; doubling below takes over
        const int kshift = pivot ? 1 : 6;                       // bits below the part of the key that orders / separates
        const int kbits = pivot ? lb + 6 : 6 + ebits;    // bits below the group id
        DevBuf<u32> clamped(1);
        if (pivot) {
            dev_memset(clamped.get(), 0, 4, st);
            DevBuf<u32> hslot(G + 1);                           // list slot of every group's first member
            const u32 *ap = apos.get(); const u8 *a_ = ah.get(); const u32 *gi = gidx.get(); const DP *sp = sa.get();
            const u32 *dp = adep.get();
            u64 *kp = k2.get(); DP *vp = v.get(); u32 *hs = hslot.get(); u32 *cl = clamped.get();
            const int lbv = lb; const RunTableT<DP> rtv = rt;
            launch_for(m, st, PFB_LAMBDA(u64 a) { if (a_[a]) hs[gi[a]] = (u32)a; });
            launch_for(m, st, PFB_LAMBDA(u64 a) {
                u64 gid = (u64)gi[a] + a_[a] - 1;
                DP idx = sp[ap[a]];
                vp[a] = idx;
                u32 h = hs[gid];
                u64 pk = pivot_key(dict, D, (u64)idx + dp[a], (u64)sp[ap[h]] + dp[h], lbv, rtv);
                if (!(pk & 1) && ((pk >> (lbv + 4)) & 3) == 1) atomic_add(cl, 1u);
                kp[a] = (gid << (lbv + 6)) | pk;
            });
        } else {
            const u32 *ap = apos.get(); const u8 *a_ = ah.get(); const u32 *gi = gidx.get(); const DP *sp = sa.get();
            const u32 *dp = adep.get();
            u64 *kp = k2.get(); DP *vp = v.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) {
                u64 gid = (u64)gi[a] + a_[a] - 1;
                DP idx = sp[ap[a]];
                vp[a] = idx;
                EnumKey e = enum_scan(dict, D, (u64)idx + dp[a], L2);
                kp[a] = (gid << (6 + ebits)) | (e.code << 6) | ((u64)e.consumed << 1) | (e.closed ? 1ull : 0ull);
            });
        }
        timer.lap(pivot ? "K6 round: pivot keys" : "K6 round: window keys");
        // Order every group by its members' keys.  Tie groups are SMALL where this loop matters (a pangenome dictionary: the
        // variants of one phrase, a handful of members, but 10^8 of them), so a radix sort of the whole list - six passes
        // over (key, value) pairs - is the wrong tool: every member of a group of at most SMALL_GROUP members ranks itself by
        // counting the smaller keys of its group (one pass, the group's keys come from L1/L2).  Larger groups (poly-A runs,
        // tandem repeats) are compacted and take the radix sort as before.
        DevBuf<u8> hl(m);                                       // 1 = first, 2 = last member with its key inside the group
        {
            static const u32 SMALL_GROUP = getenv("PFB_SMALL_GROUP") ? (u32)atoi(getenv("PFB_SMALL_GROUP")) : 48u;   // test knob: 0 = radix sort only
            DevBuf<u32> hslot2(G + 1);
            DevBuf<u8> bigf(m);
            {
                const u8 *a_ = ah.get(); const u32 *gi = gidx.get(); u32 *hs = hslot2.get();
                const u32 m32 = (u32)m; const u64 G_ = G;
                launch_for(m + 1, st, PFB_LAMBDA(u64 a) { if (a == m32) hs[G_] = m32; else if (a_[a]) hs[gi[a]] = (u32)a; });
                const u64 *kp = k2.get(); const DP *vp = v.get(); u64 *ko = k2b.get(); DP *vo = vb.get(); u8 *hp = hl.get(); u8 *bf = bigf.get();
                u32 *srcp = srcslot.get();
                const u32 cap = SMALL_GROUP;
                launch_for(m, st, PFB_LAMBDA(u64 a) {
                    const u64 gid = (u64)gi[a] + a_[a] - 1;
                    const u32 s = hs[gid], e = hs[gid + 1];
                    if (e - s > cap) { bf[a] = 1; return; }
                    bf[a] = 0;
                    const u64 mine = kp[a] >> kshift;
                    u32 less = 0, eqb = 0, eqa = 0;
                    for (u32 j = s; j < e; ++j) {
                        const u64 o = kp[j] >> kshift;
                        less += o < mine ? 1u : 0u;
                        if (o == mine) { if (j < (u32)a) ++eqb; else if (j > (u32)a) ++eqa; }
                    }
                    const u32 dest = s + less + eqb;               // stable: equal keys keep their order
                    ko[dest] = kp[a]; vo[dest] = vp[a]; srcp[dest] = (u32)a;
                    hp[dest] = (u8)((eqb == 0 ? 1 : 0) | (eqa == 0 ? 2 : 0));
                });
            }
            timer.lap("K6 round: rank inside small groups");
            DevBuf<u32> bscratch, bslots;
            const u64 mb = compact_indices(bigf.get(), m, bscratch, bslots, ws);
            if (mb) {
                DevBuf<u64> bk(mb), bkb(mb);
                DevBuf<DP> bv(mb), bvb(mb);
                {
                    const u32 *bs = bslots.get(); const u64 *kp = k2.get(); const DP *vp = v.get(); u64 *bkp = bk.get(); DP *bvp = bv.get();
                    launch_for(mb, st, PFB_LAMBDA(u64 i) { bkp[i] = kp[bs[i]]; bvp[i] = vp[bs[i]]; });
                }
                sort_pairs(bk, bkb, bv, bvb, mb, kshift, kbits + gbits, ws.sc, st);
                {
                    // whole groups were taken and the group id leads the key: the i-th sorted pair belongs into the i-th slot
                    const u32 *bs = bslots.get(); const u64 *bkp = bk.get(); const DP *bvp = bv.get(); u64 *ko = k2b.get(); DP *vo = vb.get(); u8 *hp = hl.get();
                    u32 *srcp = srcslot.get();
                    launch_for(mb, st, PFB_LAMBDA(u64 i) {
                        const u64 k = bkp[i] >> kshift;
                        const bool head = i == 0 || k != (bkp[i - 1] >> kshift);
                        const bool last = i + 1 == mb || k != (bkp[i + 1] >> kshift);
                        const u32 dest = bs[i];
                        ko[dest] = bkp[i]; vo[dest] = bvp[i]; srcp[dest] = 0xFFFFFFFFu;      // the sort does not say where a pair came from: looked up again
                        hp[dest] = (u8)((head ? 1 : 0) | (last ? 2 : 0));
                    });
                }
            }
            k2.swap(k2b); v.swap(vb);
            timer.lap("K6 round: large groups (radix sort)");
        }
        if (pivot) {
            ++pivot_rounds;
            if ((u64)read_scalar(clamped.get(), st) * 8 > m) pivot_ok = false;      // mostly self-overlapping repeats
        }
        DevBuf<u8> act2(m);
        {
            const u32 *ap = apos.get(); const u64 *kp = k2.get(); const DP *vp = v.get(); const u8 *hp = hl.get();
            DP *sp = sa.get(); u8 *ac = act2.get(); u32 *dp = adep.get(); u32 *cp = cnt.get();
            const WordLookupT<DP> w3 = wl; const u64 w64 = (u64)w; const int lbw = lb;
            const u32 *srcp = srcslot.get(); const u32 *tc = tcnt.get(); const u8 *tm_ = tmeta.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) {
                const bool head = hp[a] & 1, last = (hp[a] & 2) != 0;
                u32 j = ap[a];
                sp[j] = vp[a];
                const u32 from = srcp[a];
                u32 c, mb_;                                                // what travels with the suffix: count, valid/full bits,
                if (from != 0xFFFFFFFFu) { c = tc[from]; mb_ = tm_[from] & (u32)(0x0f | M_VALID | M_FULL); }      // and the preceding-byte code
                else { u32 ib, sl; const u32 vf = suffix_class<DP>(w3, vp[a], w64, c, ib, sl); mb_ = (u32)prev_code(dict, vp[a]) | vf; }
                cp[j] = c;
                mt[j] = (u8)(mb_ | (head ? M_HEAD : 0));
                dp[a] += pivot ? pivot_consumed(kp[a], lbw) : ((u32)(kp[a] >> 1) & 31u);   // slot a stays inside its old group: same depth so far
                ac[a] = (!(head && last) && !(kp[a] & 1)) ? 1 : 0;
            });
        }
        timer.lap("K6 round: write back + classify");
        DevBuf<u32> scratch, sel, apos2, adep2;
        u64 m2 = compact_indices(act2.get(), m, scratch, sel, ws);
        apos2.alloc(m2); adep2.alloc(m2);
        if (m2) {
            const u32 *sl = sel.get(); const u32 *ap = apos.get(); const u32 *dp = adep.get(); u32 *a2 = apos2.get(); u32 *d2 = adep2.get();
            launch_for(m2, st, PFB_LAMBDA(u64 a) { a2[a] = ap[sl[a]]; d2[a] = dp[sl[a]]; });
        }
        apos = std::move(apos2);
        adep = std::move(adep2);
        m = m2;
        ++rounds;
        timer.lap(pivot ? "K6 round: compact open list (pivot round ends)" : "K6 round: compact open list (extension round ends)");
    }
    if (m && sh) { err = "dictionary repeats too long for the range-partitioned suffix sort (use the single-GPU path)"; return PFB_ERR_LIMIT; }
    if (m) {
        // long repeats: hand the partially sorted array to prefix doubling.  Every open group is
        // sorted on at least h = min depth symbols, which is all the doubling step needs.
        DevBuf<u32> hmin(1);
        dev_memset(hmin.get(), 0xff, 4, st);
        {
            const u32 *dp = adep.get(); u32 *hm = hmin.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) { atomic_min(hm, dp[a]); });
        }
        u64 h = read_scalar(hmin.get(), st);
        apos.release(); adep.release();
        DevBuf<u32> rank(D);
        DevBuf<u8> fl(D);
        {
            u8 *f = fl.get();
            launch_for(D, st, PFB_LAMBDA(u64 j) { f[j] = (mt[j] & M_HEAD) ? 1 : 0; });
        }
        // prefix doubling runs over ALL D < 2^32 suffixes (never inside a key range): 32-bit positions, whatever DP is
        DevBuf<u32> sa_tmp;
        DevBuf<u32> &sa32 = sa_as_u32(sa, sa_tmp, D, st);
        ranks_from_flags(sa32, fl, rank, D, ws);
        const WordLookupT<DP> w2 = wl;
        u32 dr = 0;
        int rc = prefix_doubling(sa32, rank, fl, D, h, PFB_LAMBDA(u32 idx, u64 hh) {
            return w2.suflen(w2.word((DP)idx), (DP)idx) < hh;   // EndOfWord already inside the compared prefix
        }, ws, &dr, err);
        if (rc) return rc;
        sa_from_u32(sa, sa_tmp, D, st);
        rounds += dr;
        const DP *sp = sa.get(); const u8 *f = fl.get();
        launch_for(D, st, PFB_LAMBDA(u64 j) { mt[j] = (u8)(prev_code(dict, sp[j]) | (f[j] ? M_HEAD : 0)); });
        *reclassify = true;
        timer.lap("K6 prefix doubling");
    }
    tm.dict_sort_rounds = rounds;
    return PFB_OK;
}

// number of positions x in [0, len) with b[x] != b[x-1] (b[-1] = 0): the run count r of src/pfbwt-f.cpp:304-346.
// 16 bytes per thread, compared with the same bytes shifted by one position (SWAR), one atomic per CTA.
#ifndef PFB_EMU
__global__ void __launch_bounds__(256) k_count_runs_kernel(const u8 *__restrict__ b, u64 len, u64 *__restrict__ out) {
    __shared__ u32 wsum[8];
    u32 cnt = 0;
    const u64 nchunks = (len + 15) / 16;
    for (u64 c = (u64)blockIdx.x * 256 + threadIdx.x; c < nchunks; c += (u64)gridDim.x * 256) {
        const u64 base = c * 16;
        u32 prev = base ? (u32)b[base - 1] : 0u;
        if (base + 16 <= len) {
            const uint4 v = *reinterpret_cast<const uint4 *>(b + base);            // the buffer is 16-byte aligned
            const u32 w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const u32 d = w4[k] ^ ((w4[k] << 8) | prev);
                cnt += (u32)__popc((d | ((d & 0x7f7f7f7fu) + 0x7f7f7f7fu)) & 0x80808080u);   // non-zero bytes
                prev = w4[k] >> 24;
            }
        } else {
            for (u64 x = base; x < len; ++x) { cnt += b[x] != prev ? 1u : 0u; prev = b[x]; }
        }
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        u64 t = 0;
        for (int k = 0; k < 8; ++k) t += wsum[k];
        if (t) atomicAdd((unsigned long long *)out, (unsigned long long)t);
    }
}
#endif
inline u64 count_runs(const u8 *b, u64 len, Workspace &ws) {
#ifdef PFB_EMU
    (void)ws;
    u64 r = 0;
    for (u64 x = 0; x < len; ++x) r += b[x] != (x ? b[x - 1] : (u8)0) ? 1 : 0;
    return r;
#else
    DevBuf<u64> out(1);
    dev_memset(out.get(), 0, 8, ws.st);
    k_count_runs_kernel<<<148 * 8, 256, 0, ws.st>>>(b, len, out.get());
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
    return read_scalar(out.get(), ws.st);
#endif
}

// what the emission kernels look positions up with: 32-bit positions and their tables, or nothing (ibase[] is filled)
inline void set_emit_lookup(EmitArgs &ea, const u32 *sa, const WordLookupT<u32> &wl) { ea.sa = sa; ea.wl = wl; }
inline void set_emit_lookup(EmitArgs &ea, const u64 *, const WordLookupT<u64> &) { ea.sa = nullptr; ea.wl = WordLookupT<u32>(); }

template <int U>
PFB_HD inline void store_sa(u8 *sa_out, u64 x, u64 v) {
    if (U == 4) reinterpret_cast<u32 *>(sa_out)[x] = (u32)v;
    else reinterpret_cast<u64 *>(sa_out)[x] = v;
}

// ===========================================================================
// build_bwt (include/pfbwt.hpp:96-223, src/pfbwt-f.cpp:298-328)
// ===========================================================================
// bucket / carry: ONE GPU builds the BWT one key range after the other (rank `bucket->rank` of `bucket->world` ranges, no
// communicator) so that the per-suffix work arrays are 1/world of the full ones: the -m mode.  *carry
// hands what a later bucket needs from the earlier ones.
struct BucketCarry {
    u64 off = 0;                               // BWT positions emitted so far
    u64 r = 0, ssa = 0, esa = 0;               // runs, .ssa pairs, .esa pairs so far
    u64 easy = 0, hard = 0;
    bool have_last = false; u8 last_byte = 0; u64 last_sa = 0;     // the last position emitted: the next bucket decides whether it ends a run
};
// word tables of the position -> word lookup: (start | first ilist slot - 1 << 32) packed for 32-bit positions, two arrays for 64-bit
struct WordTables { DevBuf<u64> wc; DevBuf<u32> cm; DevBuf<u32> wsamp; };
inline WordLookupT<u32> make_word_lookup(WordTables &t, const DeviceParse &pp, u32 tag, Workspace &ws) {
    (void)tag;
    pfbStream_t st = ws.st;
    const u64 D = pp.dsize, Dw = pp.Dw;
    t.wc.alloc(Dw + 2); t.wsamp.alloc((D + 127) / 128);
    DevBuf<u32> cum(Dw + 1);
    exclusive_sum<u32, u32>(pp.occ.get(), cum.get(), Dw, ws.sc, st);
    const u32 *occ = pp.occ.get(); const u64 *wstart = pp.wstart.get(); u32 *cm = cum.get(); u64 *wp = t.wc.get();
    launch_for(Dw + 2, st, PFB_LAMBDA(u64 d) {
        u64 c = d < Dw ? cm[d] : (Dw ? (u64)cm[Dw - 1] + occ[Dw - 1] : 0);
        u64 s = d <= Dw ? wstart[d] : 0xFFFFFFFFull;
        wp[d] = s | (c << 32);
    });
    u32 *wsm = t.wsamp.get();
    const u32 Dw32 = (u32)Dw;
    launch_for((D + 127) / 128, st, PFB_LAMBDA(u64 b) { wsm[b] = word_of(wstart, Dw32 + 1, b * 128); });
    WordLookupT<u32> wl;
    wl.wc = t.wc.get(); wl.wsamp = t.wsamp.get(); wl.Dw = (u32)Dw;
    return wl;
}
inline WordLookupT<u64> make_word_lookup(WordTables &t, const DeviceParse &pp, u64 tag, Workspace &ws) {
    (void)tag;
    pfbStream_t st = ws.st;
    const u64 D = pp.dsize, Dw = pp.Dw;
    t.wc.alloc(Dw + 2); t.cm.alloc(Dw + 2); t.wsamp.alloc((D + 127) / 128);
    DevBuf<u32> cum(Dw + 1);
    exclusive_sum<u32, u32>(pp.occ.get(), cum.get(), Dw, ws.sc, st);
    const u32 *occ = pp.occ.get(); const u64 *wstart = pp.wstart.get(); const u32 *cs = cum.get(); u64 *wp = t.wc.get(); u32 *cp = t.cm.get();
    launch_for(Dw + 2, st, PFB_LAMBDA(u64 d) {
        cp[d] = d < Dw ? cs[d] : (Dw ? cs[Dw - 1] + occ[Dw - 1] : 0u);
        wp[d] = d <= Dw ? wstart[d] : ~0ull;
    });
    u32 *wsm = t.wsamp.get();
    const u32 Dw32 = (u32)Dw;
    launch_for((D + 127) / 128, st, PFB_LAMBDA(u64 b) { wsm[b] = word_of(wstart, Dw32 + 1, b * 128); });
    WordLookupT<u64> wl;
    wl.ws = t.wc.get(); wl.cm = t.cm.get(); wl.wsamp = t.wsamp.get(); wl.Dw = (u32)Dw;
    return wl;
}

template <int U, class DP>
inline int build_bwt_t(const DeviceParse &pp, const Params &prm, bool want_sa, bool want_rssa, DeviceBwt &out, Workspace &ws,
                       StageTimes &tm, std::string &err, Comm *comm, const ShardSpec *bucket = nullptr, BucketCarry *carry = nullptr) {
    pfbStream_t st = ws.st;
    const u64 D = pp.dsize, Dw = pp.Dw, n = pp.n, P = pp.P;
    const u64 w = (u64)prm.w;
    const bool any_sa = want_sa || want_rssa;
    if (!pp.has_parse_bwt) { err = "parse BWT (bwlast/ilist) not available"; return PFB_ERR_ARG; }
    if (any_sa && pp.bwsai.size() == 0) { err = "bwsai not available"; return PFB_ERR_ARG; }
    if (sizeof(DP) == 4 && D > 0xFFFFFFF0ull) { err = "internal: 32-bit positions for a dictionary of 2^32 bytes or more"; return PFB_ERR_INTERNAL; }
    if (D > 0xFFFFFFF0ull && !comm && !(bucket && carry)) { err = "a dictionary of 2^32 bytes or more is built in key ranges (bucketed build)"; return PFB_ERR_LIMIT; }
    Timer timer(st);

    // word tables: wc[d] = (start, first ilist slot - 1), sample table for position -> word
    WordTables wtab;
    const WordLookupT<DP> wl = make_word_lookup(wtab, pp, DP(0), ws);
    timer.lap(sizeof(DP) == 8 ? "K7 word tables (64-bit dictionary positions)" : "K7 word tables");

    // K6 dictionary suffix sort -------------------------------------------------------
    DevBuf<DP> sa;
    DevBuf<u8> meta;
    DevBuf<u32> cnt, ibase, slen;
    const bool wide = sizeof(DP) == 8;     // the emission kernels never see 64-bit positions: ibase[] replaces their lookup
    bool reclassify = false;
    ShardSpec spec;
    if (comm) { spec.rank = comm->rank; spec.world = comm->world; spec.comm = comm; }
    else if (bucket && carry) { spec = *bucket; spec.comm = nullptr; }
    else carry = nullptr;
    const bool ranged = comm || carry;
    u64 N = D;                          // sorted entries held by this rank (all D without a communicator)
    int rc;
    try {
        rc = sort_dict_suffixes<DP>(pp, (int)w, sa, meta, cnt, wl, &reclassify, ws, tm, timer, err, ranged ? &spec : nullptr, &N);
    } catch (const Error &e) {
        if (!comm) throw;
        rc = e.code; err = e.what();
    }
    if (comm) {
        // a failure on one rank (a tie its key range cannot close, out of memory) must not leave the others waiting in the
        // next exchange: agree on it here
        u64 f = rc ? 1 : 0;
        std::vector<u64> all = comm->gather_u64(&f, 1, st);
        for (int r = 0; r < comm->world && !rc; ++r)
            if (all[(size_t)r]) { err = "rank " + std::to_string(r) + " failed in the dictionary suffix sort"; rc = PFB_ERR_ARG; }
    }
    if (rc) return rc;
    tm.dict_sort = timer.stop();

    // K7 tile offsets, hard groups ------------------------------------------------------
    // Without SA output everything the emission needs per suffix came with the sort (meta, cnt);
    // with SA output (or after prefix doubling) a lookup pass adds ilist base and suffix length.
    const u64 nt = tiles_of(N);
    DevBuf<u64> tilesum(nt + 1), tilebase(nt + 1);
    if (any_sa || wide) { ibase.alloc(padded(N)); slen.alloc(padded(N)); }
    if (N == 0) {
    } else if (any_sa || reclassify || wide) {
        ClassifyArgsT<DP> ca;
        ca.sa = sa.get(); ca.meta = meta.get(); ca.wl = wl; ca.D = N; ca.w = w;
        ca.cnt = cnt.get(); ca.tilesum = tilesum.get(); ca.ibase = ibase.get(); ca.sl = slen.get();
        k_classify<DP>(ca, any_sa || wide, st);
        timer.lap("K7 classify (word lookup)");
    } else {
        k_tile_sums(cnt.get(), N, tilesum.get(), st);
        timer.lap("K7 tile sums");
    }
    dev_memset(tilesum.get() + nt, 0, 8, st);
    exclusive_sum<u64, u64>(tilesum.get(), tilebase.get(), nt + 1, ws.sc, st);
    const u64 total = read_scalar(tilebase.get() + nt, st);     // BWT positions emitted by this rank
    u64 slice_off = 0;
    std::vector<u64> totals(1, total);
    if (comm) {
        // the ranks' key ranges are consecutive, so their outputs are consecutive slices of the BWT
        totals = comm->gather_u64(&total, 1, st);
        for (int r = 0; r < comm->rank; ++r) slice_off += totals[(size_t)r];
    }
    if (carry) slice_off = carry->off;
    if (!carry || spec.rank + 1 == spec.world) {
        u64 sum = carry ? carry->off : 0;
        for (u64 t_ : totals) sum += t_;
        if (sum != n + 1) { err = "internal: emitted positions " + std::to_string(sum) + " != n+1 = " + std::to_string(n + 1); return PFB_ERR_INTERNAL; }
    }
    const u64 nhg_cap = N ? k_group_classes(meta.get(), N, any_sa, ws) : 0;   // number of hard groups (sizes the work list)
    timer.lap("K7 group classes");
    tm.groups = timer.stop();

    // K8 emission ---------------------------------------------------------------------
    // (one pass through a do-while so that a failure on one rank - an exception included - is agreed on by all ranks before
    // the next exchange instead of leaving them waiting in it)
    rc = PFB_OK;
    try { do {
        out.n = n; out.off = slice_off; out.len = total;
        out.bwt.alloc(total + 16);
        if (any_sa) out.sa.alloc(total * (u64)U + 16);
        DevBuf<u32> hard_t(nhg_cap + 1), ctr32(2);
        DevBuf<u64> hard_x(nhg_cap + 1), ctr64(1);
        dev_memset(ctr32.get(), 0, 8, st);
        dev_memset(ctr64.get(), 0, 8, st);
        EmitArgs ea;
        ea.meta = meta.get(); ea.cnt = cnt.get(); ea.ibase = ibase.get(); ea.sl = slen.get();
        set_emit_lookup(ea, sa.get(), wl);
        ea.tilebase = tilebase.get(); ea.D = N;
        ea.ilist = pp.ilist.get(); ea.bwlast = pp.bwlast.get(); ea.bwsai = pp.bwsai.get();
        ea.bwt = out.bwt.get(); ea.sa_out = out.sa.get();
        ea.small_lim = HARD_SMALL; ea.mid_lim = HARD_MID;
        if (const char *e = getenv("PFB_HARD_LIMITS")) {   // test knob "small,mid": push groups down the larger-group paths
            unsigned a_ = HARD_SMALL, b_ = HARD_MID;
            if (sscanf(e, "%u,%u", &a_, &b_) == 2) { ea.small_lim = a_ < HARD_SMALL ? a_ : HARD_SMALL; ea.mid_lim = b_ < HARD_MID ? b_ : HARD_MID; }
        }
        ea.hard_t = hard_t.get(); ea.hard_x = hard_x.get(); ea.n_hard_groups = ctr32.get(); ea.n_hard_occ = ctr64.get();
    #ifdef PFB_EMU
        if (N) emu_emit<U>(ea, any_sa);
    #else
        if (N) {
            if (any_sa) {
                static bool sa_attr_set = false;
                if (!sa_attr_set) {
                    PFB_CUDA_CHECK(cudaFuncSetAttribute(k_emit_sa_kernel<U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EmitSaSmem<U>::BYTES));
                    sa_attr_set = true;
                }
                k_emit_sa_kernel<U><<<(unsigned)nt, 256, EmitSaSmem<U>::BYTES, st>>>(ea);
            } else k_emit_kernel<U, false><<<(unsigned)nt, 256, 0, st>>>(ea);
            PFB_CUDA_CHECK(cudaGetLastError());
            launch_stats().own++;
        }
    #endif
        timer.lap("K8 emit easy");
        const u64 nhg = read_scalar(ctr32.get(), st);
        const u64 nhard = read_scalar(ctr64.get(), st);
        if (nhg != nhg_cap) { err = "internal: hard group count mismatch"; rc = PFB_ERR_INTERNAL; break; }
        if (nhg) {
            DevBuf<u32> biglist(nhg);
            u32 *nbigp = ctr32.get() + 1;
    #ifdef PFB_EMU
            emu_emit_hard<U>(ea, any_sa, (u32)nhg, biglist.get(), nbigp);
    #else
            const unsigned blocks = (unsigned)((nhg + 31) / 32);
            if (any_sa) k_emit_hard_kernel<U, true><<<blocks, 256, 0, st>>>(ea, (u32)nhg, biglist.get(), nbigp);
            else k_emit_hard_kernel<U, false><<<blocks, 256, 0, st>>>(ea, (u32)nhg, biglist.get(), nbigp);
            PFB_CUDA_CHECK(cudaGetLastError());
            launch_stats().own++;
    #endif
            timer.lap("K8 hard small groups");
            const u64 nmid = read_scalar(nbigp, st);
            if (nmid) {
                DevBuf<u32> biglist2(nmid), nbig2(1);
                dev_memset(nbig2.get(), 0, 4, st);
    #ifdef PFB_EMU
                emu_emit_hard_mid<U>(ea, any_sa, biglist.get(), (u32)nmid, biglist2.get(), nbig2.get());
    #else
                static bool attr_set = false;
                if (!attr_set) {
                    PFB_CUDA_CHECK(cudaFuncSetAttribute(k_emit_hard_mid_kernel<U, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MID_SMEM));
                    PFB_CUDA_CHECK(cudaFuncSetAttribute(k_emit_hard_mid_kernel<U, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MID_SMEM));
                    attr_set = true;
                }
                const unsigned mblocks = (unsigned)((nmid + 7) / 8);
                if (any_sa) k_emit_hard_mid_kernel<U, true><<<mblocks, 256, MID_SMEM, st>>>(ea, biglist.get(), (u32)nmid, biglist2.get(), nbig2.get());
                else k_emit_hard_mid_kernel<U, false><<<mblocks, 256, MID_SMEM, st>>>(ea, biglist.get(), (u32)nmid, biglist2.get(), nbig2.get());
                PFB_CUDA_CHECK(cudaGetLastError());
                launch_stats().own++;
    #endif
                timer.lap("K8 hard mid groups");
                const u64 nbig = read_scalar(nbig2.get(), st);
                if (nbig) {
                    rc = emit_big_groups<U>(ea, any_sa, biglist2.get(), nbig, P, ws, err);
                    if (rc) break;
                }
            }
            timer.lap("K8 hard big groups");
        }
        out.hard = nhard;                 // "# hard cases" of include/pfbwt.hpp:188
        out.easy = total - nhard;
        u8 *sa_out0 = out.sa.get();
        if (any_sa && slice_off == 0 && total) launch_for(1, st, PFB_LAMBDA(u64) { store_sa<U>(sa_out0, 0, n); });   // src/pfbwt-f.cpp:301
        sa.release(); meta.release(); cnt.release(); ibase.release(); slen.release();
    } while (0); }
    catch (const Error &e) { if (!comm) throw; rc = e.code; err = e.what(); }
    if (comm) {
        u64 f = rc ? 1 : 0;
        std::vector<u64> all = comm->gather_u64(&f, 1, st);
        for (int r = 0; r < comm->world && !rc; ++r)
            if (all[(size_t)r]) { err = "rank " + std::to_string(r) + " failed in the BWT assembly"; rc = PFB_ERR_ARG; }
    }
    if (rc) return rc;
    u8 *bwt = out.bwt.get();
    u8 *sa_out = out.sa.get();
    tm.emit = timer.stop();

    // K9 run boundaries + run-length SA samples (src/pfbwt-f.cpp:304-328) -------------------
    if (comm || carry) {
        // Slices meet at rank (or bucket) boundaries: a run start needs the byte before the slice, a run end the
        // byte after it.  Ranks: one small all-gather of (first byte, last byte) per rank supplies both.  Buckets come
        // one after the other: the byte before is carried over, and whether a bucket's last position ends a run is decided
        // by the next bucket, which then writes that .esa pair first (`pend`).
        u8 prevb = 0, nextb = 0;
        bool has_next = false, defer_last = false;
        u64 pend = 0, pend_pos = 0, pend_sa = 0;
        if (comm) {
            u64 edge[2] = {0, 0};
            if (total) { edge[0] = read_scalar(bwt, st); edge[1] = read_scalar(bwt + (total - 1), st); }
            std::vector<u64> edges = comm->gather_u64(edge, 2, st);
            for (int r = comm->rank - 1; r >= 0; --r) if (totals[(size_t)r]) { prevb = (u8)edges[2 * (size_t)r + 1]; break; }
            for (int r = comm->rank + 1; r < comm->world; ++r) if (totals[(size_t)r]) { nextb = (u8)edges[2 * (size_t)r]; has_next = true; break; }
        } else {
            prevb = carry->have_last ? carry->last_byte : (u8)0;
            defer_last = spec.rank + 1 < spec.world;
            const u8 first = total ? read_scalar(bwt, st) : (u8)0;
            if (carry->have_last && ((total && first != carry->last_byte) || (!total && !defer_last))) { pend = 1; pend_pos = carry->off - 1; pend_sa = carry->last_sa; }
        }
        DevBuf<u8> rf(total + 1), ef(total + 1);
        u8 *rp = rf.get(); u8 *ep = ef.get();
        launch_for(total, st, PFB_LAMBDA(u64 x) {
            rp[x] = bwt[x] != (x ? bwt[x - 1] : prevb) ? 1 : 0;
            ep[x] = (x + 1 < total) ? (bwt[x + 1] != bwt[x] ? 1 : 0) : (defer_last ? 0 : ((!has_next || nextb != bwt[x]) ? 1 : 0));
        });
        DevBuf<u64> ridx, eidx;
        u64 cnts[2] = {0, 0};
        if (!want_rssa) {
            cnts[0] = cnts[1] = total ? reduce_sum<u8, u64>(rf.get(), total, ws.sc, st) : 0;
        } else if (total) {
            ridx.alloc(total); eidx.alloc(total);
            exclusive_sum<u8, u64>(rf.get(), ridx.get(), total, ws.sc, st);
            exclusive_sum<u8, u64>(ef.get(), eidx.get(), total, ws.sc, st);
            cnts[0] = read_scalar(ridx.get() + (total - 1), st) + read_scalar(rf.get() + (total - 1), st);
            cnts[1] = read_scalar(eidx.get() + (total - 1), st) + read_scalar(ef.get() + (total - 1), st);
        }
        u64 r = 0, so = 0, eo = 0;
        if (comm) {
            std::vector<u64> all = comm->gather_u64(cnts, 2, st);
            for (int q = 0; q < comm->world; ++q) {
                r += all[2 * (size_t)q];
                if (q < comm->rank) { so += all[2 * (size_t)q]; eo += all[2 * (size_t)q + 1]; }
            }
        } else { r = carry->r + cnts[0]; so = carry->ssa; eo = carry->esa; }
        out.r = r;
        if (want_rssa) {
            out.ssa_off = so; out.ssa_cnt = cnts[0]; out.esa_off = eo; out.esa_cnt = pend + cnts[1];
            out.ssa.alloc(2 * cnts[0] * (u64)U + 16);
            out.esa.alloc(2 * (pend + cnts[1]) * (u64)U + 16);
            u8 *ssa = out.ssa.get(); u8 *esa = out.esa.get(); const u64 *ri = ridx.get(); const u64 *ei = eidx.get();
            launch_for(total, st, PFB_LAMBDA(u64 x) {
                u64 v = U == 4 ? (u64)reinterpret_cast<const u32 *>(sa_out)[x] : reinterpret_cast<const u64 *>(sa_out)[x];
                if (rp[x]) { store_sa<U>(ssa, 2 * ri[x], slice_off + x); store_sa<U>(ssa, 2 * ri[x] + 1, v); }
                if (ep[x]) { store_sa<U>(esa, 2 * (ei[x] + pend), slice_off + x); store_sa<U>(esa, 2 * (ei[x] + pend) + 1, v); }
            });
            if (pend) launch_for(1, st, PFB_LAMBDA(u64) { store_sa<U>(esa, 0, pend_pos); store_sa<U>(esa, 1, pend_sa); });
        }
        if (carry) {
            carry->r = r; carry->ssa += cnts[0]; carry->esa += pend + cnts[1];
            carry->easy += out.easy; carry->hard += out.hard;
            if (total) {
                carry->have_last = true;
                carry->last_byte = read_scalar(bwt + (total - 1), st);
                if (any_sa) carry->last_sa = U == 4 ? (u64)read_scalar(reinterpret_cast<const u32 *>(sa_out) + (total - 1), st)
                                                    : read_scalar(reinterpret_cast<const u64 *>(sa_out) + (total - 1), st);
            }
            carry->off += total;
            out.easy = carry->easy; out.hard = carry->hard;
        }
    } else if (!want_rssa) {
        out.r = count_runs(bwt, n + 1, ws);
    } else {
        DevBuf<u8> rf(n + 1);
        u8 *rp = rf.get();
        launch_for(n + 1, st, PFB_LAMBDA(u64 x) { rp[x] = bwt[x] != (x ? bwt[x - 1] : (u8)0) ? 1 : 0; });
        {
            DevBuf<u64> ridx(n + 1);
            exclusive_sum<u8, u64>(rf.get(), ridx.get(), n + 1, ws.sc, st);
            u64 r = read_scalar(ridx.get() + n, st) + read_scalar(rf.get() + n, st);
            out.r = r;
            out.ssa.alloc(2 * r * (u64)U);
            out.esa.alloc(2 * r * (u64)U);
            u8 *ssa = out.ssa.get(); u8 *esa = out.esa.get(); const u64 *ri = ridx.get();
            launch_for(n + 1, st, PFB_LAMBDA(u64 x) {
                if (U == 4) {
                    const u32 *S4 = reinterpret_cast<const u32 *>(sa_out);
                    u32 *s4 = reinterpret_cast<u32 *>(ssa); u32 *e4 = reinterpret_cast<u32 *>(esa);
                    if (rp[x]) {
                        u64 k = ri[x];
                        s4[2 * k] = (u32)x; s4[2 * k + 1] = S4[x];
                        if (x) { e4[2 * (k - 1)] = (u32)(x - 1); e4[2 * (k - 1) + 1] = S4[x - 1]; }
                    }
                    if (x == n) { e4[2 * (r - 1)] = (u32)n; e4[2 * (r - 1) + 1] = S4[n]; }
                } else {
                    const u64 *S8 = reinterpret_cast<const u64 *>(sa_out);
                    u64 *s8 = reinterpret_cast<u64 *>(ssa); u64 *e8 = reinterpret_cast<u64 *>(esa);
                    if (rp[x]) {
                        u64 k = ri[x];
                        s8[2 * k] = x; s8[2 * k + 1] = S8[x];
                        if (x) { e8[2 * (k - 1)] = x - 1; e8[2 * (k - 1) + 1] = S8[x - 1]; }
                    }
                    if (x == n) { e8[2 * (r - 1)] = n; e8[2 * (r - 1) + 1] = S8[n]; }
                }
            });
        }
    }
    if (!comm && !carry) { out.ssa_off = out.esa_off = 0; out.ssa_cnt = out.esa_cnt = want_rssa ? out.r : 0; }
    timer.lap("K9 runs");
    tm.runs = timer.stop();
    trace_log().dump("build_bwt");
    return PFB_OK;
}

inline int build_bwt(const DeviceParse &pp, const Params &prm, bool want_sa, bool want_rssa, DeviceBwt &out, Workspace &ws,
                     StageTimes &tm, std::string &err, Comm *comm = nullptr, const ShardSpec *bucket = nullptr, BucketCarry *carry = nullptr) {
    // positions inside the dictionary: 32-bit below 2^32 bytes, 64-bit above (PFB_DICT64=1: test knob, 64-bit at any size)
    static const bool force64 = getenv("PFB_DICT64") && atoi(getenv("PFB_DICT64")) != 0;
    if (pp.dsize > 0xFFFFFFF0ull || force64)
        return prm.bits == 64 ? build_bwt_t<8, u64>(pp, prm, want_sa, want_rssa, out, ws, tm, err, comm, bucket, carry)
                              : build_bwt_t<4, u64>(pp, prm, want_sa, want_rssa, out, ws, tm, err, comm, bucket, carry);
    return prm.bits == 64 ? build_bwt_t<8, u32>(pp, prm, want_sa, want_rssa, out, ws, tm, err, comm, bucket, carry)
                          : build_bwt_t<4, u32>(pp, prm, want_sa, want_rssa, out, ws, tm, err, comm, bucket, carry);
}

}  // namespace pfb
