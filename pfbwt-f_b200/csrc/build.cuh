// build.cuh -- BWT build from a device-resident parse: dictionary suffix sort,
// group detection and segmented emission of .bwt/.sa/.ssa/.esa.
// Replaces PrefixFreeBWT::sort_dict_suffixes + generate_bwt_lcp
// (include/pfbwt.hpp:96-223) and the writer lambda of run_pfbwt
// (src/pfbwt-f.cpp:298-328).
//
// Data layout in HBM (D = dict bytes, Dw = words, P = phrases, n = |S|):
//   key[D]   u64  enumerative code of the first L symbols of the suffix, symbols consumed,
//                 closed flag, code of the preceding byte (layout: dictkeys.cuh)
//   sa[D]    u32  dictionary position of the j-th smallest suffix
//   fl[D]    u8   1 at the first member of every group of identical suffix strings
//   wc[Dw+2] u64  (word start | ilist base << 32): one 16-byte read gives both ends
//   wsamp[D/128] u32  word id at every 128th dictionary byte (position -> word)
//   cnt[D]   u32  occurrences emitted by the j-th suffix (0 when suffix length <= w)
//   off[D]   u64  exclusive sum of cnt = position in the BWT
#pragma once
#include "dictkeys.cuh"

namespace pfb {

// position -> (word, suffix length) through the sample table
struct WordLookup {
    const u64 *wc;      // Dw+2 entries, low 32 bits = word start; wc[Dw] = D-1, wc[Dw+1] = 2^32-1
    const u32 *wsamp;
    u32 Dw;
    PFB_HD inline u32 word(u32 idx) const {
        u32 d = wsamp[idx >> 7];
        while ((u32)wc[d + 1] <= idx) ++d;
        return d;
    }
    PFB_HD inline u32 start(u32 d) const { return (u32)wc[d]; }
    PFB_HD inline u32 cum(u32 d) const { return (u32)(wc[d] >> 32); }
    // length of the suffix starting at idx inside word d, EndOfWord excluded (0 for the EndOfDict slot)
    PFB_HD inline u64 suflen(u32 d, u32 idx) const { return d < Dw ? (u64)((u32)wc[d + 1] - 1 - idx) : 0; }
};

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
// ---------------------------------------------------------------------------
inline int sort_dict_suffixes(const DeviceParse &pp, DevBuf<u64> &key, DevBuf<u32> &sa, DevBuf<u8> &fl, const WordLookup &wl,
                              Workspace &ws, StageTimes &tm, Timer &timer, std::string &err) {
    pfbStream_t st = ws.st;
    const u64 D = pp.dsize;
    const u8 *dict = pp.dict.get();
    key.alloc(D); sa.alloc(D); fl.alloc(D);
    int L = 1;
    while (L < KEY_MAX_SYMS && (1ull << (2 * L)) < 128ull * D) ++L;     // ties become rare once 4^L >> D
    if (const char *e = getenv("PFB_KEY_SYMS")) { L = atoi(e); if (L < 1) L = 1; if (L > KEY_MAX_SYMS) L = KEY_MAX_SYMS; }   // test knob
    {
        DevBuf<u64> key_b(D);
        DevBuf<u32> val_b(D);
        k_suffix_keys(dict, D, L, key.get(), sa.get(), st);
        timer.lap("K6 keys");
        sort_pairs(key, key_b, sa, val_b, D, KEY_CODE_SHIFT, KEY_CODE_SHIFT + enum_bits(L), ws.sc, st);
        timer.lap("K6 radix sort");
    }
    DevBuf<u8> act(D);
    {
        const u64 *ks = key.get(); u8 *f = fl.get(); u8 *ac = act.get();
        launch_for(D, st, PFB_LAMBDA(u64 j) {
            u64 k = ks[j] >> KEY_CODE_SHIFT;
            bool head = j == 0 || k != (ks[j - 1] >> KEY_CODE_SHIFT);
            bool last = j + 1 == D || k != (ks[j + 1] >> KEY_CODE_SHIFT);
            f[j] = head ? 1 : 0;
            ac[j] = (!(head && last) && !key_closed(ks[j])) ? 1 : 0;     // tied and still open
        });
    }
    DevBuf<u32> apos, adep;
    u64 m;
    {
        DevBuf<u32> scratch;
        m = compact_indices(act.get(), D, scratch, apos, ws);
    }
    act.release();
    if (m) {
        adep.alloc(m);
        const u32 *ap = apos.get(); const u64 *ks = key.get(); u32 *dp = adep.get();
        launch_for(m, st, PFB_LAMBDA(u64 a) { dp[a] = key_consumed(ks[ap[a]]); });
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
    while (m && (rounds < max_ext_rounds || (!knob && m <= small_set && rounds < max_small_rounds))) {
        // dense group ids inside the active list (whole groups are active, so the
        // first listed member of a group is its head)
        DevBuf<u8> ah(m);
        DevBuf<u32> gidx(m);
        {
            const u32 *ap = apos.get(); const u8 *f = fl.get(); u8 *a_ = ah.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) { a_[a] = f[ap[a]]; });
        }
        u64 G = scan_flags(ah.get(), gidx.get(), m, ws);
        int gbits = bits_for(G ? G - 1 : 0);
        int L2 = KEY_MAX_SYMS;
        while (L2 > 0 && 6 + enum_bits(L2) + gbits > 64) --L2;
        if (L2 < 1) break;   // cannot happen with D < 2^32; doubling below takes over
        const int ebits = enum_bits(L2);
        DevBuf<u64> k2(m), k2b(m);
        DevBuf<u32> v(m), vb(m);
        {
            const u32 *ap = apos.get(); const u8 *a_ = ah.get(); const u32 *gi = gidx.get(); const u32 *sp = sa.get();
            const u32 *dp = adep.get();
            u64 *kp = k2.get(); u32 *vp = v.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) {
                u64 gid = (u64)gi[a] + a_[a] - 1;
                u32 idx = sp[ap[a]];
                vp[a] = idx;
                EnumKey e = enum_scan(dict, D, (u64)idx + dp[a], L2);
                kp[a] = (gid << (6 + ebits)) | (e.code << 6) | ((u64)e.consumed << 1) | (e.closed ? 1ull : 0ull);
            });
        }
        sort_pairs(k2, k2b, v, vb, m, 6, 6 + ebits + gbits, ws.sc, st);
        DevBuf<u8> act2(m);
        {
            const u32 *ap = apos.get(); const u64 *kp = k2.get(); const u32 *vp = v.get();
            u32 *sp = sa.get(); u8 *f = fl.get(); u8 *ac = act2.get(); u64 *ks = key.get(); u32 *dp = adep.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) {
                u64 k = kp[a] >> 6;
                bool head = a == 0 || k != (kp[a - 1] >> 6);
                bool last = a + 1 == m || k != (kp[a + 1] >> 6);
                u32 j = ap[a];
                sp[j] = vp[a];
                ks[j] = (ks[j] & ~15ull) | prev_code(dict, vp[a]);   // the preceding-byte code travels with the suffix
                f[j] = head ? 1 : 0;
                dp[a] += (u32)(kp[a] >> 1) & 31u;                      // slot a stays inside its old group: same depth so far
                ac[a] = (!(head && last) && !(kp[a] & 1)) ? 1 : 0;
            });
        }
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
        timer.lap("K6 extension round");
    }
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
        ranks_from_flags(sa, fl, rank, D, ws);
        const WordLookup w2 = wl;
        u32 dr = 0;
        int rc = prefix_doubling(sa, rank, fl, D, h, PFB_LAMBDA(u32 idx, u64 hh) {
            return w2.suflen(w2.word(idx), idx) < hh;   // EndOfWord already inside the compared prefix
        }, ws, &dr, err);
        if (rc) return rc;
        rounds += dr;
        const u32 *sp = sa.get(); u64 *ks = key.get();
        launch_for(D, st, PFB_LAMBDA(u64 j) { ks[j] = (ks[j] & ~15ull) | prev_code(dict, sp[j]); });
        timer.lap("K6 prefix doubling");
    }
    tm.dict_sort_rounds = rounds;
    return PFB_OK;
}

template <int U>
PFB_HD inline void store_sa(u8 *sa_out, u64 x, u64 v) {
    if (U == 4) reinterpret_cast<u32 *>(sa_out)[x] = (u32)v;
    else reinterpret_cast<u64 *>(sa_out)[x] = v;
}

// ===========================================================================
// build_bwt (include/pfbwt.hpp:96-223, src/pfbwt-f.cpp:298-328)
// ===========================================================================
template <int U>
inline int build_bwt_t(const DeviceParse &pp, const Params &prm, bool want_sa, bool want_rssa, DeviceBwt &out, Workspace &ws,
                       StageTimes &tm, std::string &err) {
    pfbStream_t st = ws.st;
    const u64 D = pp.dsize, Dw = pp.Dw, n = pp.n, P = pp.P;
    const u64 w = (u64)prm.w;
    const bool any_sa = want_sa || want_rssa;
    if (!pp.has_parse_bwt) { err = "parse BWT (bwlast/ilist) not available"; return PFB_ERR_ARG; }
    if (any_sa && pp.bwsai.size() == 0) { err = "bwsai not available"; return PFB_ERR_ARG; }
    if (D > 0xFFFFFFF0ull) { err = "dictionary larger than 2^32 bytes is not supported yet"; return PFB_ERR_LIMIT; }
    Timer timer(st);
    const u8 *dict = pp.dict.get();

    // word tables: wc[d] = (start, first ilist slot - 1), sample table for position -> word
    DevBuf<u64> wc(Dw + 2);
    DevBuf<u32> wsamp((D + 127) / 128);
    {
        DevBuf<u32> cum(Dw + 1);
        exclusive_sum<u32, u32>(pp.occ.get(), cum.get(), Dw, ws.sc, st);
        const u32 *occ = pp.occ.get(); const u32 *wstart = pp.wstart.get(); u32 *cm = cum.get(); u64 *wp = wc.get();
        launch_for(Dw + 2, st, PFB_LAMBDA(u64 d) {
            u64 c = d < Dw ? cm[d] : (Dw ? (u64)cm[Dw - 1] + occ[Dw - 1] : 0);
            u64 s = d <= Dw ? (u64)wstart[d] : 0xFFFFFFFFull;
            wp[d] = s | (c << 32);
        });
        u32 *wsm = wsamp.get();
        const u32 Dw32 = (u32)Dw;
        launch_for((D + 127) / 128, st, PFB_LAMBDA(u64 b) { wsm[b] = word_of(wstart, Dw32 + 1, (u32)(b * 128)); });
    }
    WordLookup wl;
    wl.wc = wc.get(); wl.wsamp = wsamp.get(); wl.Dw = (u32)Dw;

    // K6 dictionary suffix sort -------------------------------------------------------
    DevBuf<u64> key;
    DevBuf<u32> sa;
    DevBuf<u8> fl;
    timer.lap("K7 word tables");
    int rc = sort_dict_suffixes(pp, key, sa, fl, wl, ws, tm, timer, err);
    if (rc) return rc;
    tm.dict_sort = timer.stop();

    // K7 counts, offsets, groups ----------------------------------------------------
    DevBuf<u32> wordv(D), cnt(D);
    DevBuf<u64> off(D);
    {
        const u32 *sp = sa.get(); u32 *wv = wordv.get(); u32 *cp = cnt.get();
        launch_for(D, st, PFB_LAMBDA(u64 t) {
            u32 idx = sp[t];
            u32 d = wl.word(idx);
            bool valid = wl.suflen(d, idx) > w;                 // include/pfbwt.hpp:114
            wv[t] = d;
            cp[t] = valid ? wl.cum(d + 1) - wl.cum(d) : 0u;
        });
    }
    timer.lap("K7 classify (word lookup)");
    exclusive_sum<u32, u64>(cnt.get(), off.get(), D, ws.sc, st);
    timer.lap("K7 offsets scan");
    {
        u64 total = read_scalar(off.get() + (D - 1), st) + read_scalar(cnt.get() + (D - 1), st);
        if (total != n + 1) { err = "internal: emitted positions " + std::to_string(total) + " != n+1 = " + std::to_string(n + 1); return PFB_ERR_INTERNAL; }
    }
    // members of multi-word groups: head position, then "chars differ" per group (BWT-only mode)
    // cls[t]: 0 nothing to emit, 1 easy (direct), 2 hard (merge by ilist position)
    DevBuf<u8> cls(D);
    DevBuf<u32> head(D);
    {
        DevBuf<u32> hd(D);
        const u8 *f = fl.get(); u32 *hp = hd.get();
        launch_for(D, st, PFB_LAMBDA(u64 j) { hp[j] = f[j] ? (u32)j : 0; });
        inclusive_max(hd.get(), head.get(), D, ws.sc, st);
    }
    timer.lap("K7 group heads");
    {
        DevBuf<u8> gdiff;
        if (!any_sa) {
            gdiff.alloc(D);
            dev_memset(gdiff.get(), 0, D, st);
            const u8 *f = fl.get(); const u32 *hd = head.get(); const u64 *ks = key.get(); const u32 *cp = cnt.get(); u8 *gd = gdiff.get();
            launch_for(D, st, PFB_LAMBDA(u64 t) {
                if (f[t] || !cp[t]) return;
                u32 h = hd[t];
                if ((ks[t] & 15) != (ks[h] & 15)) gd[h] = 1;    // include/pfbwt.hpp:146
            });
        }
        const u8 *f = fl.get(); const u32 *hd = head.get(); const u32 *cp = cnt.get(); const u8 *gd = gdiff.get(); u8 *cl = cls.get();
        launch_for(D, st, PFB_LAMBDA(u64 t) {
            u32 c = cp[t];
            bool single = f[t] && (t + 1 == D || f[t + 1]);
            bool hard = c && !single && (any_sa || gd[hd[t]]);
            cl[t] = !c ? 0 : (hard ? 2 : 1);
        });
    }
    timer.lap("K7 group class");
    tm.groups = timer.stop();

    // K8 emission ---------------------------------------------------------------------
    out.n = n;
    out.bwt.alloc(n + 1);
    if (any_sa) out.sa.alloc((n + 1) * (u64)U);
    u8 *bwt = out.bwt.get();
    u8 *sa_out = out.sa.get();
    const u32 *ilist = pp.ilist.get(); const u8 *bwlast = pp.bwlast.get(); const u64 *bwsai = pp.bwsai.get();
    const u32 SMALL = 8;
    {   // easy members: their occurrences are already ordered by ilist position
        const u64 bigcap = (n + 1) / (SMALL + 1) + 1 < D ? (n + 1) / (SMALL + 1) + 1 : D;
        DevBuf<u32> biglist(bigcap), bigcount(1);
        dev_memset(bigcount.get(), 0, 4, st);
        const u32 *sp = sa.get(); const u32 *wv = wordv.get(); const u32 *cp = cnt.get(); const u8 *cl = cls.get();
        const u64 *of = off.get(); const u64 *ks = key.get(); const u8 *f = fl.get();
        u32 *bl = biglist.get(); u32 *bc = bigcount.get();
        launch_for(D, st, PFB_LAMBDA(u64 t) {
            if (cl[t] != 1) return;
            u32 c = cp[t];
            if (c > SMALL) { bl[atomic_add(bc, 1u)] = (u32)t; return; }
            u32 idx = sp[t], d = wv[t];
            bool full = idx == wl.start(d) && f[t] && (t + 1 == D || f[t + 1]);   // include/pfbwt.hpp:116 (EASY1 is never grouped)
            u64 sl = wl.suflen(d, idx);
            u8 pc = prev_byte(ks[t]);
            u64 x = of[t];
            u32 base = 1 + wl.cum(d);
            for (u32 j = 0; j < c; ++j) {
                if (full || any_sa) {
                    u32 ii = ilist[base + j];
                    bwt[x + j] = full ? bwlast[ii] : pc;          // include/pfbwt.hpp:119,151
                    if (any_sa) store_sa<U>(sa_out, x + j, bwsai[ii] - sl);
                } else bwt[x + j] = pc;
            }
        });
        timer.lap("K8 easy small");
        u64 nbig = read_scalar(bigcount.get(), st);
        // frequent words: 32 logical lanes per member, unit-stride writes
        launch_for(nbig * 32, st, PFB_LAMBDA(u64 i) {
            u64 t = bl[i >> 5];
            u32 lane = (u32)(i & 31);
            u32 c = cp[t], idx = sp[t], d = wv[t];
            bool full = idx == wl.start(d) && f[t] && (t + 1 == D || f[t + 1]);   // include/pfbwt.hpp:116 (EASY1 is never grouped)
            u64 sl = wl.suflen(d, idx);
            u8 pc = prev_byte(ks[t]);
            u64 x = of[t];
            u32 base = 1 + wl.cum(d);
            for (u32 j = lane; j < c; j += 32) {
                if (full || any_sa) {
                    u32 ii = ilist[base + j];
                    bwt[x + j] = full ? bwlast[ii] : pc;
                    if (any_sa) store_sa<U>(sa_out, x + j, bwsai[ii] - sl);
                } else bwt[x + j] = pc;
            }
        });
    }
    timer.lap("K8 easy big");
    u64 nhard = 0;
    {   // hard groups: (dense group, ilist position) keys of every occurrence, one radix sort, scatter
        DevBuf<u8> hf(D);
        {
            const u8 *cl = cls.get(); u8 *h = hf.get();
            launch_for(D, st, PFB_LAMBDA(u64 t) { h[t] = cl[t] == 2 ? 1 : 0; });
        }
        DevBuf<u32> hl;
        u64 mh;
        {
            DevBuf<u32> scratch;
            mh = compact_indices(hf.get(), D, scratch, hl, ws);
        }
        hf.release();
        timer.lap("K8 hard list");
        if (mh) {
            DevBuf<u32> hc(mh), ghead(mh), gid(mh);
            DevBuf<u64> hoff(mh);
            DevBuf<u8> gh(mh);
            {
                const u32 *hp = hl.get(); const u32 *cp = cnt.get(); const u8 *f = fl.get(); u32 *c_ = hc.get(); u8 *g = gh.get(); u32 *gd = ghead.get();
                launch_for(mh, st, PFB_LAMBDA(u64 a) {
                    u32 t = hp[a];
                    c_[a] = cp[t];
                    g[a] = f[t];                       // whole groups are listed: a group's head is its first listed member
                    gd[a] = f[t] ? (u32)a : 0;
                });
            }
            exclusive_sum<u32, u64>(hc.get(), hoff.get(), mh, ws.sc, st);
            nhard = read_scalar(hoff.get() + (mh - 1), st) + read_scalar(hc.get() + (mh - 1), st);
            u64 nhg = scan_flags(gh.get(), gid.get(), mh, ws);
            {
                DevBuf<u32> tmp(mh);
                inclusive_max(ghead.get(), tmp.get(), mh, ws.sc, st);
                ghead.swap(tmp);
            }
            const int ibits = bits_for(P);           // ilist values are in [0, P]
            const int gbits = bits_for(nhg ? nhg - 1 : 0);
            if (ibits + gbits > 64) { err = "internal: hard key overflow"; return PFB_ERR_INTERNAL; }
            DevBuf<u64> hkey(nhard), hkey_b(nhard);
            DevBuf<u32> hval(nhard), hval_b(nhard);
            {
                const u32 *hp = hl.get(); const u32 *wv = wordv.get(); const u32 *c_ = hc.get(); const u64 *ho = hoff.get();
                const u32 *gi = gid.get(); const u8 *g = gh.get();
                u64 *hk = hkey.get(); u32 *hv = hval.get();
                launch_for(mh * 32, st, PFB_LAMBDA(u64 i) {
                    u64 a = i >> 5;
                    u32 lane = (u32)(i & 31);
                    u32 t = hp[a];
                    u32 c = c_[a];
                    u64 gg = (u64)gi[a] + g[a] - 1;
                    u64 e = ho[a];
                    u32 base = 1 + wl.cum(wv[t]);
                    for (u32 j = lane; j < c; j += 32) {
                        hk[e + j] = (gg << ibits) | (u64)ilist[base + j];
                        hv[e + j] = (u32)a;
                    }
                });
            }
            timer.lap("K8 hard keys");
            sort_pairs(hkey, hkey_b, hval, hval_b, nhard, 0, ibits + gbits, ws.sc, st);
            timer.lap("K8 hard sort");
            {
                const u64 *hk = hkey.get(); const u32 *hv = hval.get(); const u64 *ho = hoff.get(); const u64 *of = off.get();
                const u32 *hp = hl.get(); const u32 *gd = ghead.get();
                const u32 *sp = sa.get(); const u32 *wv = wordv.get(); const u64 *ks = key.get();
                const u64 imask = ibits >= 64 ? ~0ull : ((1ull << ibits) - 1);
                launch_for(nhard, st, PFB_LAMBDA(u64 e) {
                    u32 a = hv[e];
                    u32 t = hp[a];
                    u32 ah = gd[a];
                    u64 x = of[hp[ah]] + (e - ho[ah]);
                    bwt[x] = prev_byte(ks[t]);                      // include/pfbwt.hpp:173-178
                    if (any_sa) {
                        u32 ii = (u32)(hk[e] & imask);
                        u64 sl = wl.suflen(wv[t], sp[t]);
                        store_sa<U>(sa_out, x, bwsai[ii] - sl);
                    }
                });
            }
        }
    }
    timer.lap("K8 hard scatter");
    out.hard = nhard;                 // "# hard cases" of include/pfbwt.hpp:188
    out.easy = n + 1 - nhard;
    if (any_sa) launch_for(1, st, PFB_LAMBDA(u64) { store_sa<U>(sa_out, 0, n); });   // src/pfbwt-f.cpp:301
    key.release(); sa.release(); fl.release(); wordv.release(); cnt.release(); off.release(); cls.release(); head.release();
    tm.emit = timer.stop();

    // K9 run boundaries + run-length SA samples (src/pfbwt-f.cpp:304-328) -------------------
    {
        DevBuf<u8> rf(n + 1);
        u8 *rp = rf.get();
        launch_for(n + 1, st, PFB_LAMBDA(u64 x) { rp[x] = bwt[x] != (x ? bwt[x - 1] : (u8)0) ? 1 : 0; });
        if (!want_rssa) {
            out.r = reduce_sum<u8, u64>(rf.get(), n + 1, ws.sc, st);
        } else {
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
    timer.lap("K9 runs");
    tm.runs = timer.stop();
    trace_log().dump("build_bwt");
    return PFB_OK;
}

inline int build_bwt(const DeviceParse &pp, const Params &prm, bool want_sa, bool want_rssa, DeviceBwt &out, Workspace &ws,
                     StageTimes &tm, std::string &err) {
    return prm.bits == 64 ? build_bwt_t<8>(pp, prm, want_sa, want_rssa, out, ws, tm, err)
                          : build_bwt_t<4>(pp, prm, want_sa, want_rssa, out, ws, tm, err);
}

}  // namespace pfb
