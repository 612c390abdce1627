// shard.inl -- methods of pfb::Engine (included inside the struct): the parse step of a single job
// spread over the ranks of a communicator.
//
// Rank g holds the records S_g of a contiguous piece of the job's record list; S = S_0 S_1 ... S_{G-1}.
//   1. trigger positions are position-local (a w-byte window), so a rank needs only the w+1 bytes
//      before its piece: one tiny all-gather of every rank's tail bytes;
//   2. a phrase belongs to the rank that holds its first byte.  The phrase that straddles the cut
//      g | g+1 ends at the first trigger of a later rank: ranks all-gather their "heads"
//      (bytes up to and including their first trigger), rank g appends what it needs to its text;
//   3. every rank parses its piece (parse_piece: same kernels as the single-GPU path) into a local
//      sorted dictionary + parse;
//   4. merge_pfp-style merge (include/pfparser.hpp:194-263, src/merge_pfp.cpp:97-170): the distinct words of
//      the local dictionaries are range-partitioned to owner ranks by one all-to-all over NVLink (contiguous
//      slices: a local dictionary is sorted), every owner sorts and deduplicates its 1/G of the union
//      (sort_strings with duplicate groups), the owners return the global ranks (reverse all-to-all) and
//      all-gather their pieces of the global dictionary; local ranks are renamed and the renamed parses /
//      last / sai arrays are all-gathered.  After this every rank holds the whole parse (dictionary
//      replicated, as the range-partitioned suffix sort of build.cuh needs it).
// Error handling is collective: a rank that fails still takes part in the next exchange and all
// ranks return an error together.

    float shard_exchange_ms = 0, shard_merge_ms = 0;

    int shard_fail_together(int rc, const char *what, pfbStream_t st) {
        u64 v = rc ? 1 : 0;
        std::vector<u64> all = comm->gather_u64(&v, 1, st);
        for (int r = 0; r < comm->world; ++r)
            if (all[(size_t)r]) {
                if (rc) return rc;
                return fail(PFB_ERR_ARG, std::string("rank ") + std::to_string(r) + " failed in " + what);
            }
        return PFB_OK;
    }

    // first trigger position inside S (index into S) given the w+1 bytes stored right before it; ~0 if none
    u64 shard_first_trigger() {
        pfbStream_t st = ws.st;
        const u64 w = (u64)prm.w, len = text.n;
        const u8 *X = text.S() - (w + 1);
        const DivTest dt = DivTest::make(prm.p);
        u64 done = 0, win = 1u << 16;
        std::vector<u32> host;
        while (done < len) {
            const u64 cntp = win < len - done ? win : len - done;
            const u64 qmax = w + cntp;                       // positions w+1 .. w+cntp of X+done are S[done .. done+cntp)
            const u64 nwords = qmax / 32 + 1;
            DevBuf<u32> trig(nwords), cnt(nwords);
            k_trigger_words(X + done, qmax, (int)w, dt, trig.get(), cnt.get(), nwords, st, w + 1);
            // the smallest set position is reduced on the device: one scalar comes back, not the bitmap
            DevBuf<u64> first(1);
            dev_memset(first.get(), 0xff, 8, st);
            {
                const u32 *tg = trig.get(); u64 *fp = first.get();
                launch_for(nwords, st, PFB_LAMBDA(u64 t) { const u32 b = tg[t]; if (b) atomic_min(fp, t * 32 + (u64)(ffs32(b) - 1)); });
            }
            const u64 q = read_scalar(first.get(), st);
            if (q != ~0ull) return done + q - (w + 1);
            done += cntp;
            win *= 8;
        }
        return ~0ull;
    }

    int shard_parse() {
        try {
            if (!comm) return fail(PFB_ERR_ARG, "no communicator attached to this context");
            pfbStream_t st = ws.st;
            const int G = comm->world, g = comm->rank;
            const u64 w = (u64)prm.w;
            Timer xt(st);
            float t_x = 0, t_m = 0;

            // ---- 0. text on the device, normalised; agree on sizes ---------------------------------
            int rc = PFB_OK;
            if (!text_ready) { text.tail_room = 4096; rc = upload_text(); }
            u64 len = rc ? 0 : text.n;
            // a rank may hold fewer than w bytes, or none at all (a byte range that covers only a header line): it then owns
            // no phrase of its own making, but its bytes still travel as tail / head of its neighbours
            if (!rc && chunk_open && g == G - 1) rc = fail(PFB_ERR_ARG, "the last rank's last record is still open (pfb_add_chunk with ends_record = 0)");
            u64 hdr[6] = {rc ? 1ull : 0ull, len, 0, 0, 0, 0};
            const u64 tl = len < w + 1 ? len : w + 1;
            if (!rc && tl) {
                u8 tail[32];
                memset(tail, 0, sizeof(tail));
                d2h(tail + (32 - tl), text.S() + (len - tl), tl, st);
                stream_sync(st);
                memcpy(&hdr[2], tail, 32);
            }
            std::vector<u64> all = comm->gather_u64(hdr, 6, st);
            for (int r = 0; r < G; ++r)
                if (all[6 * (size_t)r]) return rc ? rc : fail(PFB_ERR_ARG, "rank " + std::to_string(r) + " could not load its records");
            std::vector<u64> lens((size_t)G), starts((size_t)G + 1, 0);
            for (int r = 0; r < G; ++r) { lens[(size_t)r] = all[6 * (size_t)r + 1]; starts[(size_t)r + 1] = starts[(size_t)r] + lens[(size_t)r]; }
            const u64 n_total = starts[(size_t)G];
            shard_start = starts[(size_t)g];
            if (n_total == 0) return fail(PFB_ERR_ARG, "empty input");
            if (prm.bits == 32 && n_total > 0xFFFFFFFFull) return fail(PFB_ERR_TOO_LONG_32, "input too long, please use 64-bit version");

            // ---- 1. the w+1 bytes of T' before this rank's records -----------------------------------
            if (g > 0) {
                std::vector<u8> pre(1, DOLLAR);
                for (int r = 0; r < g; ++r) {
                    const u64 tr = lens[(size_t)r] < w + 1 ? lens[(size_t)r] : w + 1;
                    const u8 *tb = (const u8 *)&all[6 * (size_t)r + 2];
                    pre.insert(pre.end(), tb + (32 - tr), tb + 32);
                }
                h2d(text.S() - (w + 1), pre.data() + (pre.size() - (w + 1)), w + 1, st);
            }

            // ---- 2. first trigger, heads ------------------------------------------------------------------
            const u64 NONE = ~0ull;
            u64 f = g > 0 ? shard_first_trigger() : NONE;
            u64 fh[2] = {f, g == 0 ? 0ull : (f == NONE ? len : f + 1)};
            std::vector<u64> fhs = comm->gather_u64(fh, 2, st);
            std::vector<u64> hcnt((size_t)G), hoff((size_t)G + 1, 0);
            for (int r = 0; r < G; ++r) { hcnt[(size_t)r] = fhs[2 * (size_t)r + 1]; hoff[(size_t)r + 1] = hoff[(size_t)r] + hcnt[(size_t)r]; }
            DevBuf<u8> heads(hoff[(size_t)G] + 1);
            comm->all_gather_v(text.S(), heads.get(), hcnt.data(), st);
            const bool own = g == 0 || f != NONE;
            u64 rlen = 0;
            bool dollar_end = true;
            int hlast = g;
            for (int r = g + 1; r < G && own; ++r) {
                rlen += hcnt[(size_t)r];
                hlast = r;
                if (fhs[2 * (size_t)r] != NONE) { dollar_end = false; break; }
            }
            if (own) {
                if (rlen + w + 16 > w + 64 + text.tail_room) {          // rare: a very long straddling phrase
                    DevBuf<u8> nb(TextOnDevice::FRONT + len + rlen + w + 64);
                    d2d(nb.get(), text.buf.get(), TextOnDevice::FRONT + len, st);
                    stream_sync(st);
                    text.buf = std::move(nb);
                    text.tail_room = rlen;
                }
                if (rlen) d2d(text.S() + len, heads.get() + hoff[(size_t)g + 1], hoff[(size_t)hlast + 1] - hoff[(size_t)g + 1], st);
                if (dollar_end) { u8 *e = text.S() + len + rlen; launch_for(w, st, PFB_LAMBDA(u64 i) { e[i] = DOLLAR; }); }
            }
            heads.release();
            t_x += xt.stop();

            // ---- 3. local parse of the piece ---------------------------------------------------------------
            DeviceParse lp;
            u64 base = 0;                        // T' index of the first byte of the piece
            if (own) {
                PieceSpec ps;
                const u8 *Tp;
                if (g == 0) { Tp = text.Tp(); ps.qmin = w + 1; base = 0; }
                else { Tp = text.S() + f - (w - 1); ps.qmin = w; base = 1 + shard_start + f - (w - 1); }
                ps.dollar_end = dollar_end;
                ps.mem_lo = text.buf.get(); ps.mem_hi = text.buf.get() + text.buf.size();
                ps.tlen = (u64)((text.S() + len + rlen + (dollar_end ? w : 0)) - Tp);
                rc = parse_piece(Tp, ps, prm, lp, ws, tm, err);
                if (rc) fail(rc, err);
                else if (base) {
                    u64 *sp = lp.sai.get();
                    launch_for(lp.P, st, PFB_LAMBDA(u64 k) { sp[k] += base; });
                }
            }
            xt.stop();
            rc = shard_fail_together(rc, "the local parse", st);
            if (rc) return rc;
            rc = shard_merge(lp, n_total, t_x, t_m);
            shard_exchange_ms = t_x; shard_merge_ms = t_m;
            return rc;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // ---- 4. merge of the ranks' parses -----------------------------------------------------------------
    // The distinct words of all ranks are RANGE-PARTITIONED to owner ranks by an all-to-all (the hash-partitioned phrase
    // exchange of the north star, with an order-preserving key instead of a hash so that an owner's words are one interval
    // of the global dictionary): key = the first 21 symbols of a word (pack_syms, 63 bits, order preserving; equal keys go to
    // the same owner).  A local dictionary is sorted, so what a rank sends to owner d is one contiguous slice of its
    // dictionary image - no scatter.  Every owner sorts and deduplicates 1/G of the union (sort_strings with duplicate
    // groups) instead of every rank sorting all of it; global rank = words of the owners before + local rank.  The owners
    // send the ranks back (the reverse all-to-all) and all-gather their piece of the global dictionary, which every rank
    // needs whole for the range-partitioned suffix sort of build.cuh.
    int shard_merge(DeviceParse &lp, u64 n_total, float &t_x, float &t_m) {
        pfbStream_t st = ws.st;
        const int G = comm->world, g = comm->rank;
        Timer xt(st);
        const u64 Wl = lp.Dw;                                                   // local distinct words
        // -- splitters from a sample of every rank's keys
        const u64 SAMP = 256;
        DevBuf<u64> kw(Wl + 1);
        if (Wl) {
            const u8 *dp = lp.dict.get(); const u64 *wsp = lp.wstart.get(); u64 *kp = kw.get();
            launch_for(Wl, st, PFB_LAMBDA(u64 j) { kp[j] = pack_syms(dp + wsp[j], wsp[j + 1] - 1 - wsp[j], 0, 21); });
        }
        std::vector<u64> bound((size_t)G + 1, 0);
        {
            DevBuf<u64> smp(SAMP), allsmp(SAMP * (u64)G);
            const u64 *kp = kw.get(); u64 *sp = smp.get();
            launch_for(SAMP, st, PFB_LAMBDA(u64 i) { sp[i] = Wl ? kp[(2 * i + 1) * Wl / (2 * SAMP)] : ~0ull; });      // Wl < 2^32: no overflow
            comm->all_gather(smp.get(), allsmp.get(), SAMP * 8, st);
            std::vector<u64> hs(SAMP * (size_t)G);
            d2h(hs.data(), allsmp.get(), hs.size() * 8, st);
            stream_sync(st);
            std::sort(hs.begin(), hs.end());
            size_t real = hs.size();
            while (real && hs[real - 1] == ~0ull) --real;                      // ranks without words sent padding
            for (int d = 1; d < G; ++d) bound[(size_t)d] = real ? hs[real * (size_t)d / (size_t)G] : ~0ull;
            bound[(size_t)G] = ~0ull;
            for (int d = 1; d <= G; ++d) if (bound[(size_t)d] < bound[(size_t)d - 1]) bound[(size_t)d] = bound[(size_t)d - 1];
        }
        // -- slice of the local (sorted) dictionary for every owner: words [cut[d], cut[d+1])
        std::vector<u64> cut((size_t)G + 1, 0), cutb((size_t)G + 1, 0);
        {
            DevBuf<u64> db((size_t)G + 1), dc(2 * ((size_t)G + 1));
            h2d(db.get(), bound.data(), ((size_t)G + 1) * 8, st);
            const u64 *kp = kw.get(); const u64 *bp = db.get(); const u64 *wsp = lp.wstart.get(); u64 *cp = dc.get();
            const u64 G1 = (u64)G + 1;
            launch_for(G1, st, PFB_LAMBDA(u64 d) {
                u64 lo = 0, hi = Wl;                                            // first word with key >= bound[d] (all words for d = G)
                if (d + 1 == G1) lo = Wl;
                else while (lo < hi) { const u64 mid = lo + (hi - lo) / 2; if (kp[mid] < bp[d]) lo = mid + 1; else hi = mid; }
                cp[d] = lo; cp[G1 + d] = Wl ? wsp[lo] : 0;
            });
            std::vector<u64> hc(2 * ((size_t)G + 1));
            d2h(hc.data(), dc.get(), hc.size() * 8, st);
            stream_sync(st);
            for (int d = 0; d <= G; ++d) { cut[(size_t)d] = hc[(size_t)d]; cutb[(size_t)d] = hc[(size_t)G + 1 + (size_t)d]; }
            cut[0] = 0; cutb[0] = 0;
        }
        kw.release();
        // -- count matrices (words and bytes, source-major) and the phrase counts
        std::vector<u64> row(2 * (size_t)G + 1);
        for (int d = 0; d < G; ++d) { row[(size_t)d] = cut[(size_t)d + 1] - cut[(size_t)d]; row[(size_t)G + (size_t)d] = cutb[(size_t)d + 1] - cutb[(size_t)d]; }
        row[2 * (size_t)G] = lp.P;
        std::vector<u64> rows = comm->gather_u64(row.data(), 2 * G + 1, st);
        std::vector<u64> mw((size_t)G * G), mb((size_t)G * G), cP((size_t)G), oP((size_t)G + 1, 0);
        for (int s_ = 0; s_ < G; ++s_) {
            for (int d = 0; d < G; ++d) {
                mw[(size_t)s_ * G + d] = rows[(size_t)s_ * (2 * G + 1) + d];
                mb[(size_t)s_ * G + d] = rows[(size_t)s_ * (2 * G + 1) + G + d];
            }
            cP[(size_t)s_] = rows[(size_t)s_ * (2 * G + 1) + 2 * G];
            oP[(size_t)s_ + 1] = oP[(size_t)s_] + cP[(size_t)s_];
        }
        const u64 P = oP[(size_t)G];
        if (prm.bits == 32 && P > 0x7FFFFFFEull) return fail(PFB_ERR_LIMIT, "Input containing more than 2^31-2 phrases! Please use 64 bit version");
        if (P > 0xFFFFFFFEull) return fail(PFB_ERR_LIMIT, "Input containing more than 2^32-2 phrases! This is currently a hard limit");
        u64 M = 0, MB = 0;                                                      // words / bytes this owner receives
        for (int s_ = 0; s_ < G; ++s_) { M += mw[(size_t)s_ * G + g]; MB += mb[(size_t)s_ * G + g]; }
        {
            u64 tw = 0;
            for (u64 x : mw) tw += x;
            if (tw > 0xFFFFFFF0ull) return fail(PFB_ERR_LIMIT, "more than 2^32 dictionary words across the ranks");
        }
        auto scaled = [&](const std::vector<u64> &c, u64 k) { std::vector<u64> v(c); for (auto &x : v) x *= k; return v; };
        // -- the exchange: bytes, word lengths, occurrence counts
        DevBuf<u8> ub(MB + 1);
        DevBuf<u32> ulen(M + 1), uocc(M + 1), slen(Wl + 1);
        if (Wl) {
            const u64 *wsp = lp.wstart.get(); u32 *lp_ = slen.get();
            launch_for(Wl, st, PFB_LAMBDA(u64 j) { lp_[j] = (u32)(wsp[j + 1] - wsp[j]); });      // word + EndOfWord
        }
        comm->all_to_all_v(lp.dict.get(), ub.get(), mb.data(), st);
        comm->all_to_all_v(slen.get(), ulen.get(), scaled(mw, 4).data(), st);
        comm->all_to_all_v(lp.occ.get(), uocc.get(), scaled(mw, 4).data(), st);
        slen.release();
        t_x += xt.stop();
        // -- this owner's share of the union: sort, groups of identical words collapse into one global word
        DevBuf<u64> ustart(M + 1);
        dev_memset(ulen.get() + M, 0, 4, st);
        exclusive_sum<u32, u64>(ulen.get(), ustart.get(), M + 1, ws.sc, st);
        DevBuf<u32> order(M + 1);
        { u32 *op = order.get(); launch_for(M, st, PFB_LAMBDA(u64 j) { op[j] = (u32)j; }); }
        DevBuf<u8> head;
        u64 Wg = 0;                                                             // distinct words of this owner
        DevBuf<u32> gidx(M + 1), map(M + 1);
        if (M) {
            WordStr wsr;
            wsr.wstart = ustart.get();
            int rc = sort_strings(ub.get(), wsr, order, M, ws, &tm.phrase_sort_rounds, err, &head);
            if (rc) return fail(rc, err);
            Wg = scan_flags(head.get(), gidx.get(), M, ws);                     // gidx[j] + head[j] = 1-based rank of order[j] inside this owner
        }
        DevBuf<u32> headpos(Wg + 1), occg(Wg + 1), leng(Wg + 1);
        dev_memset(occg.get(), 0, (Wg + 1) * 4, st);
        if (M) {
            const u32 *op = order.get(); const u32 *gi = gidx.get(); const u8 *hd = head.get(); const u32 *uo = uocc.get(); const u32 *ul = ulen.get();
            u32 *mp = map.get(); u32 *hp = headpos.get(); u32 *oc = occg.get(); u32 *lg = leng.get();
            launch_for(M, st, PFB_LAMBDA(u64 j) {
                const u32 r = gi[j] + hd[j];               // 1-based
                mp[op[j]] = r;
                if (hd[j]) { hp[r - 1] = op[j]; lg[r - 1] = ul[op[j]]; }
                atomic_add(&oc[r - 1], uo[op[j]]);
            });
        }
        // this owner's piece of the global dictionary image (include/pfbwt_io.hpp:71-82)
        DevBuf<u64> pstart(Wg + 1);
        dev_memset(leng.get() + Wg, 0, 4, st);
        exclusive_sum<u32, u64>(leng.get(), pstart.get(), Wg + 1, ws.sc, st);
        const u64 PB = read_scalar(pstart.get() + Wg, st);                      // bytes of the piece
        DevBuf<u8> piece(PB + 1);
        if (Wg) {
            const u32 *hp = headpos.get(); const u64 *us = ustart.get(); const u64 *ps_ = pstart.get(); const u8 *src = ub.get(); u8 *dst = piece.get();
            launch_for(Wg * 32, st, PFB_LAMBDA(u64 i) {
                const u64 r = i / 32, lane = i % 32;
                const u64 s0 = us[hp[r]], len = us[hp[r] + 1] - s0;             // includes the EndOfWord byte
                const u64 d0 = ps_[r];
                for (u64 q = lane; q < len; q += 32) dst[d0 + q] = src[s0 + q];
            });
        }
        t_m += xt.stop();
        // -- sizes of the pieces; ranks back to the senders; the pieces to everybody
        u64 pz[2] = {Wg, PB};
        std::vector<u64> pzs = comm->gather_u64(pz, 2, st);
        std::vector<u64> cW((size_t)G), cB((size_t)G), oW((size_t)G + 1, 0), oB((size_t)G + 1, 0);
        for (int r = 0; r < G; ++r) {
            cW[(size_t)r] = pzs[2 * (size_t)r]; cB[(size_t)r] = pzs[2 * (size_t)r + 1];
            oW[(size_t)r + 1] = oW[(size_t)r] + cW[(size_t)r]; oB[(size_t)r + 1] = oB[(size_t)r] + cB[(size_t)r];
        }
        const u64 Dw = oW[(size_t)G], dsize = oB[(size_t)G] + 1;
        if (dsize >= DICT_MAX_BYTES) return fail(PFB_ERR_LIMIT, "dictionary larger than 2^40 bytes is not supported");
        {
            u32 *mp = map.get(); const u32 add = (u32)oW[(size_t)g];
            launch_for(M, st, PFB_LAMBDA(u64 j) { mp[j] += add; });              // global 1-based rank of every received word
        }
        DevBuf<u32> gmap(Wl + 1);                                               // global rank of local word j
        {
            std::vector<u64> back((size_t)G * G);                               // transposed: owner d sends rank s what s sent d
            for (int s_ = 0; s_ < G; ++s_) for (int d = 0; d < G; ++d) back[(size_t)d * G + s_] = mw[(size_t)s_ * G + d] * 4;
            comm->all_to_all_v(map.get(), gmap.get(), back.data(), st);
        }
        parse = DeviceParse();
        parse.n = n_total; parse.P = P; parse.Dw = Dw; parse.dsize = dsize;
        parse.dict.alloc(dsize);
        parse.occ.alloc(Dw ? Dw : 1);
        DevBuf<u32> alllen(Dw + 1);
        comm->all_gather_v(piece.get(), parse.dict.get(), cB.data(), st);
        comm->all_gather_v(leng.get(), alllen.get(), scaled(cW, 4).data(), st);
        comm->all_gather_v(occg.get(), parse.occ.get(), scaled(cW, 4).data(), st);
        t_x += xt.stop();
        dev_memset(alllen.get() + Dw, 0, 4, st);
        parse.wstart.alloc(Dw + 1);
        exclusive_sum<u32, u64>(alllen.get(), parse.wstart.get(), Dw + 1, ws.sc, st);       // wstart[Dw] = dsize - 1
        { u8 *dict = parse.dict.get(); launch_for(1, st, PFB_LAMBDA(u64) { dict[dsize - 1] = ENDOFDICT; }); }
        piece.release(); ub.release(); ustart.release(); order.release(); head.release(); gidx.release(); map.release();
        // local ranks -> global ranks, then the whole parse on every rank
        DevBuf<u32> lp2(lp.P + 1);
        {
            const u32 *mp = gmap.get(); const u32 *pr = lp.parse.get(); u32 *o = lp2.get();
            launch_for(lp.P, st, PFB_LAMBDA(u64 k) { o[k] = mp[pr[k] - 1]; });
        }
        t_m += xt.stop();
        parse.parse.alloc(P + 1);
        parse.last.alloc(P + 1);
        parse.sai.alloc(P + 1);
        comm->all_gather_v(lp2.get(), parse.parse.get(), scaled(cP, 4).data(), st);
        comm->all_gather_v(lp.last.get(), parse.last.get(), cP.data(), st);
        comm->all_gather_v(lp.sai.get(), parse.sai.get(), scaled(cP, 8).data(), st);
        dev_memset(parse.parse.get() + P, 0, 4, st);
        stream_sync(st);
        t_x += xt.stop();
        parsed = true; built = false;
        return PFB_OK;
    }
