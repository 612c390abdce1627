// markers.cuh -- the marker-array join on the device: what src/mps_to_ma.cpp does with
// write_marker_array (include/marker_array.hpp:137-174) over an rle_window_arr (include/rle_window_array.hpp:15-51,118-130).
//
// Input  <mps>: a stream of u64 records  [first, last, marker, marker, ..., 2^64-1]  - a run of text positions and the
//               variant markers within reach of it (MarkerPositionsWriter, include/marker_array.hpp:62-134), ascending.
//        SA:    the suffix array in BWT order, u64 entries (the reference reads `pfbwt-f64 --stdout sa`).
// Output <ma>:  for every maximal run of BWT positions i whose text positions SA[i] carry the SAME non-empty marker list
//               (compared by content, marker_array.hpp:148 vec_eq):  [first i, last i, markers..., 2^64-1].
//
// The reference walks the SA serially and answers every at(SA[i]) with rank/select on three sd_vectors.  Here:
//   load_marker_runs   the record stream is cut at its delimiters with one flag pass + one scan (record r = words between
//                      delimiter r-1 and r): start[r], end[r], value slice voff[r]..voff[r+1]; the checks the reference makes
//                      (a run that starts or ends at 0 aborts it, rle_window_array.hpp:28-31) and the ones it relies on
//                      silently (ascending, disjoint, at least one marker) are made on the device
//   marker_join_chunk  one thread per BWT position: binary search of SA[i] in start[] (the tables are small and stay in
//                      L2), run id or none; a position is the LAST of an output run when its list differs from its right
//                      neighbour's - the run's first position is the most recent list change to its left (one max-scan) -
//                      the output words of every run are placed by one exclusive scan of the run sizes and written by the
//                      thread of the run's last position.  The SA is taken in chunks (from HBM after pfb_build_bwt, or from a
//                      file / stdin): a run that crosses a chunk border is written by the chunk holding its last position,
//                      its first position travels in MarkerCarry.
// (included by engine.hpp)
#pragma once

namespace pfb {

static const u64 MARKER_DELIM = ~0ull;
static const u32 MARKER_NONE = 0xFFFFFFFFu;

struct MarkerRuns {
    DevBuf<u64> start, end, voff, vals;      // R, R, R+1, total values
    u64 R = 0;
    DevBuf<u8> raw;                          // the file image (vals points into a copy, raw is dropped after loading)
};

// 0 ok; rc != 0 with err set (PFB_ERR_IO: unreadable / malformed, PFB_ERR_ARG: the reference's own refusal)
inline int load_marker_runs(const std::string &path, MarkerRuns &mr, int device, Workspace &ws, std::string &err) {
    pfbStream_t st = ws.st;
    u64 bytes = 0;
    std::string e;
    if (read_whole_file_to_dev(path, mr.raw, bytes, 64, device, st, e)) { err = e; return PFB_ERR_IO; }
    if (bytes % 8) { err = "bad mps file (size is not a multiple of 8)"; return PFB_ERR_IO; }
    const u64 nw = bytes / 8;
    mr.R = 0;
    if (nw == 0) { mr.start.alloc(1); mr.end.alloc(1); mr.voff.alloc(1); mr.vals.alloc(1); dev_memset(mr.voff.get(), 0, 8, st); return PFB_OK; }
    const u64 *in = reinterpret_cast<const u64 *>(mr.raw.get());
    DevBuf<u8> isd(nw);
    DevBuf<u32> rid(nw);
    {
        u8 *f = isd.get();
        launch_for(nw, st, PFB_LAMBDA(u64 i) { f[i] = in[i] == MARKER_DELIM ? 1 : 0; });
    }
    const u64 R = scan_flags(isd.get(), rid.get(), nw, ws);        // rid[i] = delimiters before word i = record of word i
    if (R == 0) { err = "bad mps file (no complete record)"; return PFB_ERR_IO; }
    if (R >= 0xFFFFFFF0ull) { err = "more than 2^32 marker runs"; return PFB_ERR_LIMIT; }
    mr.R = R;
    mr.start.alloc(R); mr.end.alloc(R); mr.voff.alloc(R + 1);
    DevBuf<u64> dpos(R), nval(R + 1);
    DevBuf<u32> bad(4);
    dev_memset(bad.get(), 0, 16, st);
    {
        const u8 *f = isd.get(); const u32 *ri = rid.get(); u64 *dp = dpos.get();
        launch_for(nw, st, PFB_LAMBDA(u64 i) { if (f[i]) dp[ri[i]] = i; });
        u64 *sp = mr.start.get(); u64 *ep = mr.end.get(); u64 *nv = nval.get(); u32 *b = bad.get();
        launch_for(R + 1, st, PFB_LAMBDA(u64 r) {
            if (r == R) { nv[r] = 0; return; }
            const u64 b0 = r ? dp[r - 1] + 1 : 0, len = dp[r] - b0;       // words of the record without its delimiter
            if (len < 3) { b[0] = 1; sp[r] = ep[r] = 0; nv[r] = 0; return; }     // two keys and at least one marker
            sp[r] = in[b0]; ep[r] = in[b0 + 1]; nv[r] = len - 2;
            if (in[b0] == 0 || in[b0 + 1] == 0) b[1] = 1;                  // rle_window_array.hpp:28-31 (rs = re = 0 there)
            if (in[b0] > in[b0 + 1]) b[2] = 1;
        });
    }
    exclusive_sum<u64, u64>(nval.get(), mr.voff.get(), R + 1, ws.sc, st);
    const u64 total = read_scalar(mr.voff.get() + R, st);
    mr.vals.alloc(total ? total : 1);
    {
        const u64 *sp = mr.start.get(); const u64 *ep = mr.end.get(); const u64 *dp = dpos.get(); const u64 *vo = mr.voff.get();
        u64 *vp = mr.vals.get(); u32 *b = bad.get();
        launch_for(R, st, PFB_LAMBDA(u64 r) {
            if (r && sp[r] <= ep[r - 1]) b[3] = 1;                         // ascending and disjoint: what rank/select assumes
            const u64 b0 = (r ? dp[r - 1] + 1 : 0) + 2, k = vo[r + 1] - vo[r];
            for (u64 j = 0; j < k; ++j) vp[vo[r] + j] = in[b0 + j];
        });
    }
    u32 hb[4];
    d2h(hb, bad.get(), 16, st);
    stream_sync(st);
    mr.raw.release();
    if (hb[0]) { err = "bad mps file (a record needs two keys and at least one marker)"; return PFB_ERR_IO; }
    if (hb[1]) { err = "warning equal run start at 0 or run end at 0"; return PFB_ERR_ARG; }
    if (hb[2] || hb[3]) { err = "bad mps file (runs must be ascending and disjoint)"; return PFB_ERR_IO; }
    return PFB_OK;
}

struct MarkerTables { const u64 *start, *end, *voff, *vals; u32 R; };
// run holding text position s, or MARKER_NONE (rle_window_arr::at: #starts <= s == #ends < s + 1)
PFB_HD inline u32 marker_run_of(const MarkerTables &t, u64 s) {
    if (!t.R || s < t.start[0]) return MARKER_NONE;
    u32 lo = 0, hi = t.R;
    while (hi - lo > 1) { const u32 mid = lo + ((hi - lo) >> 1); if (t.start[mid] <= s) lo = mid; else hi = mid; }
    return s <= t.end[lo] ? lo : MARKER_NONE;
}
// same marker list (by content: two runs may carry equal lists, marker_array.hpp:148)
PFB_HD inline bool marker_same(const MarkerTables &t, u32 a, u32 b) {
    if (a == b) return true;
    if (a == MARKER_NONE || b == MARKER_NONE) return false;
    const u64 ka = t.voff[a + 1] - t.voff[a], kb = t.voff[b + 1] - t.voff[b];
    if (ka != kb) return false;
    for (u64 j = 0; j < ka; ++j) if (t.vals[t.voff[a] + j] != t.vals[t.voff[b] + j]) return false;
    return true;
}

struct MarkerCarry {
    u64 done = 0;                 // BWT positions consumed so far
    u32 prev = MARKER_NONE;       // run of the last position consumed
    u64 first = 0;                // first BWT position of the output run that position belongs to (if prev != NONE)
    u64 words = 0;                // output words written so far
};

// One chunk of the SA: m entries of U bytes at sa (device), BWT positions carry.done .. carry.done + m.  has_next / next:
// the SA entry after the chunk (decides whether the chunk's last position ends a run).  The output words of the runs that
// END in this chunk are appended to out (device, reallocated as needed); *nout = how many.
template <class SAT>
inline int marker_join_chunk(const MarkerRuns &mr, const SAT *sa, u64 m, bool has_next, u64 next, MarkerCarry &carry, DevBuf<u64> &out,
                             u64 *nout, Workspace &ws) {
    pfbStream_t st = ws.st;
    *nout = 0;
    if (m == 0) return PFB_OK;
    MarkerTables t;
    t.start = mr.start.get(); t.end = mr.end.get(); t.voff = mr.voff.get(); t.vals = mr.vals.get(); t.R = (u32)mr.R;
    DevBuf<u32> rec(m + 1);
    {
        u32 *rp = rec.get();
        launch_for(m + 1, st, PFB_LAMBDA(u64 i) { rp[i] = i < m ? marker_run_of(t, (u64)sa[i]) : (has_next ? marker_run_of(t, next) : MARKER_NONE); });
    }
    // change[i] = 1 + global position when the list of position i differs from its left neighbour's (a run starts here)
    DevBuf<u64> chg(m), firstp(m), sz(m + 1), off(m + 1);
    const u32 prev = carry.prev; const u64 base = carry.done;
    {
        const u32 *rp = rec.get(); u64 *cp = chg.get(); u64 *zp = sz.get();
        launch_for(m + 1, st, PFB_LAMBDA(u64 i) {
            if (i == m) { zp[i] = 0; return; }
            const u32 left = i ? rp[i - 1] : prev;
            cp[i] = marker_same(t, left, rp[i]) ? 0 : base + i + 1;
            const bool last = !marker_same(t, rp[i], rp[i + 1]) || (i + 1 == m && !has_next);
            zp[i] = (last && rp[i] != MARKER_NONE) ? 3 + (t.voff[rp[i] + 1] - t.voff[rp[i]]) : 0;
        });
    }
    inclusive_max(chg.get(), firstp.get(), m, ws.sc, st);
    exclusive_sum<u64, u64>(sz.get(), off.get(), m + 1, ws.sc, st);
    const u64 total = read_scalar(off.get() + m, st);
    if (total) {
        if (out.size() < total) out.alloc(total + total / 8);
        const u32 *rp = rec.get(); const u64 *fp = firstp.get(); const u64 *zp = sz.get(); const u64 *op = off.get(); u64 *o = out.get();
        const u64 cfirst = carry.first;
        launch_for(m, st, PFB_LAMBDA(u64 i) {
            if (!zp[i]) return;
            const u32 r = rp[i];
            u64 *w = o + op[i];
            w[0] = fp[i] ? fp[i] - 1 : cfirst;            // no change inside the chunk up to i: the run began in an earlier chunk
            w[1] = base + i;
            const u64 k = t.voff[r + 1] - t.voff[r];
            for (u64 j = 0; j < k; ++j) w[2 + j] = t.vals[t.voff[r] + j];
            w[2 + k] = MARKER_DELIM;
        });
    }
    // what the next chunk needs: the run of the last position and where its output run began
    const u32 lastrec = read_scalar(rec.get() + (m - 1), st);
    const u64 lastfirst = read_scalar(firstp.get() + (m - 1), st);
    if (lastrec != MARKER_NONE && lastfirst) carry.first = lastfirst - 1;
    carry.prev = lastrec;
    carry.done += m;
    carry.words += total;
    *nout = total;
    return PFB_OK;
}

}  // namespace pfb
