// engine.hpp -- host-side object behind the C ABI (include/pfbwt_b200.h).  One
// Engine = one pfbwtf::PfParser (+ the PrefixFreeBWT built from it) living on one
// GPU.  Mirrors the reference's operator surface for the hot path:
//   add_fasta / finalize        include/pfparser.hpp:299-369, 484-492
//   bwt_of_parse                include/pfparser.hpp:379-467
//   save_parser / load_parser   include/pfbwt_io.hpp:211-249
//   operator+=                  include/pfparser.hpp:194-263
//   generate_bwt_lcp + writer   include/pfbwt.hpp:96-194, src/pfbwt-f.cpp:272-349
#pragma once
#include "host_io.hpp"
#include "pipeline.cuh"
#include "ingest.cuh"
#include "devio.cuh"
#include "markers.cuh"
#include <memory>
#include <thread>


namespace pfb {

struct Segment {             // a piece of S waiting to be parsed
    bool on_device = false;
    DevBuf<u8> dev;          // normalised bytes (reconstructed from a loaded parse)
    size_t host_off = 0;     // offset into Engine::raw
    u64 len = 0;
};

struct Engine {
    Params prm;
    int device = 0;
    Workspace ws;
    StageTimes tm;
    std::string err;
    int last_rc = 0;
    float run_ms = 0;

    HostBytes raw;                   // record bytes + A^w padding, as read from FASTA / caller buffers
    std::vector<Segment> segs;
    FastaResult docs;
    bool store_docs = false;
    u64 pending_n = 0;               // bytes of S accumulated in segs
    bool input_released = false;     // pfb_release_text dropped the staged input (cleared by pfb_clear)
    float ingest_ms = 0;             // device time of the FASTA record kernels so far
    u64 ingest_files = 0;            // files that took the device reader

    TextOnDevice text;
    bool text_ready = false;
    DeviceParse parse;
    bool parsed = false;
    DeviceBwt bwt;
    bool built = false;
    bool built_sa = false, built_rssa = false;

    // multi-GPU job: this context is rank comm->rank of comm->world; its records
    // are one contiguous piece of the job's record list (rank order = text order)
    std::unique_ptr<Comm> comm;
    u64 shard_start = 0;             // offset of this rank's records in S

    int fail(int rc, const std::string &m) { err = m; last_rc = rc; return rc; }
    ~Engine() { wait_saves(); }

    // ---- ingestion -----------------------------------------------------------
    void push_host_segment(size_t off, u64 len) {
        if (!len) return;
        if (!segs.empty() && !segs.back().on_device && segs.back().host_off + segs.back().len == off) segs.back().len += len;
        else { Segment s; s.host_off = off; s.len = len; segs.push_back(std::move(s)); }
        pending_n += len;
    }

    int add_fasta(const std::string &path) {
        if (chunk_open) return fail(PFB_ERR_ARG, "a record is still open (pfb_add_chunk with ends_record = 0): close it first");
        // FASTA records are split on the device (ingest.cuh; gzip and stdin are inflated / drained on the host first);
        // FASTQ, CR LF line ends and inputs too large for HBM go through the host reader below
        try {
            IngestResult ir;
            std::string e;
            const int rc = ingest_fasta_device(path, prm.w, store_docs, device, ws, ir, e);
            if (rc < 0) return fail(PFB_ERR_IO, e);
            if (rc == 0) {
                if (prm.bits == 32 && ir.total_len > 0xFFFFFFFFull)   // include/pfparser.hpp:326-332
                    return fail(PFB_ERR_TOO_LONG_32, "input too long, please use 64-bit version");
                for (u64 s : ir.starts) docs.starts.push_back(pending_n + s);
                for (auto &nm : ir.names) docs.names.push_back(nm);
                docs.nrec += ir.nrec; docs.total_len += ir.total_len;
                Segment s;
                s.on_device = true; s.len = ir.len;
                s.dev = std::move(ir.seg);
                pending_n += s.len;
                segs.push_back(std::move(s));
                ingest_ms += ir.device_ms; ingest_files++;
                invalidate();
                return PFB_OK;
            }
            if (rc == 2) {          // the stream has been taken in but needs the host reader (FASTQ, CR LF, no header, ...)
                size_t off0 = raw.n;
                FastaResult r;
                read_fasta_mem(ir.hostraw.p, ir.hostraw.n, prm.w, raw, r, store_docs, 0);
                return take_host_records(off0, r);
            }
        } catch (const Error &ex) { return fail(ex.code, ex.what()); }
        size_t off0 = raw.n;
        FastaResult r;
        std::string e;
        if (read_fasta(path, prm.w, raw, r, store_docs, 0, e)) return fail(PFB_ERR_IO, e);
        return take_host_records(off0, r);
    }
    int take_host_records(size_t off0, FastaResult &r) {
        if (prm.bits == 32 && r.total_len > 0xFFFFFFFFull)   // include/pfparser.hpp:326-332
            return fail(PFB_ERR_TOO_LONG_32, "input too long, please use 64-bit version");
        // doc starts are offsets in S of everything parsed so far
        u64 base = pending_n - (u64)0;
        for (size_t i = 0; i < r.starts.size(); ++i) docs.starts.push_back(base + (r.starts[i] - off0));
        for (auto &nm : r.names) docs.names.push_back(nm);
        docs.nrec += r.nrec; docs.total_len += r.total_len;
        push_host_segment(off0, raw.n - off0);
        invalidate();
        return PFB_OK;
    }

    // Records handed over in caller memory go straight to HBM (one H2D per record, padding
    // written by a kernel); many small records are staged through `raw` instead.
    int add_records(const u8 *seq, const u64 *rec_len, u64 nrec) {
        if (chunk_open) return fail(PFB_ERR_ARG, "a record is still open (pfb_add_chunk with ends_record = 0): close it first");
        try {
            u64 tot = 0;
            for (u64 r = 0; r < nrec; ++r) tot += rec_len[r];
            if (prm.bits == 32 && tot > 0xFFFFFFFFull) return fail(PFB_ERR_TOO_LONG_32, "input too long, please use 64-bit version");
            const u64 w = (u64)prm.w;
            if (nrec > 256) {
                size_t off0 = raw.n;
                raw.reserve(raw.n + tot + nrec * w);
                u64 src = 0;
                for (u64 r = 0; r < nrec; ++r) {
                    raw.append(seq + src, rec_len[r]);
                    raw.append_fill('A', (size_t)w);
                    src += rec_len[r];
                }
                push_host_segment(off0, raw.n - off0);
            } else if (nrec) {
                Segment s;
                s.on_device = true; s.len = tot + nrec * w;
                s.dev.alloc(s.len);
                u64 src = 0, dst = 0;
                for (u64 r = 0; r < nrec; ++r) {
                    if (rec_len[r]) h2d(s.dev.get() + dst, seq + src, rec_len[r], ws.st);
                    u8 *pad = s.dev.get() + dst + rec_len[r];
                    launch_for(w, ws.st, PFB_LAMBDA(u64 i) { pad[i] = 'A'; });
                    src += rec_len[r]; dst += rec_len[r] + w;
                }
                stream_sync(ws.st);   // caller memory may be reused after return
                pending_n += s.len;
                segs.push_back(std::move(s));
            }
            docs.nrec += nrec; docs.total_len += tot;
            invalidate();
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // A piece of ONE record: `len` bytes that continue the record the previous chunk left open (or start a new one);
    // ends_record closes it (the w filler 'A's of include/pfparser.hpp:333-337 follow).  This is the mid-record cut of a
    // multi-GPU job: rank g holds bytes [a_g, a_{g+1}) of the text whatever the record structure, and only the rank that
    // holds a record's last byte pads it.  On one context, chunks of a record simply concatenate.
    bool chunk_open = false;         // the last chunk added left its record open
    int add_chunk(const u8 *seq, u64 len, bool ends_record) {
        try {
            if (prm.bits == 32 && docs.total_len + len > 0xFFFFFFFFull) return fail(PFB_ERR_TOO_LONG_32, "input too long, please use 64-bit version");
            const u64 w = (u64)prm.w;
            const u64 tot = len + (ends_record ? w : 0);
            if (tot) {
                Segment s;
                s.on_device = true; s.len = tot;
                s.dev.alloc(tot);
                if (len) h2d(s.dev.get(), seq, len, ws.st);
                if (ends_record) { u8 *pad = s.dev.get() + len; launch_for(w, ws.st, PFB_LAMBDA(u64 i) { pad[i] = 'A'; }); }
                stream_sync(ws.st);   // caller memory may be reused after return
                pending_n += tot;
                segs.push_back(std::move(s));
            }
            if (ends_record) docs.nrec++;
            docs.total_len += len;
            chunk_open = !ends_record;
            invalidate();
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // one rank's byte range of a plain FASTA file that is split over the ranks of a multi-GPU job (ingest.cuh, range mode)
    int add_fasta_range(const std::string &path, u64 begin, u64 end) {
        try {
            IngestResult ir;
            IngestRange rg;
            rg.begin = begin; rg.end = end;
            std::string e;
            if (ingest_fasta_device(path, prm.w, store_docs, device, ws, ir, e, &rg) != 0) return fail(PFB_ERR_IO, e);
            if (prm.bits == 32 && docs.total_len + ir.total_len > 0xFFFFFFFFull) return fail(PFB_ERR_TOO_LONG_32, "input too long, please use 64-bit version");
            for (u64 s : ir.starts) docs.starts.push_back(pending_n + s);
            for (auto &nm : ir.names) docs.names.push_back(nm);
            docs.nrec += ir.nrec; docs.total_len += ir.total_len;
            if (ir.len) {
                Segment s;
                s.on_device = true; s.len = ir.len;
                s.dev = std::move(ir.seg);
                pending_n += s.len;
                segs.push_back(std::move(s));
            }
            if (rg.end > rg.begin) chunk_open = rg.open_end;
            ingest_ms += ir.device_ms; ingest_files++;
            invalidate();
            return PFB_OK;
        } catch (const Error &ex) { return fail(ex.code, ex.what()); }
    }

    void clear() {
        wait_saves();
        segs.clear(); raw.n = 0; pending_n = 0; docs = FastaResult(); input_released = false; chunk_open = false;
        text.buf.release(); text.n = 0;
        parse = DeviceParse(); bwt = DeviceBwt();
        invalidate();
    }

    void invalidate() { text_ready = false; parsed = false; built = false; streamed = false; }

    // ---- text upload: T' = Dollar . S . Dollar^w, S 16-byte aligned ---------------
    int upload_text() {
        try {
            const u64 n = pending_n;
            if (n == 0 && !comm) return fail(PFB_ERR_ARG, "empty input");     // a rank of a multi-GPU job may hold nothing
            if (input_released) return fail(PFB_ERR_ARG, "the input was released (pfb_release_text): pfb_clear and add it again");
            Timer timer(ws.st);
            text.buf.alloc(TextOnDevice::FRONT + n + (u64)prm.w + 64 + text.tail_room);
            text.n = n;
            u8 *S = text.S();
            u64 o = 0;
            for (auto &s : segs) {
                if (s.on_device) d2d(S + o, s.dev.get(), s.len, ws.st);
                else h2d(S + o, raw.p + s.host_off, s.len, ws.st);
                o += s.len;
            }
            u8 *Tp = text.Tp();
            const int w = prm.w;
            launch_for((u64)w + 1, ws.st, PFB_LAMBDA(u64 i) {
                if (i == 0) Tp[0] = DOLLAR; else Tp[n + i] = DOLLAR;
            });
            // K0 normalise
            DevBuf<u64> errpos(1);
            dev_memset(errpos.get(), 0xff, 8, ws.st);
            k_normalise(S, n, prm.non_acgt_to_a, errpos.get(), ws.st);
            u64 ep = read_scalar(errpos.get(), ws.st);
            tm.normalise = timer.stop();
            if (ep != ~0ull) {
                u8 c = read_scalar(S + ep, ws.st);
                char msg[128];
                snprintf(msg, sizeof(msg), "error, invalid character %d/%c -> %d", (int)(signed char)c, c, 5);   // include/hash.hpp:31
                return fail(PFB_ERR_INVALID_CHAR, msg);
            }
            text_ready = true;
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // ---- finalize = parse on device ---------------------------------------------------
    int parse_device() {
        try {
            if (chunk_open) return fail(PFB_ERR_ARG, "the last record is still open (pfb_add_chunk with ends_record = 0)");
            if (!text_ready) { int rc = upload_text(); if (rc) return rc; }
            parse = DeviceParse();
            int rc = parse_text(text, prm, parse, ws, tm, err);
            if (rc) return fail(rc, err);
            parsed = true; built = false;
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }
    int finalize() { wait_saves(); return comm ? shard_parse() : parse_device(); }      // collective when a communicator is attached

    int do_parse_bwt() {
        try {
            if (!parsed) return fail(PFB_ERR_ARG, "parse not finalized");
            int rc = parse_bwt(parse, prm, ws, tm, err);
            if (rc) return fail(rc, err);
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    int build(bool sa, bool rssa) {
        try {
            if (!parsed) return fail(PFB_ERR_ARG, "no parse loaded");
            if (!parse.has_parse_bwt) { int rc = do_parse_bwt(); if (rc) return rc; }
            bwt = DeviceBwt();
            u64 mem_buckets = 0;
            if (const char *e = getenv("PFB_MEM_BUCKETS")) mem_buckets = (u64)atol(e);          // test knob
            if (!comm && (parse.dsize > 0xFFFFFFF0ull || mem_buckets > 1)) return build_in_key_ranges(sa, rssa, mem_buckets);
            int rc = build_bwt(parse, prm, sa, rssa, bwt, ws, tm, err, comm.get());
            if (rc) return fail(rc, err);
            built = true; built_sa = sa; built_rssa = rssa;
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // ---- dictionaries of 2^32 bytes or more (pfbwt-f64: gsa/gsacak.h:44-58 indexes the dictionary with 64-bit uint_t).
    // Indices INTO the sorted suffix array stay 32-bit here, so such a dictionary is sorted and emitted one key range
    // after the other (each fewer than 2^32 suffixes, the loop of build_to_files) and the slices are put together in HBM:
    // the result is the same DeviceBwt a one-piece build leaves.
    int build_in_key_ranges(bool sa, bool rssa, u64 B) {
        const bool any_sa = sa || rssa;
        const u64 U = (u64)prm.bits / 8, n = parse.n;
        if (B < 2) {
            B = parse.dsize / (1ull << 30) + 1;                  // <= 2^30 suffixes per range on average: splitters balance work, not counts
#ifndef PFB_EMU
            size_t fr = 0, tot = 0;
            cudaMemGetInfo(&fr, &tot);
            const u64 avail = (u64)fr + slab_pool().idle();
            const u64 keep = (n + 1) * (1 + (any_sa ? U : 0)), est = build_workspace_estimate(any_sa, rssa);
            if (avail > keep + (est >> 6)) { const u64 byMem = est / (avail - keep) + 1; if (byMem > B) B = byMem; }
#endif
        }
        if (B > (u64)KEY_MAX_DEST) B = KEY_MAX_DEST;
        bwt.n = n; bwt.off = 0; bwt.len = n + 1;
        bwt.bwt.alloc(n + 1 + 16);
        if (any_sa) bwt.sa.alloc((n + 1) * U + 16);
        struct Piece { DevBuf<u8> ssa, esa; u64 so, sc, eo, ec; };
        std::vector<Piece> pieces;
        BucketCarry carry;
        for (u64 b = 0; b < B; ++b) {
            ShardSpec spec;
            spec.rank = (int)b; spec.world = (int)B;
            DeviceBwt out;
            int rc = build_bwt(parse, prm, sa, rssa, out, ws, tm, err, nullptr, &spec, &carry);
            if (rc) return fail(rc, err);
            if (out.len) {
                d2d(bwt.bwt.get() + out.off, out.bwt.get(), out.len, ws.st);
                if (any_sa) d2d(bwt.sa.get() + out.off * U, out.sa.get(), out.len * U, ws.st);
            }
            if (rssa) {
                Piece pc;
                pc.so = out.ssa_off; pc.sc = out.ssa_cnt; pc.eo = out.esa_off; pc.ec = out.esa_cnt;
                pc.ssa = std::move(out.ssa); pc.esa = std::move(out.esa);
                pieces.push_back(std::move(pc));
            }
            bwt.r = out.r; bwt.easy = out.easy; bwt.hard = out.hard;
            stream_sync(ws.st);
        }
        if (carry.off != n + 1) return fail(PFB_ERR_INTERNAL, "internal: key ranges emitted " + std::to_string(carry.off) + " positions, n+1 = " + std::to_string(n + 1));
        if (rssa) {
            // the last range cannot know that its last position ends the last run only if another range followed: it did not
            if (carry.ssa != bwt.r || carry.esa != bwt.r) return fail(PFB_ERR_INTERNAL, "internal: run samples of the key ranges do not add up");
            bwt.ssa.alloc(2 * bwt.r * U + 16); bwt.esa.alloc(2 * bwt.r * U + 16);
            for (auto &pc : pieces) {
                if (pc.sc) d2d(bwt.ssa.get() + 2 * pc.so * U, pc.ssa.get(), 2 * pc.sc * U, ws.st);
                if (pc.ec) d2d(bwt.esa.get() + 2 * pc.eo * U, pc.esa.get(), 2 * pc.ec * U, ws.st);
            }
            stream_sync(ws.st);
        }
        bwt.ssa_off = bwt.esa_off = 0; bwt.ssa_cnt = bwt.esa_cnt = rssa ? bwt.r : 0;
        built = true; built_sa = sa; built_rssa = rssa;
        return PFB_OK;
    }

    // ---- document array (README.md:29-33 lists it as a feature to come, so there is no reference format):
    // <prefix>.da = one uint32 per BWT position, the index of the record (line of <prefix>.docs) that holds text position
    // SA[i]; the terminator's entry SA[0] = n counts to the last record.  Needs --print-docs and an SA on the device.
    int write_doc_array(const std::string &prefix) {
        try {
            if (!built || !(built_sa || built_rssa)) return fail(PFB_ERR_ARG, "the document array needs a BWT built with the suffix array (-s / -r)");
            if (docs.starts.empty()) return fail(PFB_ERR_ARG, "the document array needs the record starts (store_docs / --print-docs)");
            if (comm) return fail(PFB_ERR_ARG, "the document array is a single-GPU output");
            const u64 nd = docs.starts.size(), len = bwt.len;
            DevBuf<u64> ds(nd);
            h2d(ds.get(), docs.starts.data(), nd * 8, ws.st);
            DevBuf<u32> da(len ? len : 1);
            const u64 *sp = ds.get(); u32 *dp = da.get(); const u8 *sa_out = bwt.sa.get();
            const int U = prm.bits / 8;
            launch_for(len, ws.st, PFB_LAMBDA(u64 x) {
                const u64 v = U == 4 ? (u64)reinterpret_cast<const u32 *>(sa_out)[x] : reinterpret_cast<const u64 *>(sa_out)[x];
                u64 lo = 0, hi = nd;                    // largest d with starts[d] <= v
                while (hi - lo > 1) { const u64 mid = (lo + hi) >> 1; if (sp[mid] <= v) lo = mid; else hi = mid; }
                dp[x] = (u32)lo;
            });
            std::string e;
            if (dev_file(prefix + ".da", da.get(), len * 4, e)) return fail(PFB_ERR_IO, e);
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // ---- marker array: the join of src/mps_to_ma.cpp / include/marker_array.hpp:137-174 on the device
    // (markers.cuh).  write_marker_array takes the SA pfb_build_bwt left in HBM - the 8 bytes per base of `--stdout sa` never
    // leave the device -, marker_join reads a .sa file or stdin ("-") of u64 entries like the reference's tool.
    static u64 marker_chunk_entries() {
        u64 c = 1ull << 27;
        if (const char *e = getenv("PFB_MA_CHUNK")) { const long v = atol(e); if (v > 0) c = (u64)v; }      // test knob
        return c;
    }
    int marker_flush(FILE *f, const DevBuf<u64> &out, u64 nout) {
        std::string e;
        if (nout && write_dev_stream(f, out.get(), nout * 8, device, ws.st, e)) return fail(PFB_ERR_IO, e);
        return PFB_OK;
    }
    int write_marker_array(const std::string &mps, const std::string &outp) {
        try {
            if (comm) return fail(PFB_ERR_ARG, "the marker array is a single-GPU output");
            if (streamed && !built) {
                // the last build went straight to the files (pfb_build_bwt_to_files): join with the .sa file it wrote
                if (!streamed_sa) return fail(PFB_ERR_ARG, "the marker array after a streamed build needs the .sa file (build with sa)");
                return prm.bits == 64 ? marker_join_t<u64>(mps, streamed_prefix + ".sa", outp) : marker_join_t<u32>(mps, streamed_prefix + ".sa", outp);
            }
            if (!built || !(built_sa || built_rssa)) return fail(PFB_ERR_ARG, "the marker array needs a BWT built with the suffix array (-s / -r)");
            MarkerRuns mr;
            int rc = load_marker_runs(mps, mr, device, ws, err);
            if (rc) return fail(rc, err);
            FILE *f = fopen(outp.empty() ? "out" : outp.c_str(), "wb");
            if (!f) return fail(PFB_ERR_IO, "could not open " + outp + " for writing");
            const u64 len = bwt.len, CH = marker_chunk_entries();
            const int U = prm.bits / 8;
            MarkerCarry carry;
            DevBuf<u64> out;
            for (u64 c0 = 0; c0 < len && !rc; c0 += CH) {
                const u64 m = len - c0 < CH ? len - c0 : CH;
                const bool has_next = c0 + m < len;
                u64 next = 0, nout = 0;
                if (U == 4) {
                    const u32 *sp = reinterpret_cast<const u32 *>(bwt.sa.get()) + c0;
                    if (has_next) next = read_scalar(sp + m, ws.st);
                    rc = marker_join_chunk<u32>(mr, sp, m, has_next, next, carry, out, &nout, ws);
                } else {
                    const u64 *sp = reinterpret_cast<const u64 *>(bwt.sa.get()) + c0;
                    if (has_next) next = read_scalar(sp + m, ws.st);
                    rc = marker_join_chunk<u64>(mr, sp, m, has_next, next, carry, out, &nout, ws);
                }
                if (!rc) rc = marker_flush(f, out, nout);
            }
            fclose(f);
            return rc;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }
    // SAT = u64: the reference tool's input (`--stdout sa` of pfbwt-f64); u32: the .sa file a 32-bit streamed build left
    int marker_join(const std::string &mps, const std::string &sa_path, const std::string &outp) { return marker_join_t<u64>(mps, sa_path, outp); }
    template <class SAT>
    int marker_join_t(const std::string &mps, const std::string &sa_path, const std::string &outp) {
        try {
            MarkerRuns mr;
            int rc = load_marker_runs(mps, mr, device, ws, err);
            if (rc) return fail(rc, err);
            const bool from_stdin = sa_path == "-";
            FILE *in = from_stdin ? stdin : fopen(sa_path.c_str(), "rb");
            if (!in) return fail(PFB_ERR_IO, "error opening " + sa_path);
            FILE *f = fopen(outp.empty() ? "out" : outp.c_str(), "wb");
            if (!f) { if (!from_stdin) fclose(in); return fail(PFB_ERR_IO, "could not open " + outp + " for writing"); }
            const u64 CH = marker_chunk_entries();
            MarkerCarry carry;
            DevBuf<u64> out;
            DevBuf<SAT> dsa(CH + 1);
            // the SA goes through the host in chunks of CH entries plus one of look-ahead (a trailing partial entry is
            // dropped like fread(&s, 8, 1, fp) drops it)
            std::vector<SAT> hbuf(CH + 1);
            u64 have = fread(hbuf.data(), sizeof(SAT), CH + 1, in);
            while (have && !rc) {
                const bool has_next = have == CH + 1;
                const u64 m = has_next ? CH : have;
                h2d(dsa.get(), hbuf.data(), have * sizeof(SAT), ws.st);
                stream_sync(ws.st);
                const u64 next = has_next ? (u64)hbuf[CH] : 0;
                u64 nout = 0;
                rc = marker_join_chunk<SAT>(mr, dsa.get(), m, has_next, next, carry, out, &nout, ws);
                if (!rc) rc = marker_flush(f, out, nout);
                if (!has_next) break;
                hbuf[0] = (SAT)next;
                have = 1 + fread(hbuf.data() + 1, sizeof(SAT), CH, in);
            }
            fclose(f);
            if (!from_stdin) fclose(in);
            return rc;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // ---- -m: workspace larger than the device memory (include/file_wrappers.hpp:35-124, include/pfbwt.hpp:208-209; the
    // reference keeps gsa/glcp/ilist in mmap'ed files).  Here the per-suffix work arrays are what does not fit, so the
    // dictionary suffixes are sorted and emitted one key range ("bucket") after the other on the same GPU - the
    // range-partitioned sort of the multi-GPU path run as a loop - and every finished slice of .bwt/.sa/.ssa/.esa is
    // streamed to its place in the output files before the next bucket starts.
    u64 stream_buckets = 0, stream_peak = 0;      // of the last build_to_files
    static u64 parse_size_bytes(const std::string &s) {        // "48G", "512M", plain bytes
        char *end = nullptr;
        double v = strtod(s.c_str(), &end);
        if (end && (*end == 'G' || *end == 'g')) v *= (double)(1ull << 30);
        else if (end && (*end == 'M' || *end == 'm')) v *= (double)(1ull << 20);
        else if (end && (*end == 'K' || *end == 'k')) v *= 1024.0;
        return v > 0 ? (u64)v : 0;
    }
    // bytes of per-suffix / per-position work and output one full build holds at its peak (measured: 111 GB at D = 3.38e9,
    // n = 3.1e9, U = 8 with SA - keys 2 x 8, positions 2 x 4, meta, counts, lookups with SA; output 1 + U per position)
    u64 build_workspace_estimate(bool any_sa, bool rssa) const {
        const u64 U = (u64)prm.bits / 8;
        const u64 wide = parse.dsize > 0xFFFFFFF0ull ? 16ull : 0ull;      // 64-bit positions: 2 x 4 more per suffix, lookups always
        return parse.dsize * ((any_sa ? 38ull : 30ull) + wide) + (parse.n + 1) * (1 + (any_sa ? U : 0) + (rssa ? 18ull : 0));
    }
    int build_to_files(const std::string &prefix, bool sa, bool rssa, u64 limit_bytes) {
        try {
            if (!parsed) return fail(PFB_ERR_ARG, "no parse loaded");
            if (comm) return fail(PFB_ERR_ARG, "the bucketed build is a single-GPU mode");
            if (!parse.has_parse_bwt) { int rc = do_parse_bwt(); if (rc) return rc; }
            const bool any_sa = sa || rssa;
            const int U = prm.bits / 8;
            if (const char *e = getenv("PFB_HBM_LIMIT")) { const u64 v = parse_size_bytes(e); if (v) limit_bytes = v; }
#ifndef PFB_EMU
            if (!limit_bytes) {
                size_t fr = 0, tot = 0;
                cudaMemGetInfo(&fr, &tot);
                limit_bytes = (u64)fr + slab_pool().idle();
            }
#endif
            const u64 est = build_workspace_estimate(any_sa, rssa);
            const u64 resident = dev_bytes_live();
            u64 B = 1;
            if (limit_bytes && est + resident > limit_bytes) {
                if (resident + (est >> 6) >= limit_bytes) return fail(PFB_ERR_NOMEM, "the memory limit does not even hold the dictionary and the parse");
                B = (est + (limit_bytes - resident) - 1) / (limit_bytes - resident);
            }
            if (parse.dsize > 0xFFFFFFF0ull && B < parse.dsize / (1ull << 30) + 1) B = parse.dsize / (1ull << 30) + 1;   // a key range holds fewer than 2^32 suffixes
            if (const char *e = getenv("PFB_BUCKETS")) { const long v = atol(e); if (v > 0) B = (u64)v; }       // test knob
            if (B > (u64)KEY_MAX_DEST) B = KEY_MAX_DEST;
            bwt = DeviceBwt();
            built = false;
            const size_t peak0 = dev_bytes_peak();
            dev_bytes_peak() = dev_bytes_live();
            BucketCarry carry;
            std::string e;
            for (const char *ext : {"bwt", "sa", "ssa", "esa"}) {
                const bool want = !strcmp(ext, "bwt") || (!strcmp(ext, "sa") && sa) || ((!strcmp(ext, "ssa") || !strcmp(ext, "esa")) && rssa);
                if (want) unlink((prefix + "." + ext).c_str());
            }
            DeviceBwt last;
            for (u64 b = 0; b < B; ++b) {
                ShardSpec spec;
                spec.rank = (int)b; spec.world = (int)B;
                DeviceBwt out;
                int rc = build_bwt(parse, prm, sa, rssa, out, ws, tm, err, nullptr, &spec, &carry);
                if (rc) return fail(rc, err);
                if (write_dev_file(prefix + ".bwt", out.bwt.get(), out.len, out.off, out.n + 1, false, device, ws.st, e)) return fail(PFB_ERR_IO, e);
                if (sa && write_dev_file(prefix + ".sa", out.sa.get(), out.len * (size_t)U, out.off * (u64)U, (out.n + 1) * (u64)U, false, device, ws.st, e)) return fail(PFB_ERR_IO, e);
                if (rssa) {
                    if (write_dev_file(prefix + ".ssa", out.ssa.get(), 2 * out.ssa_cnt * (size_t)U, 2 * out.ssa_off * (u64)U, 2 * (out.ssa_off + out.ssa_cnt) * (u64)U, false, device, ws.st, e)) return fail(PFB_ERR_IO, e);
                    if (write_dev_file(prefix + ".esa", out.esa.get(), 2 * out.esa_cnt * (size_t)U, 2 * out.esa_off * (u64)U, 2 * (out.esa_off + out.esa_cnt) * (u64)U, false, device, ws.st, e)) return fail(PFB_ERR_IO, e);
                }
                last.n = out.n; last.r = out.r; last.easy = out.easy; last.hard = out.hard;
            }
            bwt.n = last.n; bwt.r = last.r; bwt.easy = last.easy; bwt.hard = last.hard;      // sizes only: the slices went to the files
            streamed = true; streamed_sa = sa; streamed_prefix = prefix;
            stream_buckets = B; stream_peak = dev_bytes_peak();
            if (dev_bytes_peak() < peak0) dev_bytes_peak() = peak0;
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }
    bool streamed = false;            // the last build went straight to files (sizes available, arrays not)
    bool streamed_sa = false; std::string streamed_prefix;

    // drop the text and the staged input (records in HBM from the device reader / add_records, host staging) once the
    // parse exists: only dict/parse are needed downstream.  Parsing again needs pfb_clear + the input once more.
    void release_text() {
        text.buf.release(); text_ready = false;
        if (parsed && !segs.empty()) {
            for (auto &s : segs) s.dev.release();
            raw.release();
            input_released = true;
        }
    }

    // ---- copies to host -----------------------------------------------------------------
    template <class T>
    std::vector<T> to_host(const DevBuf<T> &b, size_t count) {
        std::vector<T> v(count);
        if (count) { d2h(v.data(), b.get(), count * sizeof(T), ws.st); stream_sync(ws.st); }
        return v;
    }

    // device array -> file (devio.cuh); integer arrays are brought to the file's width U on the device first
    int dev_file(const std::string &path, const void *d, size_t bytes, std::string &e) {
        return write_dev_file(path, d, bytes, 0, bytes, true, device, ws.st, e);
    }
    int dev_file_u32(const std::string &path, const DevBuf<u32> &b, size_t count, int U, std::string &e) {
        if (U == 4) return dev_file(path, b.get(), count * 4, e);
        DevBuf<u64> t(count ? count : 1);
        const u32 *s = b.get(); u64 *d = t.get();
        launch_for(count, ws.st, PFB_LAMBDA(u64 i) { d[i] = s[i]; });
        return dev_file(path, t.get(), count * 8, e);
    }
    int dev_file_u64(const std::string &path, const DevBuf<u64> &b, size_t count, int U, std::string &e) {
        if (U == 8) return dev_file(path, b.get(), count * 8, e);
        DevBuf<u32> t(count ? count : 1);
        const u64 *s = b.get(); u32 *d = t.get();
        launch_for(count, ws.st, PFB_LAMBDA(u64 i) { d[i] = (u32)s[i]; });
        return dev_file(path, t.get(), count * 4, e);
    }

    // ---- files: save_parser (include/pfbwt_io.hpp:234-249) + parse-bwt (src/pfbwt-f.cpp:229-235)
    // only_parse_bwt: just .bwlast/.ilist(/.bwsai) - run_parser writes the four parse files in its second task and the
    // parse-BWT files in its third (src/pfbwt-f.cpp:221-236), so a failing bwt_of_parse leaves the former behind
    int save_parse(const std::string &prefix, bool with_parse_bwt, bool with_sai, bool only_parse_bwt = false) {
        try {
            if (!parsed) return fail(PFB_ERR_ARG, "parse not finalized");
            const int U = prm.bits / 8;
            std::string e;
            if (only_parse_bwt) {
                if (!parse.has_parse_bwt) return fail(PFB_ERR_ARG, "parse BWT not computed");
                if (dev_file(prefix + ".bwlast", parse.bwlast.get(), parse.P + 1, e)) return fail(PFB_ERR_IO, e);
                if (dev_file_u32(prefix + ".ilist", parse.ilist, parse.P + 1, U, e)) return fail(PFB_ERR_IO, e);
                if (with_sai && dev_file_u64(prefix + ".bwsai", parse.bwsai, parse.P + 1, U, e)) return fail(PFB_ERR_IO, e);
                return PFB_OK;
            }
            if (dev_file(prefix + ".dict", parse.dict.get(), parse.dsize, e)) return fail(PFB_ERR_IO, e);
            if (dev_file_u32(prefix + ".occ", parse.occ, parse.Dw, U, e)) return fail(PFB_ERR_IO, e);
            if (dev_file(prefix + ".parse", parse.parse.get(), parse.P * 4, e)) return fail(PFB_ERR_IO, e);
            if (store_docs) {   // include/pfbwt_io.hpp:224-231
                FILE *f = fopen((prefix + ".docs").c_str(), "w");
                if (!f) return fail(PFB_ERR_IO, "could not open docs file");
                for (size_t i = 0; i < docs.starts.size(); ++i)
                    fprintf(f, "%s %lu\n", i < docs.names.size() ? docs.names[i].c_str() : "", (unsigned long)docs.starts[i]);
                fclose(f);
            }
            { FILE *f = fopen((prefix + ".n").c_str(), "w"); if (!f) return fail(PFB_ERR_IO, "could not open .n file");
              fprintf(f, "%lu\n", (unsigned long)parse.n); fclose(f); }
            if (with_parse_bwt) {
                if (!parse.has_parse_bwt) return fail(PFB_ERR_ARG, "parse BWT not computed");
                if (dev_file(prefix + ".bwlast", parse.bwlast.get(), parse.P + 1, e)) return fail(PFB_ERR_IO, e);
                if (dev_file_u32(prefix + ".ilist", parse.ilist, parse.P + 1, U, e)) return fail(PFB_ERR_IO, e);
                if (with_sai && dev_file_u64(prefix + ".bwsai", parse.bwsai, parse.P + 1, U, e)) return fail(PFB_ERR_IO, e);
            }
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // ---- the same files written in the background: the GB-sized parse files leave through a writer thread (own pinned
    // staging buffers and copy streams, devio.cuh) while the device goes on with the parse BWT and the BWT build - file
    // writing, not the device, is what a file-to-file run waits for.  The arrays written are those of `parse`, which stay
    // untouched until wait_saves() (called by everything that replaces the parse).
    struct SaveJob {
        struct Item { std::string path; const void *d = nullptr; size_t bytes = 0; };
        std::vector<Item> items;
        std::vector<DevBuf<u8>> held;        // width-converted images
        std::thread th;
        std::string err;
        int rc = 0;
    };
    std::vector<std::unique_ptr<SaveJob>> save_jobs;
    int wait_saves() {
        int rc = PFB_OK;
        for (auto &j : save_jobs) {
            if (j->th.joinable()) j->th.join();
            if (j->rc && !rc) rc = fail(PFB_ERR_IO, j->err);
        }
        save_jobs.clear();
        return rc;
    }
    template <class T, class O>
    const void *widen_into(SaveJob &job, const DevBuf<T> &b, size_t count) {          // device array as O-wide integers
        if (sizeof(T) == sizeof(O)) return b.get();
        DevBuf<u8> t((count ? count : 1) * sizeof(O));
        const T *s = b.get(); O *d = reinterpret_cast<O *>(t.get());
        launch_for(count, ws.st, PFB_LAMBDA(u64 i) { d[i] = (O)s[i]; });
        job.held.push_back(std::move(t));
        return job.held.back().get();
    }
    int save_parse_async(const std::string &prefix, int mode, bool with_sai) {          // mode as pfb_save_parse's with_parse_bwt
        try {
            if (!parsed) return fail(PFB_ERR_ARG, "parse not finalized");
            if (mode && !parse.has_parse_bwt) return fail(PFB_ERR_ARG, "parse BWT not computed");
            const int U = prm.bits / 8;
            std::unique_ptr<SaveJob> job(new SaveJob());
            job->held.reserve(4);
            auto add = [&](const char *ext, const void *d, size_t bytes) { SaveJob::Item it; it.path = prefix + "." + ext; it.d = d; it.bytes = bytes; job->items.push_back(it); };
            if (mode != 2) {
                add("dict", parse.dict.get(), parse.dsize);
                add("occ", U == 4 ? widen_into<u32, u32>(*job, parse.occ, parse.Dw) : widen_into<u32, u64>(*job, parse.occ, parse.Dw), parse.Dw * (size_t)U);
                add("parse", parse.parse.get(), parse.P * 4);
                if (store_docs) {   // include/pfbwt_io.hpp:224-231
                    FILE *f = fopen((prefix + ".docs").c_str(), "w");
                    if (!f) return fail(PFB_ERR_IO, "could not open docs file");
                    for (size_t i = 0; i < docs.starts.size(); ++i)
                        fprintf(f, "%s %lu\n", i < docs.names.size() ? docs.names[i].c_str() : "", (unsigned long)docs.starts[i]);
                    fclose(f);
                }
                FILE *f = fopen((prefix + ".n").c_str(), "w");
                if (!f) return fail(PFB_ERR_IO, "could not open .n file");
                fprintf(f, "%lu\n", (unsigned long)parse.n);
                fclose(f);
            }
            if (mode) {
                add("bwlast", parse.bwlast.get(), parse.P + 1);
                add("ilist", U == 4 ? widen_into<u32, u32>(*job, parse.ilist, parse.P + 1) : widen_into<u32, u64>(*job, parse.ilist, parse.P + 1), (parse.P + 1) * (size_t)U);
                if (with_sai) add("bwsai", U == 8 ? widen_into<u64, u64>(*job, parse.bwsai, parse.P + 1) : widen_into<u64, u32>(*job, parse.bwsai, parse.P + 1), (parse.P + 1) * (size_t)U);
            }
            stream_sync(ws.st);                 // the images are complete; the writer does not touch the pipeline's stream
            SaveJob *jp = job.get();
            const int dev = device;
            jp->th = std::thread([jp, dev]() {
                for (auto &it : jp->items) {
                    std::string e;
                    if (write_dev_file(it.path, it.d, it.bytes, 0, it.bytes, true, dev, 0, e, /*sync_producer=*/false)) { jp->rc = -1; jp->err = e; return; }
                }
            });
            save_jobs.push_back(std::move(job));
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // a file of `count` integers of the binary's width U -> device array of the pipeline's width, converted on the device
    // (0 ok - present says whether the file exists -, -1 wrong size)
    template <class TD>
    int load_uint_file(const std::string &path, u64 count, DevBuf<TD> &dst, bool &present) {
        const int U = prm.bits / 8;
        DevBuf<u8> rawb;
        u64 sz = 0;
        std::string e2;
        present = read_whole_file_to_dev(path, rawb, sz, 64, device, ws.st, e2) == 0;
        if (!present) return 0;
        if (sz != count * (u64)U) return -1;
        dst.alloc(count ? count : 1);
        TD *d = dst.get();
        if (U == 4) { const u32 *s4 = reinterpret_cast<const u32 *>(rawb.get()); launch_for(count, ws.st, PFB_LAMBDA(u64 i) { d[i] = (TD)s4[i]; }); }
        else { const u64 *s8 = reinterpret_cast<const u64 *>(rawb.get()); launch_for(count, ws.st, PFB_LAMBDA(u64 i) { d[i] = (TD)s8[i]; }); }
        stream_sync(ws.st);
        return 0;
    }

    // ---- load: PrefixFreeBWT ctor (include/pfbwt.hpp:64-81) / load_parser (include/pfbwt_io.hpp:211-222)
    int load_parse(const std::string &prefix, bool with_parse_bwt, bool with_sai) {
        try {
            wait_saves();
            const int U = prm.bits / 8;
            std::string e;
            // the dictionary (GBs) goes from the file straight to HBM; it is checked and cut into words there
            parse = DeviceParse();
            u64 dsize = 0;
            if (read_whole_file_to_dev(prefix + ".dict", parse.dict, dsize, 64, device, ws.st, e)) return fail(PFB_ERR_IO, "bad dict file\n");
            if (dsize < 2) return fail(PFB_ERR_IO, "bad dict file\n");
            if (dsize >= DICT_MAX_BYTES) return fail(PFB_ERR_LIMIT, "dictionary larger than 2^40 bytes is not supported");
            parse.dsize = dsize;
            {
                const u8 *dp = parse.dict.get();
                DevBuf<u32> st3(3);        // [0] byte outside the alphabet, [1] bad tail
                dev_memset(st3.get(), 0, 12, ws.st);
                u32 *sp = st3.get();
                DevBuf<u8> isend(dsize);
                DevBuf<u32> rank(dsize);
                u8 *ie = isend.get();
                launch_for(dsize, ws.st, PFB_LAMBDA(u64 i) {
                    const u8 c = dp[i];
                    if (i + 1 == dsize) { ie[i] = 0; if (c != ENDOFDICT) sp[1] = 1u; return; }
                    if (!(c == 1 || c == 2 || c == '-' || c == 'A' || c == 'C' || c == 'G' || c == 'N' || c == 'T')) sp[0] = 1u;
                    if (i + 2 == dsize && c != ENDOFWORD) sp[1] = 1u;          // the last word must end right before EndOfDict
                    ie[i] = c == ENDOFWORD ? 1 : 0;
                });
                const u64 Dw = scan_flags(isend.get(), rank.get(), dsize, ws);
                u32 bad[3];
                d2h(bad, st3.get(), 12, ws.st);
                stream_sync(ws.st);
                if (bad[1] || Dw == 0) return fail(PFB_ERR_IO, "bad dict file\n");
                if (bad[0]) return fail(PFB_ERR_IO, "dict file contains a byte outside the parser's alphabet");
                parse.Dw = Dw;
                parse.wstart.alloc(Dw + 1);
                u64 *wsp = parse.wstart.get();
                const u32 *rk = rank.get();
                launch_for(dsize, ws.st, PFB_LAMBDA(u64 i) {
                    if (i == 0) wsp[0] = 0;
                    if (ie[i]) wsp[rk[i] + 1] = i + 1;
                });
            }
            // .parse / .occ / .ilist / .bwsai go from their files straight to HBM (devio.cuh) and are brought to the device
            // widths there: no element-wise loops on a host core
            {
                DevBuf<u8> rawp;
                u64 psz = 0;
                if (read_whole_file_to_dev(prefix + ".parse", rawp, psz, 64, device, ws.st, e)) return fail(PFB_ERR_IO, "bad parse file\n");
                parse.P = psz / 4;
                parse.parse.alloc(parse.P + 1);
                d2d(parse.parse.get(), rawp.get(), parse.P * 4, ws.st);
                dev_memset(parse.parse.get() + parse.P, 0, 4, ws.st);
                stream_sync(ws.st);
            }
            bool have_occ = false;
            if (load_uint_file(prefix + ".occ", parse.Dw, parse.occ, have_occ)) have_occ = false;      // wrong size: recount below
            if (!have_occ) { parse.occ.alloc(parse.Dw); dev_memset(parse.occ.get(), 0, parse.Dw * 4, ws.st); }
            // n from <prefix>.n (src/pfbwt-f.cpp:255-270) or from the phrase lengths
            {
                FILE *f = fopen((prefix + ".n").c_str(), "r");
                unsigned long nn = 0;
                if (f && fscanf(f, "%lu", &nn) == 1) parse.n = nn;
                else {
                    DevBuf<u64> tot(1);
                    dev_memset(tot.get(), 0, 8, ws.st);
                    const u32 *pr_ = parse.parse.get(); const u64 *wsr = parse.wstart.get(); u64 *tp = tot.get();
                    const u64 Dw_ = parse.Dw, w_ = (u64)prm.w;
                    launch_for(parse.P, ws.st, PFB_LAMBDA(u64 k) {
                        const u64 d = (u64)pr_[k] - 1;
                        if (d < Dw_) atomic_add(tp, (u64)(wsr[d + 1] - 1 - wsr[d]) - (k ? w_ : 0));
                    });
                    parse.n = read_scalar(tot.get(), ws.st) - 1 - (u64)prm.w;
                }
                if (f) fclose(f);
            }
            stream_sync(ws.st);
            if (with_parse_bwt) {
                u64 bsz = 0;
                if (read_whole_file_to_dev(prefix + ".bwlast", parse.bwlast, bsz, 64, device, ws.st, e)) return fail(PFB_ERR_IO, e);
                bool have = false;
                if (bsz != parse.P + 1 || load_uint_file(prefix + ".ilist", parse.P + 1, parse.ilist, have) || !have) return fail(PFB_ERR_IO, "bwlast/ilist size mismatch");
                if (with_sai) {
                    if (load_uint_file(prefix + ".bwsai", parse.P + 1, parse.bwsai, have)) return fail(PFB_ERR_IO, "bwsai size mismatch");
                    if (!have) return fail(PFB_ERR_IO, "error opening " + prefix + ".bwsai");
                }
                parse.has_parse_bwt = true;
            }
            {
                // a stale or mismatched file set must end in "bad parse file", not in out-of-range device reads later: every
                // rank of .parse in 1..Dw, .occ = the histogram of .parse, every .ilist entry a position of the parse BWT
                DevBuf<u32> bad(1), hist(parse.Dw);
                dev_memset(bad.get(), 0, 4, ws.st);
                dev_memset(hist.get(), 0, parse.Dw * 4, ws.st);
                const u32 *pr_ = parse.parse.get(); const u32 *oc = parse.occ.get(); u32 *hp = hist.get(); u32 *bp = bad.get();
                const u64 P_ = parse.P, Dw_ = parse.Dw;
                launch_for(P_, ws.st, PFB_LAMBDA(u64 k) {
                    const u32 r = pr_[k];
                    if (r < 1 || (u64)r > Dw_) bp[0] = 1u; else atomic_add(&hp[r - 1], 1u);
                });
                if (have_occ) launch_for(Dw_, ws.st, PFB_LAMBDA(u64 d) { if (hp[d] != oc[d]) bp[0] = 1u; });
                else d2d(parse.occ.get(), hist.get(), Dw_ * 4, ws.st);           // no .occ file: the counts of the parse
                if (with_parse_bwt) {
                    const u32 *il = parse.ilist.get();
                    launch_for(P_ + 1, ws.st, PFB_LAMBDA(u64 i) { if ((u64)il[i] > P_) bp[0] = 1u; });
                }
                if (P_ == 0 || read_scalar(bad.get(), ws.st)) { parse = DeviceParse(); return fail(PFB_ERR_IO, "bad parse file\n"); }
            }
            docs = FastaResult();
            if (store_docs) {          // load_doc_info, include/pfbwt_io.hpp:177-205 ("<name> <start>" per line)
                FILE *f = fopen((prefix + ".docs").c_str(), "r");
                if (!f) return fail(PFB_ERR_IO, "error opening doc file " + prefix + ".docs");
                char *line = NULL;
                size_t cap = 0;
                while (getline(&line, &cap, f) >= 0) {
                    char *save = NULL;
                    char *tok = strtok_r(line, " ", &save);
                    if (!tok) { free(line); fclose(f); return fail(PFB_ERR_IO, "error parsing name in doc file line"); }
                    std::string nm(tok);
                    tok = strtok_r(NULL, " ", &save);
                    if (!tok) { free(line); fclose(f); return fail(PFB_ERR_IO, "error parsing size in doc file line: " + nm); }
                    docs.names.push_back(nm);
                    docs.starts.push_back(strtoull(tok, NULL, 10));
                }
                free(line);
                fclose(f);
                docs.nrec = docs.starts.size();
            }
            parsed = true; built = false;
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // .last (P bytes = last_) and .sai (P x U = sai_): extensions include/utils.h:22,24 defines and only
    // tests/test_parser.cpp:126,139 of the reference reads; written on request only
    int save_last_sai(const std::string &prefix, bool with_last, bool with_sai) {
        try {
            if (!parsed) return fail(PFB_ERR_ARG, "parse not finalized");
            if (parse.last.size() < parse.P || parse.sai.size() < parse.P) return fail(PFB_ERR_ARG, "last / sai are not held by a parse loaded from files");
            std::string e;
            if (with_last && dev_file(prefix + ".last", parse.last.get(), parse.P, e)) return fail(PFB_ERR_IO, e);
            if (with_sai && dev_file_u64(prefix + ".sai", parse.sai, parse.P, prm.bits / 8, e)) return fail(PFB_ERR_IO, e);
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

    // ---- operator+=(include/pfparser.hpp:194-263): the result is defined as the parse of
    // the concatenated record list, so the rhs parse is expanded back to its text on the
    // device (phrases minus their w-byte overlaps) and queued as a device segment.
    int append_parse_of(Engine &rhs) {
        try {
            if (!rhs.parsed) return fail(PFB_ERR_ARG, "rhs parse not finalized");
            if (rhs.prm.w != prm.w) return fail(PFB_ERR_ARG, "invalid w");
            if (rhs.prm.p != prm.p) return fail(PFB_ERR_ARG, "invalid p");
            // a parser that was loaded from files holds no text: its own parse becomes the first segment
            if (&rhs != this && parsed && segs.empty() && pending_n == 0) { int rc0 = append_parse_of(*this); if (rc0) return rc0; }
            const DeviceParse &rp = rhs.parse;
            const u64 P = rp.P; const int w = prm.w;
            DevBuf<u64> clen(P + 1), coff(P + 1);
            {
                const u32 *pr = rp.parse.get(); const u64 *wsr = rp.wstart.get(); u64 *cl = clen.get();
                launch_for(P + 1, ws.st, PFB_LAMBDA(u64 k) {
                    if (k == P) { cl[k] = 0; return; }
                    u32 d = pr[k] - 1;
                    u64 len = wsr[d + 1] - 1 - wsr[d];
                    cl[k] = k ? len - (u64)w : len;
                });
            }
            exclusive_sum<u64, u64>(clen.get(), coff.get(), P + 1, ws.sc, ws.st);
            u64 tlen = read_scalar(coff.get() + P, ws.st);      // |T'| = n + w + 1
            if (tlen < (u64)w + 1) return fail(PFB_ERR_ARG, "rhs parser malformed");
            u64 n = tlen - (u64)w - 1;
            DevBuf<u8> tp(tlen);
            {
                const u32 *pr = rp.parse.get(); const u64 *wsr = rp.wstart.get(); const u8 *dict = rp.dict.get();
                const u64 *co = coff.get(); u8 *out = tp.get();
                launch_for(P * 32, ws.st, PFB_LAMBDA(u64 i) {
                    u64 k = i / 32, lane = i % 32;
                    u32 d = pr[k] - 1;
                    u64 len = wsr[d + 1] - 1 - wsr[d];
                    u64 skip = k ? (u64)w : 0;
                    const u8 *src = dict + wsr[d] + skip;
                    u8 *dst = out + co[k];
                    for (u64 j = lane; j < len - skip; j += 32) dst[j] = src[j];
                });
            }
            Segment s;
            s.on_device = true; s.len = n;
            s.dev.alloc(n);
            d2d(s.dev.get(), tp.get() + 1, n, ws.st);
            stream_sync(ws.st);
            for (size_t i = 0; i < rhs.docs.starts.size(); ++i) docs.starts.push_back(rhs.docs.starts[i] + pending_n);
            for (auto &nm : rhs.docs.names) docs.names.push_back(nm);
            segs.push_back(std::move(s));
            pending_n += n;
            invalidate();
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }

#include "shard.inl"

    // ---- output files: src/pfbwt-f.cpp:298-349 -------------------------------------------------
    int write_bwt_files(const std::string &prefix, bool sa, bool rssa, const std::string &stdout_ext) {
        try {
            if (!built) return fail(PFB_ERR_ARG, "BWT not built");
            std::string e;
            const int U = prm.bits / 8;
            // a rank of a multi-GPU job writes its slice at its offset of the common file
            auto emit = [&](const std::string &ext, const DevBuf<u8> &b, size_t bytes, u64 off, u64 final_size) -> int {
                if (comm) {
                    if (!stdout_ext.empty()) { e = "--stdout is not available in a multi-GPU job"; return -1; }
                    return write_dev_file(prefix + "." + ext, b.get(), bytes, off, final_size, false, device, ws.st, e);
                }
                if (stdout_ext == ext) return write_dev_stream(stdout, b.get(), bytes, device, ws.st, e);
                return write_dev_file(prefix + "." + ext, b.get(), bytes, 0, bytes, true, device, ws.st, e);
            };
            if (emit("bwt", bwt.bwt, bwt.len, bwt.off, bwt.n + 1)) return fail(PFB_ERR_IO, e);
            if (sa) { if (!built_sa && !built_rssa) return fail(PFB_ERR_ARG, "SA not built");
                      if (emit("sa", bwt.sa, bwt.len * (size_t)U, bwt.off * (u64)U, (bwt.n + 1) * (u64)U)) return fail(PFB_ERR_IO, e); }
            if (rssa) {
                if (!built_rssa) return fail(PFB_ERR_ARG, "run-length SA not built");
                if (emit("ssa", bwt.ssa, 2 * bwt.ssa_cnt * (size_t)U, 2 * bwt.ssa_off * (u64)U, 2 * bwt.r * (u64)U)) return fail(PFB_ERR_IO, e);
                if (emit("esa", bwt.esa, 2 * bwt.esa_cnt * (size_t)U, 2 * bwt.esa_off * (u64)U, 2 * bwt.r * (u64)U)) return fail(PFB_ERR_IO, e);
            }
            return PFB_OK;
        } catch (const Error &e) { return fail(e.code, e.what()); }
    }
};

}  // namespace pfb
