// ingest.cuh -- FASTA ingestion on the device.
//
// The reference reads records with kseq (include/kseq.h:178-216, driven by PfParser::add_fasta,
// include/pfparser.hpp:299-337): header lines are skipped, line ends are dropped, every record is
// followed by w filler 'A's.  host_io.hpp does that byte by byte on one core; here the raw bytes of a
// plain FASTA file go to HBM as they are read (devio.cuh: several readers, each pread()ing into its own pinned buffer)
// and three kernels over 64-byte tiles do the record work:
//   A  kind of the last line that starts inside each tile (header / sequence) -> running maximum
//      over the tiles = kind of the line every tile starts in
//   B  per tile: bytes kept, headers started                                 -> two exclusive sums
//   C  kept bytes to their place in S, w x 'A' after every record, header offsets and record starts
// Everything kseq treats specially beyond plain FASTA - FASTQ ('@' records, '+' lines), '\r' before a
// line end - is only DETECTED here; such inputs take the host reader (host_io.hpp), so behaviour is the
// same for every input.  gzip files and stdin are inflated / drained by zlib on the host and then split here
// as well.  PFB_INGEST=host keeps the host reader for all inputs.
#pragma once
#include "pipeline.cuh"
#include "devio.cuh"
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

namespace pfb {

static const u64 ING_TILE = 64;

struct IngestResult {
    DevBuf<u8> seg;                 // records + padding, ready to be copied into S
    u64 len = 0;                    // bytes in seg
    u64 nrec = 0, total_len = 0;    // records, sum of record lengths (no padding)
    std::vector<u64> starts;        // offset of every record in seg (store_docs)
    std::vector<std::string> names; // header up to the first white space (store_docs)
    HostBytes hostraw;              // the whole inflated / drained stream when the input was not a plain regular file
    float device_ms = 0;
};

// one tile: calls hdr(i) at every header start and keep(i, c) for every sequence byte, in order.
// `kind` is the kind of the line the tile starts in (1 sequence, 2 header); returns the kind of the last line
// that starts inside the tile, 0 when none does.  *bad is raised for what only the host reader handles.
template <class FH, class FK>
PFB_HD inline u32 ingest_walk_tile(const u8 *R, u64 m, u64 p0, u64 t, u32 kind, u32 *bad, FH hdr, FK keep) {
    const u64 base = t * ING_TILE;
    u64 lo = base, hi = base + ING_TILE;
    if (hi > m) hi = m;
    if (lo < p0) lo = p0;
    if (lo >= hi) return 0;
    // eight 8-byte loads per tile (R is 512-byte aligned and has ING_TILE spare bytes behind m); the bytes of a word
    // are walked with the loop unrolled, the words are not (keeps the three kernels small)
    u32 started = 0;
    u8 prev = lo == p0 ? (u8)'\n' : R[lo - 1];
#pragma unroll 1
    for (int q = 0; q < (int)ING_TILE / 8; ++q) {
        const u64 wbase = base + 8ull * (u64)q;
        if (wbase + 8 <= lo || wbase >= hi) continue;
        u64 wd;
#ifdef __CUDA_ARCH__
        wd = *reinterpret_cast<const u64 *>(R + wbase);
#else
        memcpy(&wd, R + wbase, 8);
#endif
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const u64 i = wbase + (u64)b;
            const u8 c = (u8)(wd >> (8 * b));
            if (i >= lo && i < hi) {
                if (prev == '\n') {                      // kseq.h:194-198: first character of a line
                    if (c == '>') { kind = 2; hdr(i); }
                    else { kind = 1; if (c == '+' || c == '@') *bad = 1u; }
                    started = kind;
                }
                if (c == '\r') *bad = 1u;
                if (kind == 1 && c != '\n') keep(i, c);
                prev = c;
            }
        }
    }
    return started ? kind : 0;
}

// 0 = done; 1 = not a case for the device reader, nothing consumed (caller uses read_fasta on the path);
// 2 = not a case for the device reader, but the stream has been read: its bytes are in out.hostraw (caller uses
// read_fasta_mem); -1 = error (err set)
// Range mode (rg != nullptr; pfb_add_fasta_range): only bytes [rg->begin, rg->end) of a plain file are taken, both ends
// moved forward to a line start.  A range that does not start with a header continues the record of the range before it
// (its bytes before the first header carry no new record; the w filler 'A's follow them when a header shows up in the
// range); the range's last record is left open unless the next range starts with a header or the file ends.  FASTQ,
// CR LF, gzip and streams are refused in this mode (-1).
struct IngestRange { u64 begin = 0, end = 0; bool open_end = false; /* out */ };

// first line start at or after x (x itself when the byte before it is a line end)
inline int ingest_line_start(int fd, u64 size, u64 x, u64 &out) {
    if (x == 0) { out = 0; return 0; }
    if (x >= size) { out = size; return 0; }
    char buf[4096];
    u64 pos = x - 1;
    while (pos < size) {
        ssize_t got = pread(fd, buf, sizeof(buf), (off_t)pos);
        if (got < 0) return -1;
        if (got == 0) break;
        const void *q = memchr(buf, '\n', (size_t)got);
        if (q) { out = pos + (u64)((const char *)q - buf) + 1; return 0; }
        pos += (u64)got;
    }
    out = size;
    return 0;
}

inline int ingest_fasta_device(const std::string &path, int w, bool store_docs, int device, Workspace &ws, IngestResult &out, std::string &err,
                               IngestRange *rg = nullptr) {
    static const bool host_only = getenv("PFB_INGEST") && !strcmp(getenv("PFB_INGEST"), "host");
    if (host_only && !rg) return 1;
    // a regular uncompressed file is read straight into HBM by several readers; anything else (gzip, stdin, a pipe) is
    // inflated / drained by zlib into host memory first and uploaded in one piece - the records are split on the device
    // either way ("gzip-inflate stays on CPU, record splitting on device")
    int fd = -1;
    bool stream = path == "-";
    u64 m = 0, fbase = 0;                                    // bytes taken, file offset of the first one
    u32 lead = 0;                                            // range mode: the range starts inside a record
    bool closed = true;                                      // the last record ends with the bytes taken
    if (rg) {
        if (stream) { err = "a byte range needs a regular file"; return -1; }
        fd = open(path.c_str(), O_RDONLY);
        struct stat sb;
        if (fd < 0 || fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) { if (fd >= 0) close(fd); err = "failed to open file!\n"; return -1; }
        const u64 size = (u64)sb.st_size;
        u64 a = 0, b = 0;
        u8 c0[2] = {0, 0};
        if (size < 2 || pread(fd, c0, 2, 0) != 2 || c0[0] != '>') { close(fd); err = "a byte range needs a plain FASTA file that starts with a header"; return -1; }
        if (ingest_line_start(fd, size, rg->begin, a) || ingest_line_start(fd, size, rg->end, b)) { close(fd); err = "error reading input"; return -1; }
        if (a >= b) { close(fd); out.len = 0; rg->open_end = b < size; rg->begin = a; rg->end = b; return 0; }     // nothing in this range
        u8 ca = 0, cb = 0;
        if (pread(fd, &ca, 1, (off_t)a) != 1 || (b < size && pread(fd, &cb, 1, (off_t)b) != 1)) { close(fd); err = "error reading input"; return -1; }
        lead = (a > 0 && ca != '>') ? 1u : 0u;
        closed = b >= size || cb == '>';
        rg->begin = a; rg->end = b; rg->open_end = !closed;
        m = b - a; fbase = a;
    } else if (!stream) {
        fd = open(path.c_str(), O_RDONLY);
        if (fd < 0) return 1;                                // the host reader reports the error its way
        struct stat sb;
        if (fstat(fd, &sb) != 0) { close(fd); return 1; }
        u8 magic[2] = {0, 0};
        if (!S_ISREG(sb.st_mode) || sb.st_size < 2 || pread(fd, magic, 2, 0) != 2 || (magic[0] == 0x1f && magic[1] == 0x8b)) {
            close(fd); fd = -1; stream = true;
        } else m = (u64)sb.st_size;
    }
    u64 p0 = ~0ull;                                          // kseq.h:183-187: first '>' or '@' anywhere
    u8 first_marker = 0;
    auto look_for_marker = [&](const u8 *b, u64 off, u64 len) {
        if (p0 != ~0ull) return;
        for (u64 i = 0; i < len; ++i) if (b[i] == '>' || b[i] == '@') { p0 = off + i; first_marker = b[i]; return; }
    };
    if (stream) {
        gzFile fp = path == "-" ? gzdopen(fileno(stdin), "r") : gzopen(path.c_str(), "r");
        if (fp == NULL) return 1;
        gzbuffer(fp, 1 << 20);
        HostBytes &hb = out.hostraw;
        const size_t BUFSZ = 16u << 20;
        for (;;) {
            hb.reserve(hb.n + BUFSZ);
            int got = gzread(fp, hb.p + hb.n, (unsigned)BUFSZ);
            if (got < 0) { gzclose(fp); err = "error reading input"; return -1; }
            if (got == 0) break;
            hb.n += (size_t)got;
        }
        gzclose(fp);
        m = hb.n;
        if (m < 2) return 2;
        look_for_marker(hb.p, 0, m);
        if (p0 == ~0ull || first_marker != '>') return 2;
    }
    if (!rg && m >= 2) {
        // kseq.h:188: a header marker as the very last byte of the input starts no record (its name read hits EOF)
        u8 tail2[2] = {0, 0};
        bool have = false;
        if (stream) { tail2[0] = out.hostraw.p[m - 2]; tail2[1] = out.hostraw.p[m - 1]; have = true; }
        else have = pread(fd, tail2, 2, (off_t)(m - 2)) == 2;
        if (have && tail2[1] == '>' && tail2[0] == '\n') --m;
    }
    const int decline = rg ? -1 : stream ? 2 : 1;
    if (rg) err = "a byte range needs a plain FASTA file (no FASTQ records, no CR LF line ends) that fits the device";

    pfbStream_t st = ws.st;
#ifndef PFB_EMU
    {
        // raw bytes + records + 40 bytes of tile bookkeeping per 64 bytes: leave very large inputs to the host reader
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || (double)m * 2.8 + (256u << 20) > (double)free_b + (double)slab_pool().idle()) {
            if (fd >= 0) close(fd);
            return decline;
        }
    }
#endif
    DevBuf<u8> R(m + ING_TILE);
    if (stream) {
        h2d(R.get(), out.hostraw.p, m, st);
    } else {
#ifdef PFB_EMU
        u64 off = 0;
        while (off < m) {
            ssize_t got = pread(fd, R.get() + off, (size_t)(m - off < (1u << 26) ? m - off : (1u << 26)), (off_t)(fbase + off));
            if (got < 0) { close(fd); err = "error reading input"; return -1; }
            if (got == 0) break;
            look_for_marker(R.get() + off, off, (u64)got);
            off += (u64)got;
        }
        if (off != m) { close(fd); return rg ? -1 : 1; }
#else
        // the first header must show up in the first 16 MB chunk; otherwise the host reader takes the file
        std::string e;
        stream_sync(st);             // R may reuse a block that work queued on st still reads; the copies run on other streams
        const int rc = read_file_to_dev(fd, m, R.get(), device, [&](const u8 *b, size_t len) { look_for_marker(b, 0, (u64)len); }, e, fbase);
        if (rc) { close(fd); if (rc < 0) { err = e; return -1; } return rg ? -1 : 1; }
#endif
        if (rg) p0 = 0;                                      // the range starts at a line start: a header, or sequence of the open record
        else if (p0 == ~0ull || first_marker != '>') { close(fd); stream_sync(st); return 1; }
    }
    const u32 kind0 = lead ? 1u : 2u;                        // kind of the line the first tile starts in

    Timer timer(st);
    const u64 ntile = (m + ING_TILE - 1) / ING_TILE;
    const u8 *Rp = R.get();
    DevBuf<u32> bad(1);
    dev_memset(bad.get(), 0, 4, st);
    u32 *badp = bad.get();
    // A: kind of the last line starting in each tile, tagged with the tile number for the running maximum
    DevBuf<u64> tag(ntile), ent(ntile);
    {
        u64 *tg = tag.get();
        launch_for(ntile, st, PFB_LAMBDA(u64 t) {
            u32 k = ingest_walk_tile(Rp, m, p0, t, kind0, badp, [](u64) {}, [](u64, u8) {});
            tg[t] = k ? ((t + 1) << 2) | k : 0ull;
        });
    }
    inclusive_max<u64>(tag.get(), ent.get(), ntile, ws.sc, st);
    // B: counts
    DevBuf<u32> keptc(ntile), hdrc(ntile);
    DevBuf<u64> koff(ntile), hoff(ntile);
    {
        const u64 *en = ent.get();
        u32 *kc = keptc.get(), *hc = hdrc.get();
        launch_for(ntile, st, PFB_LAMBDA(u64 t) {
            u32 kind = t ? (u32)(en[t - 1] & 3) : kind0;
            if (!kind) kind = kind0;                         // nothing started yet: still before the first header
            u32 nk = 0, nh = 0;
            u32 dummy = 0;
            ingest_walk_tile(Rp, m, p0, t, kind, &dummy, [&](u64) { ++nh; }, [&](u64, u8) { ++nk; });
            kc[t] = nk; hc[t] = nh;
        });
    }
    exclusive_sum<u32, u64>(keptc.get(), koff.get(), ntile, ws.sc, st);
    exclusive_sum<u32, u64>(hdrc.get(), hoff.get(), ntile, ws.sc, st);
    const u64 K = read_scalar(koff.get() + (ntile - 1), st) + read_scalar(keptc.get() + (ntile - 1), st);
    const u64 Hh = read_scalar(hoff.get() + (ntile - 1), st) + read_scalar(hdrc.get() + (ntile - 1), st);
    const u64 H = Hh + lead;                                 // records seen (the open one a range starts in included)
    if (read_scalar(bad.get(), st) || H == 0) { if (fd >= 0) close(fd); return decline; }
    if (rg) err.clear();
    // C: write
    out.nrec = closed ? H : H - 1; out.total_len = K; out.len = K + out.nrec * (u64)w;
    out.seg.alloc(out.len ? out.len : 1);
    DevBuf<u64> hpos(store_docs ? H + 1 : 1), dstart(store_docs ? H + 1 : 1);
    {
        const u64 *en = ent.get(), *ko = koff.get(), *ho = hoff.get();
        u8 *o = out.seg.get();
        u64 *hp = hpos.get(), *ds = dstart.get();
        const u64 W = (u64)w;
        const bool docs = store_docs;
        launch_for(ntile, st, PFB_LAMBDA(u64 t) {
            u32 kind = t ? (u32)(en[t - 1] & 3) : kind0;
            if (!kind) kind = kind0;
            u64 k = ko[t], h = ho[t] + lead;
            u32 dummy = 0;
            // kept bytes are gathered in a word and leave as aligned 8-byte stores (bytes at the ragged ends)
            u64 acc = 0, dst = k + W * (h - 1);
            u32 na = 0;
            auto flush_bytes = [&]() {
                for (u32 b = 0; b < na; ++b) o[dst + b] = (u8)(acc >> (8 * b));
                dst += na; na = 0; acc = 0;
            };
            ingest_walk_tile(Rp, m, p0, t, kind, &dummy,
                [&](u64 i) {
                    flush_bytes();
                    ++h;                                     // record h starts; record h-1 (if any) is complete
                    if (h >= 2) { u8 *pad = o + k + W * (h - 2); for (u64 j = 0; j < W; ++j) pad[j] = 'A'; }
                    if (docs) { hp[h - 1 - lead] = i; ds[h - 1 - lead] = k + W * (h - 1); }
                    dst = k + W * (h - 1);
                },
                [&](u64, u8 c) {
                    acc |= (u64)c << (8 * na);
                    ++na; ++k;
                    if (((dst + na) & 7) == 0) {
                        if (na == 8) { *reinterpret_cast<u64 *>(o + dst) = acc; dst += 8; na = 0; acc = 0; }
                        else flush_bytes();
                    }
                });
            flush_bytes();
        });
        u8 *pad = o + K + W * (H - 1);
        if (closed) launch_for(W, st, PFB_LAMBDA(u64 j) { pad[j] = 'A'; });
    }
    if (store_docs && Hh) {
        std::vector<u64> hp(Hh);
        out.starts.resize(Hh);
        d2h(hp.data(), hpos.get(), Hh * 8, st);
        d2h(out.starts.data(), dstart.get(), Hh * 8, st);
        stream_sync(st);
        out.names.resize(Hh);
        char buf[256];
        for (u64 r = 0; r < Hh; ++r) {                       // name = header up to the first white space (kseq.h:188)
            std::string &nm = out.names[r];
            u64 pos = fbase + hp[r] + 1;
            bool done = false;
            if (stream) {
                const u8 *hp_ = out.hostraw.p;
                while (pos < m && !isspace(hp_[pos])) nm.push_back((char)hp_[pos++]);
                continue;
            }
            while (!done && pos < fbase + m) {
                ssize_t got = pread(fd, buf, sizeof(buf), (off_t)pos);
                if (got <= 0) break;
                for (ssize_t i = 0; i < got; ++i) {
                    if (isspace((unsigned char)buf[i])) { done = true; break; }
                    nm.push_back(buf[i]);
                }
                pos += (u64)got;
            }
        }
    }
    out.device_ms = timer.stop();
    if (fd >= 0) close(fd);
    out.hostraw.release();
    return 0;
}

}  // namespace pfb
