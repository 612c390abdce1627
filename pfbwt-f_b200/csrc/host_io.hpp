// host_io.hpp -- host side of the drop-in boundary: FASTA/FASTQ (plain or gzip,
// file or stdin) reader with kseq semantics (include/kseq.h:178-216 as driven by
// PfParser::add_fasta, include/pfparser.hpp:299-337); used for FASTQ / CR LF inputs and as the
// checker of the device reader (ingest.cuh).  The writers of the reference's on-disk formats
//  stream from HBM: devio.cuh.
#pragma once
#include <zlib.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <unistd.h>
#include <fcntl.h>

namespace pfb {

// growable host byte buffer (pinned when the CUDA build asks for it)
struct HostBytes {
    uint8_t *p = nullptr;
    size_t n = 0, cap = 0;
    bool pinned = false;
    void *(*alloc_fn)(size_t) = nullptr;
    void (*free_fn)(void *) = nullptr;
    ~HostBytes() { release(); }
    void release() {
        if (p) { if (free_fn) free_fn(p); else free(p); }
        p = nullptr; n = cap = 0;
    }
    void reserve(size_t c) {
        if (c <= cap) return;
        size_t nc = cap ? cap : (1u << 20);
        while (nc < c) nc = nc + nc / 2 + 4096;
        uint8_t *q = (uint8_t *)(alloc_fn ? alloc_fn(nc) : malloc(nc));
        if (!q) { fprintf(stderr, "out of host memory\n"); exit(1); }
        if (p) { memcpy(q, p, n); if (free_fn) free_fn(p); else free(p); }
        p = q; cap = nc;
    }
    void append(const uint8_t *s, size_t len) { reserve(n + len); memcpy(p + n, s, len); n += len; }
    void append_fill(uint8_t c, size_t len) { reserve(n + len); memset(p + n, c, len); n += len; }
};

struct FastaResult {
    uint64_t nrec = 0;
    uint64_t total_len = 0;                 // sum of record lengths (no padding)
    std::vector<std::string> names;
    std::vector<uint64_t> starts;           // doc_starts_ (include/pfparser.hpp:321-325)
};

// The reader is a state machine over the byte stream; it appends every record's bytes followed by w filler
// 'A's to `out` (the 'A's are the inter-record padding of include/pfparser.hpp:333-337).  `base_n` is the
// number of text bytes already held by the parser (for doc starts).
struct FastaMachine {
    int w;
    HostBytes &out;
    FastaResult &res;
    bool store_docs;
    uint64_t base_n;
    enum { SEEK_HDR, NAME, COMMENT, LINE_START, SEQ_LINE, PLUS_LINE, QUAL } state = SEEK_HDR;
    uint64_t seq_len = 0, qual_len = 0;
    std::string name;
    bool in_record = false;
    bool hdr_fresh = false;          // a header marker was the last byte seen so far
    bool stopped = false;            // kseq_read returned an error (-2): add_fasta's loop ends, the record and everything behind it are dropped
    size_t rec_out_start = 0;        // out.n when the current record began
    size_t names_before = 0;
    uint8_t qlast = 0;               // last quality character taken (a trailing CR is trimmed line by line, kseq.h:141)
    FastaMachine(int w_, HostBytes &o, FastaResult &r, bool docs, uint64_t base) : w(w_), out(o), res(r), store_docs(docs), base_n(base) {}
    void finish_record() {
        out.append_fill('A', (size_t)w);
        res.total_len += seq_len;
        res.nrec++;
        in_record = false;
    }
    void begin_record() {
        if (store_docs) {
            uint64_t nb = base_n + out.n;      // text bytes before this record (all already padded)
            res.starts.push_back(nb);
        }
        name.clear(); seq_len = 0; qual_len = 0; in_record = true; hdr_fresh = true;
        rec_out_start = out.n; names_before = res.names.size(); qlast = 0;
    }
    void drop_record() {             // kseq.h:209-213: no / short / long quality string -> -2
        out.n = rec_out_start;
        if (store_docs) { if (!res.starts.empty()) res.starts.pop_back(); if (res.names.size() > names_before) res.names.resize(names_before); }
        in_record = false;
        stopped = true;
    }
    void qual_line_done() {
        if (qual_len > 1 && qlast == '\r') { --qual_len; qlast = 0; }
    }
    void feed(const uint8_t *b, size_t e) {
        size_t i = 0;
        while (i < e && !stopped) {
            switch (state) {
            case SEEK_HDR: {   // kseq.h:183-187: skip to the next '>' or '@' anywhere
                while (i < e && b[i] != '>' && b[i] != '@') ++i;
                if (i < e) { ++i; begin_record(); state = NAME; }
                break;
            }
            case NAME: {       // name = up to first whitespace
                hdr_fresh = false;
                while (i < e && !isspace(b[i])) { if (store_docs) name.push_back((char)b[i]); ++i; }
                if (i < e) {
                    if (store_docs) res.names.push_back(name);
                    state = b[i] == '\n' ? LINE_START : COMMENT;
                    ++i;
                }
                break;
            }
            case COMMENT: {
                const uint8_t *q = (const uint8_t *)memchr(b + i, '\n', e - i);
                if (q) { i = (size_t)(q - b) + 1; state = LINE_START; } else i = e;
                break;
            }
            case LINE_START: { // kseq.h:194-198
                uint8_t c = b[i];
                if (c == '>' || c == '@') { finish_record(); ++i; begin_record(); state = NAME; }
                else if (c == '+') { ++i; state = PLUS_LINE; }
                else if (c == '\n') { ++i; }
                else state = SEQ_LINE;
                break;
            }
            case SEQ_LINE: {
                const uint8_t *q = (const uint8_t *)memchr(b + i, '\n', e - i);
                size_t le = q ? (size_t)(q - b) : e;
                out.append(b + i, le - i);
                seq_len += le - i;
                i = le;
                if (q) {
                    // kseq.h:139: a trailing '\r' is dropped when the string is longer than 1
                    if (seq_len > 1 && out.p[out.n - 1] == '\r') { out.n--; seq_len--; }
                    ++i; state = LINE_START;
                }
                break;
            }
            case PLUS_LINE: {  // kseq.h:207
                const uint8_t *q = (const uint8_t *)memchr(b + i, '\n', e - i);
                if (q) { i = (size_t)(q - b) + 1; state = seq_len ? QUAL : SEEK_HDR; if (!seq_len) finish_record(); } else i = e;
                break;
            }
            case QUAL: {       // kseq.h:209-213: whole quality lines until at least as long as the sequence, then the lengths must agree
                const uint8_t *q = (const uint8_t *)memchr(b + i, '\n', e - i);
                const size_t le = q ? (size_t)(q - b) : e;
                if (le > i) { qual_len += le - i; qlast = b[le - 1]; }
                i = le;
                if (q) {
                    ++i;
                    qual_line_done();
                    if (qual_len >= seq_len) {
                        if (qual_len == seq_len) { finish_record(); state = SEEK_HDR; }
                        else drop_record();
                    }
                }
                break;
            }
            }
        }
    }
    void finish() {
        if (in_record && hdr_fresh) {          // kseq.h:188: a header marker as the very last byte gives no record (the name read hits EOF)
            if (store_docs && !res.starts.empty()) res.starts.pop_back();
            in_record = false;
            return;
        }
        if (stopped) return;
        if (in_record && state == PLUS_LINE) { drop_record(); return; }            // EOF inside the '+' line: no quality string
        if (in_record && state == QUAL) {
            qual_line_done();
            if (qual_len == seq_len) finish_record(); else drop_record();
            return;
        }
        if (state == NAME && store_docs && in_record) res.names.push_back(name);
        if (in_record) {
            if (seq_len > 1 && state == SEQ_LINE && out.p[out.n - 1] == '\r') { out.n--; seq_len--; }
            finish_record();
        }
    }
};

// plain or gzip, file or stdin ("-").  Returns 0, or -1 with err set.
inline int read_fasta(const std::string &path, int w, HostBytes &out, FastaResult &res, bool store_docs, uint64_t base_n,
                      std::string &err) {
    gzFile fp = path == "-" ? gzdopen(fileno(stdin), "r") : gzopen(path.c_str(), "r");
    if (fp == NULL) { err = "failed to open file!\n"; return -1; }
    gzbuffer(fp, 1 << 20);
    size_t BUFSZ = 8u << 20;
    if (const char *e = getenv("PFB_READ_CHUNK")) { long v = atol(e); if (v > 0) BUFSZ = (size_t)v; }   // test knob: states across chunk ends
    std::vector<uint8_t> buf(BUFSZ);
    FastaMachine M(w, out, res, store_docs, base_n);
    for (;;) {
        int got = gzread(fp, buf.data(), (unsigned)BUFSZ);
        if (got < 0) { err = "error reading input"; gzclose(fp); return -1; }
        if (got == 0) break;
        M.feed(buf.data(), (size_t)got);
    }
    M.finish();
    gzclose(fp);
    return 0;
}

// the same over bytes already in memory (a stream the device reader took in and then handed back)
inline void read_fasta_mem(const uint8_t *data, size_t len, int w, HostBytes &out, FastaResult &res, bool store_docs, uint64_t base_n) {
    FastaMachine M(w, out, res, store_docs, base_n);
    M.feed(data, len);
    M.finish();
}

// ------------------------------------------------------------ small files --
inline int read_file(const std::string &path, std::vector<uint8_t> &out, std::string &err) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { err = "error opening " + path; return -1; }
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize((size_t)sz);
    if (sz && fread(out.data(), 1, (size_t)sz, f) != (size_t)sz) { fclose(f); err = "error reading " + path; return -1; }
    fclose(f);
    return 0;
}
}  // namespace pfb
