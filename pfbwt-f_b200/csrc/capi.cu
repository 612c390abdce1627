// capi.cu -- the C ABI of libpfbwt_b200.so (include/pfbwt_b200.h) over pfb::Engine.
// Compiled by nvcc for sm_100a only; there is no host fallback in this library.
#include "../../include/pfbwt_b200.h"
#include "engine.hpp"
#include <new>

namespace pfb {
size_t &dev_bytes_live() { static size_t v = 0; return v; }
size_t &dev_bytes_peak() { static size_t v = 0; return v; }
SlabPool &slab_pool() { static SlabPool *p = new SlabPool(); return *p; }   // leaked on purpose: outlives every static DevBuf
}
static int g_live_ctx = 0;

struct pfb_ctx {
    pfb::Engine eng;
};

static std::string g_create_err;

static void *pinned_alloc(size_t n) {
    void *p = nullptr;
    if (cudaMallocHost(&p, n) != cudaSuccess) return nullptr;
    return p;
}
static void pinned_free(void *p) { cudaFreeHost(p); }

template <class T>
static int copy_out(pfb_ctx *c, const pfb::DevBuf<T> &b, size_t count, void *out, bool ok, const char *what) {
    if (!ok || b.size() < count) return c->eng.fail(PFB_ERR_ARG, std::string(what) + " not available");
    if (!out) return c->eng.fail(PFB_ERR_ARG, "null output buffer");
    try {
        if (count) { pfb::d2h(out, b.get(), count * sizeof(T), c->eng.ws.st); pfb::stream_sync(c->eng.ws.st); }
    } catch (const pfb::Error &e) { return c->eng.fail(e.code, e.what()); }
    return PFB_OK;
}
// device u32 / u64 arrays leave as bits/8-byte integers
template <class T>
static int copy_out_uint(pfb_ctx *c, const pfb::DevBuf<T> &b, size_t count, void *out, bool ok, const char *what) {
    if (!ok || b.size() < count) return c->eng.fail(PFB_ERR_ARG, std::string(what) + " not available");
    if (!out) return c->eng.fail(PFB_ERR_ARG, "null output buffer");
    const size_t U = (size_t)c->eng.prm.bits / 8;
    try {
        if (U == sizeof(T)) { if (count) { pfb::d2h(out, b.get(), count * sizeof(T), c->eng.ws.st); pfb::stream_sync(c->eng.ws.st); } }
        else if (count) {       // widened / narrowed on the device, then one copy
            cudaStream_t st = c->eng.ws.st;
            const T *s = b.get();
            if (U == 4) {
                pfb::DevBuf<uint32_t> t(count);
                uint32_t *d = t.get();
                pfb::launch_for(count, st, [=] __device__(pfb::u64 i) { d[i] = (uint32_t)s[i]; });
                pfb::d2h(out, t.get(), count * 4, st);
                pfb::stream_sync(st);
            } else {
                pfb::DevBuf<uint64_t> t(count);
                uint64_t *d = t.get();
                pfb::launch_for(count, st, [=] __device__(pfb::u64 i) { d[i] = (uint64_t)s[i]; });
                pfb::d2h(out, t.get(), count * 8, st);
                pfb::stream_sync(st);
            }
        }
    } catch (const pfb::Error &e) { return c->eng.fail(e.code, e.what()); }
    return PFB_OK;
}

extern "C" {

pfb_ctx *pfb_create(int device, int bits) {
    g_create_err.clear();
    if (bits != 32 && bits != 64) { g_create_err = "bits must be 32 or 64"; return nullptr; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_err = std::string("no CUDA device available (") + cudaGetErrorString(e) + "); libpfbwt_b200 has no CPU path";
        return nullptr;
    }
    if (device < 0 || device >= ndev) { g_create_err = "invalid device ordinal"; return nullptr; }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_err = cudaGetErrorString(e); return nullptr; }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    if (prop.major < 10) {
        g_create_err = "libpfbwt_b200 is built for sm_100a (Blackwell B200) only; found sm_" + std::to_string(prop.major) +
                       std::to_string(prop.minor);
        return nullptr;
    }
    pfb_ctx *c = new (std::nothrow) pfb_ctx();
    if (!c) { g_create_err = "out of memory"; return nullptr; }
    c->eng.device = device;
    c->eng.prm.bits = bits;
    if ((e = cudaStreamCreateWithFlags(&c->eng.ws.st, cudaStreamNonBlocking)) != cudaSuccess) {
        g_create_err = cudaGetErrorString(e); delete c; return nullptr;
    }
    ++g_live_ctx;
    pfb::alloc_stream() = c->eng.ws.st;
    c->eng.raw.alloc_fn = pinned_alloc;
    c->eng.raw.free_fn = pinned_free;
    c->eng.raw.pinned = true;
    return c;
}

void pfb_destroy(pfb_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->eng.device);
    cudaStream_t st = c->eng.ws.st;
    cudaStreamSynchronize(st);
    c->eng.clear();
    c->eng.comm.reset();
    c->eng.ws.sc.buf.release();
    c->eng.raw.release();
    cudaStreamSynchronize(st);
    cudaStreamDestroy(st);
    delete c;
    if (--g_live_ctx == 0) pfb::slab_pool().trim();   // last context gone: hand the slabs back
}

const char *pfb_last_error(const pfb_ctx *c) { return c ? c->eng.err.c_str() : g_create_err.c_str(); }
const char *pfb_create_error(void) { return g_create_err.c_str(); }

#define CTX_GUARD(c)                       \
    if (!(c)) return PFB_ERR_ARG;          \
    cudaSetDevice((c)->eng.device);        \
    pfb::alloc_stream() = (c)->eng.ws.st;

int pfb_set_params(pfb_ctx *c, const pfb_parse_params *p) {
    CTX_GUARD(c);
    if (!p) return c->eng.fail(PFB_ERR_ARG, "null params");
    if (p->w > 32) return c->eng.fail(PFB_ERR_ARG, "window size w must be < 32!");   // include/pfparser.hpp:371-376
    if (p->w < 1 || p->w > 31) return c->eng.fail(PFB_ERR_ARG, "window size w must be in 1..31");
    if (p->p < 1) return c->eng.fail(PFB_ERR_ARG, "modulus p must be >= 1");
    if (p->non_acgt_to_a && p->trim_non_acgt)
        return c->eng.fail(PFB_ERR_ARG, "cannot have both --non-acgt-to-a and --trim-non-acgt options enabled at same time");
    c->eng.prm.w = (int)p->w;
    c->eng.prm.p = p->p;
    c->eng.prm.get_sai = p->get_sai != 0;
    c->eng.prm.non_acgt_to_a = p->non_acgt_to_a != 0;
    c->eng.prm.verbose = p->verbose != 0;
    c->eng.store_docs = p->store_docs != 0;
    return PFB_OK;
}

int pfb_add_fasta(pfb_ctx *c, const char *path) { CTX_GUARD(c); return c->eng.add_fasta(path ? path : "-"); }
int pfb_add_records(pfb_ctx *c, const uint8_t *seq, const uint64_t *rec_len, uint64_t nrec) {
    CTX_GUARD(c);
    return c->eng.add_records(seq, rec_len, nrec);
}
int pfb_add_chunk(pfb_ctx *c, const uint8_t *seq, uint64_t len, int ends_record) {
    CTX_GUARD(c);
    if (len && !seq) return c->eng.fail(PFB_ERR_ARG, "null chunk");
    return c->eng.add_chunk(seq, len, ends_record != 0);
}
int pfb_add_fasta_range(pfb_ctx *c, const char *path, uint64_t begin, uint64_t end) {
    CTX_GUARD(c);
    if (!path) return c->eng.fail(PFB_ERR_ARG, "null path");
    return c->eng.add_fasta_range(path, begin, end);
}
int pfb_finalize(pfb_ctx *c) { CTX_GUARD(c); return c->eng.finalize(); }
int pfb_parse_bwt(pfb_ctx *c) { CTX_GUARD(c); return c->eng.do_parse_bwt(); }
int pfb_append(pfb_ctx *l, pfb_ctx *r) {
    CTX_GUARD(l);
    if (!r) return l->eng.fail(PFB_ERR_ARG, "null rhs");
    if (r->eng.device != l->eng.device) return l->eng.fail(PFB_ERR_ARG, "contexts live on different devices");
    cudaStreamSynchronize(r->eng.ws.st);
    return l->eng.append_parse_of(r->eng);
}
int pfb_save_parse(pfb_ctx *c, const char *prefix, int with_parse_bwt, int with_sai) {
    CTX_GUARD(c);
    if (!prefix) return c->eng.fail(PFB_ERR_ARG, "null prefix");
    return c->eng.save_parse(prefix, with_parse_bwt != 0, with_sai != 0, with_parse_bwt == 2);
}
int pfb_save_parse_async(pfb_ctx *c, const char *prefix, int with_parse_bwt, int with_sai) {
    CTX_GUARD(c);
    if (!prefix) return c->eng.fail(PFB_ERR_ARG, "null prefix");
    return c->eng.save_parse_async(prefix, with_parse_bwt, with_sai != 0);
}
int pfb_wait_saves(pfb_ctx *c) { CTX_GUARD(c); return c->eng.wait_saves(); }
int pfb_save_last_sai(pfb_ctx *c, const char *prefix, int with_last, int with_sai) {
    CTX_GUARD(c);
    if (!prefix) return c->eng.fail(PFB_ERR_ARG, "null prefix");
    return c->eng.save_last_sai(prefix, with_last != 0, with_sai != 0);
}
int pfb_load_parse(pfb_ctx *c, const char *prefix, int with_parse_bwt, int with_sai) {
    CTX_GUARD(c);
    return c->eng.load_parse(prefix, with_parse_bwt != 0, with_sai != 0);
}
int pfb_build_bwt(pfb_ctx *c, int sa, int rssa) { CTX_GUARD(c); return c->eng.build(sa != 0, rssa != 0); }
int pfb_build_bwt_to_files(pfb_ctx *c, const char *prefix, int sa, int rssa, uint64_t hbm_limit_bytes) {
    CTX_GUARD(c);
    if (!prefix) return c->eng.fail(PFB_ERR_ARG, "null prefix");
    return c->eng.build_to_files(prefix, sa != 0, rssa != 0, hbm_limit_bytes);
}
uint64_t pfb_build_workspace_bytes(const pfb_ctx *c, int sa, int rssa) {
    if (!c || !c->eng.parsed) return 0;
    return c->eng.build_workspace_estimate(sa || rssa, rssa != 0) + pfb::dev_bytes_live();
}
int pfb_write_doc_array(pfb_ctx *c, const char *prefix) {
    CTX_GUARD(c);
    if (!prefix) return c->eng.fail(PFB_ERR_ARG, "null prefix");
    return c->eng.write_doc_array(prefix);
}
int pfb_write_marker_array(pfb_ctx *c, const char *mps_path, const char *out_path) {
    CTX_GUARD(c);
    if (!mps_path || !out_path) return c->eng.fail(PFB_ERR_ARG, "null path");
    return c->eng.write_marker_array(mps_path, out_path);
}
int pfb_marker_join(pfb_ctx *c, const char *mps_path, const char *sa_path_or_dash, const char *out_path) {
    CTX_GUARD(c);
    if (!mps_path || !sa_path_or_dash || !out_path) return c->eng.fail(PFB_ERR_ARG, "null path");
    return c->eng.marker_join(mps_path, sa_path_or_dash, out_path);
}
int pfb_device_free_bytes(const pfb_ctx *c, size_t *out) {
    if (!c || !out) return PFB_ERR_ARG;
    cudaSetDevice(c->eng.device);
    size_t fr = 0, tot = 0;
    if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) return PFB_ERR_CUDA;
    *out = fr + pfb::slab_pool().idle();
    return PFB_OK;
}
int pfb_write_bwt(pfb_ctx *c, const char *prefix, int sa, int rssa, const char *stdout_ext) {
    CTX_GUARD(c);
    return c->eng.write_bwt_files(prefix, sa != 0, rssa != 0, stdout_ext ? stdout_ext : "");
}

int pfb_get_sizes(const pfb_ctx *c, pfb_sizes *s) {
    if (!c || !s) return PFB_ERR_ARG;
    const pfb::Engine &e = c->eng;
    memset(s, 0, sizeof(*s));
    s->bits = e.prm.bits;
    if (e.parsed) {
        s->n = e.parse.n; s->n_phrases = e.parse.P; s->n_words = e.parse.Dw; s->dict_bytes = e.parse.dsize;
        s->has_parse_bwt = e.parse.has_parse_bwt;
    }
    if (e.built) {
        s->r = e.bwt.r; s->easy = e.bwt.easy; s->hard = e.bwt.hard;
        s->has_bwt = 1; s->has_sa = (e.built_sa || e.built_rssa); s->has_rssa = e.built_rssa;
    } else if (e.streamed) {        // pfb_build_bwt_to_files: the counts are known, the arrays went to the files
        s->r = e.bwt.r; s->easy = e.bwt.easy; s->hard = e.bwt.hard;
    }
    return PFB_OK;
}
uint64_t pfb_get_pending_bases(const pfb_ctx *c) { return c ? c->eng.pending_n : 0; }

int pfb_kat_trigger(pfb_ctx *c, const uint64_t *kmers, size_t n, uint64_t p, uint64_t *hash_out, uint8_t *divides_out) {
    CTX_GUARD(c);
    if (!kmers || !hash_out || !divides_out || p == 0) return c->eng.fail(PFB_ERR_ARG, "bad argument");
    try {
        cudaStream_t st = c->eng.ws.st;
        pfb::DevBuf<pfb::u64> k(n ? n : 1), h(n ? n : 1);
        pfb::DevBuf<pfb::u8> dv(n ? n : 1);
        if (n) pfb::h2d(k.get(), kmers, n * 8, st);
        const pfb::u64 *kp = k.get(); pfb::u64 *hp = h.get(); pfb::u8 *dp = dv.get();
        const pfb::DivTest dt = pfb::DivTest::make(p);
        pfb::launch_for(n, st, [=] __device__(pfb::u64 i) { pfb::u64 x = pfb::wang_hash(kp[i]); hp[i] = x; dp[i] = dt.divides(x) ? 1 : 0; });
        if (n) { pfb::d2h(hash_out, h.get(), n * 8, st); pfb::d2h(divides_out, dv.get(), n, st); }
        pfb::stream_sync(st);
        return PFB_OK;
    } catch (const pfb::Error &e) { return c->eng.fail(e.code, e.what()); }
}
int pfb_get_stats(const pfb_ctx *c, pfb_stats *s) {
    if (!c || !s) return PFB_ERR_ARG;
    const pfb::StageTimes &t = c->eng.tm;
    memset(s, 0, sizeof(*s));
    s->normalise_ms = t.normalise; s->trigger_ms = t.trigger; s->compact_ms = t.compact; s->fingerprint_ms = t.fingerprint;
    s->dedup_ms = t.dedup; s->phrase_sort_ms = t.phrase_sort; s->dict_build_ms = t.dict_build; s->parse_sa_ms = t.parse_sa;
    s->parse_bwt_ms = t.parse_bwt; s->dict_sort_ms = t.dict_sort; s->groups_ms = t.groups; s->emit_ms = t.emit; s->runs_ms = t.runs;
    s->dedup_attempts = t.dedup_attempts; s->phrase_sort_rounds = t.phrase_sort_rounds;
    s->parse_sa_rounds = t.parse_sa_rounds; s->dict_sort_rounds = t.dict_sort_rounds;
    s->launches_own = pfb::launch_stats().own; s->launches_lib = pfb::launch_stats().lib;
    s->dev_bytes_peak = pfb::dev_bytes_peak();
    const_cast<pfb_ctx *>(c)->eng.ws.sc.rw.collect();
    s->radix_pass_ms = pfb::launch_stats().radix_pass_ms; s->radix_passes = pfb::launch_stats().radix_passes;
    s->radix_pairs = pfb::launch_stats().radix_pairs; s->run_ms = c->eng.run_ms;
    s->shard_exchange_ms = c->eng.shard_exchange_ms; s->shard_merge_ms = c->eng.shard_merge_ms;
    s->shard_bytes_moved = c->eng.comm ? c->eng.comm->bytes_moved : 0;
    s->ingest_ms = c->eng.ingest_ms; s->ingest_files = c->eng.ingest_files;
    s->stream_buckets = c->eng.stream_buckets; s->stream_peak_bytes = c->eng.stream_peak;
    return PFB_OK;
}

int pfb_copy_dict(pfb_ctx *c, uint8_t *o) { CTX_GUARD(c); return copy_out(c, c->eng.parse.dict, c->eng.parse.dsize, o, c->eng.parsed, "dict"); }
int pfb_copy_occ(pfb_ctx *c, void *o) { CTX_GUARD(c); return copy_out_uint(c, c->eng.parse.occ, c->eng.parse.Dw, o, c->eng.parsed, "occ"); }
int pfb_copy_parse(pfb_ctx *c, uint32_t *o) { CTX_GUARD(c); return copy_out(c, c->eng.parse.parse, c->eng.parse.P, o, c->eng.parsed, "parse"); }
int pfb_copy_last(pfb_ctx *c, uint8_t *o) { CTX_GUARD(c); return copy_out(c, c->eng.parse.last, c->eng.parse.P, o, c->eng.parsed && c->eng.parse.last.size(), "last"); }
int pfb_copy_sai(pfb_ctx *c, void *o) { CTX_GUARD(c); return copy_out_uint(c, c->eng.parse.sai, c->eng.parse.P, o, c->eng.parsed && c->eng.parse.sai.size(), "sai"); }
int pfb_copy_bwlast(pfb_ctx *c, uint8_t *o) { CTX_GUARD(c); return copy_out(c, c->eng.parse.bwlast, c->eng.parse.P + 1, o, c->eng.parsed && c->eng.parse.has_parse_bwt, "bwlast"); }
int pfb_copy_ilist(pfb_ctx *c, void *o) { CTX_GUARD(c); return copy_out_uint(c, c->eng.parse.ilist, c->eng.parse.P + 1, o, c->eng.parsed && c->eng.parse.has_parse_bwt, "ilist"); }
int pfb_copy_bwsai(pfb_ctx *c, void *o) { CTX_GUARD(c); return copy_out_uint(c, c->eng.parse.bwsai, c->eng.parse.P + 1, o, c->eng.parsed && c->eng.parse.has_parse_bwt && c->eng.parse.bwsai.size(), "bwsai"); }
int pfb_copy_bwt(pfb_ctx *c, uint8_t *o) { CTX_GUARD(c); return copy_out(c, c->eng.bwt.bwt, c->eng.bwt.len, o, c->eng.built, "bwt"); }
int pfb_copy_sa(pfb_ctx *c, void *o) {
    CTX_GUARD(c);
    return copy_out(c, c->eng.bwt.sa, c->eng.bwt.len * ((size_t)c->eng.prm.bits / 8), o, c->eng.built && (c->eng.built_sa || c->eng.built_rssa), "sa");
}
int pfb_copy_ssa(pfb_ctx *c, void *o) { CTX_GUARD(c); return copy_out(c, c->eng.bwt.ssa, 2 * c->eng.bwt.ssa_cnt * ((size_t)c->eng.prm.bits / 8), o, c->eng.built && c->eng.built_rssa, "ssa"); }
int pfb_copy_esa(pfb_ctx *c, void *o) { CTX_GUARD(c); return copy_out(c, c->eng.bwt.esa, 2 * c->eng.bwt.esa_cnt * ((size_t)c->eng.prm.bits / 8), o, c->eng.built && c->eng.built_rssa, "esa"); }

int64_t pfb_get_docs(pfb_ctx *c, uint64_t *starts_out) {
    if (!c) return PFB_ERR_ARG;
    if (starts_out) for (size_t i = 0; i < c->eng.docs.starts.size(); ++i) starts_out[i] = c->eng.docs.starts[i];
    return (int64_t)c->eng.docs.starts.size();
}

int pfb_nccl_unique_id(void *out128) {
    if (!out128) return PFB_ERR_ARG;
    pfb::NcclApi &a = pfb::nccl_api();
    if (!a.load()) { g_create_err = a.err; return PFB_ERR_CUDA; }
    pfb::NcclApi::uid id;
    if (a.GetUniqueId(&id) != 0) { g_create_err = "ncclGetUniqueId failed"; return PFB_ERR_CUDA; }
    memcpy(out128, id.b, 128);
    return PFB_OK;
}
int pfb_shard_init_nccl(pfb_ctx *c, int rank, int world, const void *uid) {
    CTX_GUARD(c);
    if (world < 1 || rank < 0 || rank >= world || !uid) return c->eng.fail(PFB_ERR_ARG, "bad rank / world / id");
    try {
        std::unique_ptr<pfb::NcclComm> nc(new pfb::NcclComm());
        nc->init(rank, world, uid);
        c->eng.comm = std::move(nc);
        c->eng.invalidate();
    } catch (const pfb::Error &e) { return c->eng.fail(e.code, e.what()); }
    return PFB_OK;
}
int pfb_shard_init_host(pfb_ctx *c, int rank, int world, pfb_allgatherv_fn fn, void *user) {
    CTX_GUARD(c);
    if (world < 1 || rank < 0 || rank >= world || !fn) return c->eng.fail(PFB_ERR_ARG, "bad rank / world / callback");
    std::unique_ptr<pfb::HostCallComm> hc(new pfb::HostCallComm());
    hc->rank = rank; hc->world = world; hc->fn = fn; hc->user = user;
    c->eng.comm = std::move(hc);
    c->eng.invalidate();
    return PFB_OK;
}
int pfb_get_shard_info(const pfb_ctx *c, pfb_shard_info *s) {
    if (!c || !s) return PFB_ERR_ARG;
    const pfb::Engine &e = c->eng;
    memset(s, 0, sizeof(*s));
    s->rank = e.comm ? e.comm->rank : 0; s->world = e.comm ? e.comm->world : 1;
    s->text_start = e.shard_start;
    if (e.built) {
        s->bwt_off = e.bwt.off; s->bwt_len = e.bwt.len; s->ssa_off = e.bwt.ssa_off; s->ssa_cnt = e.bwt.ssa_cnt;
        s->esa_off = e.bwt.esa_off; s->esa_cnt = e.bwt.esa_cnt;
    }
    return PFB_OK;
}

int pfb_upload_text(pfb_ctx *c) {
    CTX_GUARD(c);
    if (c->eng.comm) c->eng.text.tail_room = 4096;      // room for the bytes up to the next rank's first trigger
    return c->eng.upload_text();
}
int pfb_run_resident(pfb_ctx *c, int sa, int rssa) {
    CTX_GUARD(c);
    if (!c->eng.text_ready) return c->eng.fail(PFB_ERR_ARG, "no text resident on the device (call pfb_upload_text)");
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, c->eng.ws.st);
    int rc = c->eng.finalize();
    if (!rc) rc = c->eng.do_parse_bwt();
    if (!rc) rc = c->eng.build(sa != 0, rssa != 0);
    cudaEventRecord(e1, c->eng.ws.st);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&c->eng.run_ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return rc;
}
int pfb_release_text(pfb_ctx *c) { CTX_GUARD(c); c->eng.release_text(); return PFB_OK; }
int pfb_clear(pfb_ctx *c) { CTX_GUARD(c); c->eng.clear(); return PFB_OK; }

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// pfb_kat_trigger_static: pfb_kat_trigger for callers that may not allocate (no counterpart in the reference).
// pfb_kat_trigger stages the k-mers, the hashes and the flags in three device buffers
// it allocates per call; here that scratch is ONE file-scope device array of fixed size, PFB_KAT_STATIC_MAX k-mers
// (65536 x 17 bytes = 1088 KB), which every GPU that loads the library holds until exit.
//   * g_kat_scratch names the array for DEVICE code.  In host code the name is only a handle: the device address comes from
//     cudaGetSymbolAddress, once per device with that device current (kat_scratch_address), and that pointer is what the
//     kernel gets.
//   * no cudaMalloc / cudaFree / allocating routine on this path: two copies in, one launch, two copies out on the context's
//     stream.
//   * every call uses the same array and finds the previous call's bytes in it.  Nothing has to start at zero: a call
//     overwrites the first n k-mer slots before the kernel reads them, and the kernel writes the first n hash / flag slots
//     before they are copied out.  Calls that overlap on one GPU (two contexts, two threads) race on the array.
static __device__ __align__(256) unsigned char g_kat_scratch[(size_t)PFB_KAT_STATIC_MAX * 17];
static const size_t KAT_OFF_HASH = (size_t)PFB_KAT_STATIC_MAX * 8, KAT_OFF_FLAG = (size_t)PFB_KAT_STATIC_MAX * 16;

__global__ void __launch_bounds__(256) k_kat_static(const pfb::u64 *__restrict__ kmers, pfb::u64 *__restrict__ hash,
                                                    pfb::u8 *__restrict__ flag, pfb::u64 n, pfb::DivTest dt) {
    for (pfb::u64 i = (pfb::u64)blockIdx.x * 256 + threadIdx.x; i < n; i += (pfb::u64)gridDim.x * 256) {
        const pfb::u64 x = pfb::wang_hash(kmers[i]);
        hash[i] = x;
        flag[i] = dt.divides(x) ? 1 : 0;
    }
}

// device address of g_kat_scratch on the CURRENT device (looked up once per device)
static cudaError_t kat_scratch_address(int device, unsigned char **out) {
    static const int MAXDEV = 64;
    static unsigned char *ptr[MAXDEV];
    static std::mutex mu;
    if (device < 0 || device >= MAXDEV) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        unsigned char *p = nullptr;
        cudaError_t rc = cudaGetSymbolAddress((void **)&p, g_kat_scratch);
        if (rc != cudaSuccess) return rc;
        ptr[device] = p;
    }
    *out = ptr[device];
    return cudaSuccess;
}

extern "C" int pfb_kat_trigger_static(pfb_ctx *c, const uint64_t *kmers, size_t n, uint64_t p, uint64_t *hash_out, uint8_t *divides_out) {
    CTX_GUARD(c);
    if (p == 0 || (n && (!kmers || !hash_out || !divides_out))) return c->eng.fail(PFB_ERR_ARG, "bad argument");
    if (n > (size_t)PFB_KAT_STATIC_MAX)
        return c->eng.fail(PFB_ERR_ARG, "pfb_kat_trigger_static: " + std::to_string(n) + " k-mers exceed the static scratch (PFB_KAT_STATIC_MAX = " +
                                            std::to_string((size_t)PFB_KAT_STATIC_MAX) + "); use pfb_kat_trigger or split the call");
    if (n == 0) return PFB_OK;
    unsigned char *base = nullptr;
    cudaError_t rc = kat_scratch_address(c->eng.device, &base);
    if (rc != cudaSuccess) return c->eng.fail(PFB_ERR_CUDA, std::string("cudaGetSymbolAddress(g_kat_scratch): ") + cudaGetErrorString(rc));
    cudaStream_t st = c->eng.ws.st;
    pfb::u64 *dk = reinterpret_cast<pfb::u64 *>(base), *dh = reinterpret_cast<pfb::u64 *>(base + KAT_OFF_HASH);
    pfb::u8 *df = base + KAT_OFF_FLAG;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if ((rc = cudaMemcpyAsync(dk, kmers, n * 8, cudaMemcpyHostToDevice, st)) == cudaSuccess) {
        k_kat_static<<<blocks, 256, 0, st>>>(dk, dh, df, n, pfb::DivTest::make(p));
        rc = cudaGetLastError();
    }
    if (rc == cudaSuccess) rc = cudaMemcpyAsync(hash_out, dh, n * 8, cudaMemcpyDeviceToHost, st);
    if (rc == cudaSuccess) rc = cudaMemcpyAsync(divides_out, df, n, cudaMemcpyDeviceToHost, st);
    if (rc == cudaSuccess) rc = cudaStreamSynchronize(st);
    if (rc != cudaSuccess) return c->eng.fail(PFB_ERR_CUDA, std::string("pfb_kat_trigger_static: ") + cudaGetErrorString(rc));
    pfb::launch_stats().own++;
    return PFB_OK;
}
