/* pfbwt_b200.h -- C ABI of libpfbwt_b200.so, the B200-native drop-in for the
 * fasta -> prefix-free parse -> BWT(+SA) hot path of alshai/pfbwt-f.
 *
 * The reference has no C ABI of its own (only header-level C++ classes); every
 * entry point below names the reference interface it replaces (paths relative
 * to the reference tree).  Plain pointers and sizes only; all buffers passed in
 * or out are HOST memory owned by the caller.  Every call is synchronous; a
 * context is bound to one CUDA device and is not thread-safe (like the
 * reference's classes).  There is no CPU fallback: pfb_create() fails when no
 * CUDA device is usable.
 *
 * Return value of int functions: 0 on success, a negative PFB_ERR_* code on
 * failure; pfb_last_error() then returns the message the reference would have
 * printed before exit(1) (same wording where the reference has one).
 */
#ifndef PFBWT_B200_H
#define PFBWT_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PFB_OK 0
#define PFB_ERR_ARG (-1)
#define PFB_ERR_INVALID_CHAR (-2)  /* include/hash.hpp:31 */
#define PFB_ERR_LIMIT (-3)         /* include/pfparser.hpp:393-405 */
#define PFB_ERR_NOMEM (-4)
#define PFB_ERR_CUDA (-5)
#define PFB_ERR_IO (-6)
#define PFB_ERR_ONE_WORD (-7)      /* include/pfparser.hpp:390-392 */
#define PFB_ERR_INTERNAL (-8)
#define PFB_ERR_TOO_LONG_32 (-9)   /* include/pfparser.hpp:326-332 */

typedef struct pfb_ctx pfb_ctx;

/* pfbwtf::PfParserParams, include/pfparser.hpp:30-44 (same field meaning) */
typedef struct {
    uint32_t w;             /* window size, default 10, 1..31            */
    uint64_t p;             /* trigger modulus, default 100              */
    int get_sai;            /* keep sai_/bwsai (-s / -r)                 */
    int store_docs;         /* --print-docs                              */
    int verbose;
    int trim_non_acgt;      /* accepted, disabled like the reference (include/pfparser.hpp:338-341) */
    int non_acgt_to_a;      /* --non-acgt-to-a                           */
} pfb_parse_params;

/* sizes of everything a context currently holds (0 when absent) */
typedef struct {
    uint64_t n;             /* PfParser::get_n(), include/pfparser.hpp:529-532  */
    uint64_t n_phrases;     /* get_parse_size()                                  */
    uint64_t n_words;       /* distinct phrases                                  */
    uint64_t dict_bytes;    /* size of the .dict image                           */
    uint64_t r;             /* BWT runs (src/pfbwt-f.cpp:346)                    */
    uint64_t easy, hard;    /* include/pfbwt.hpp:188                             */
    int bits;               /* 32 | 64: width of uint_t in the replaced binary   */
    int has_parse_bwt, has_bwt, has_sa, has_rssa;
} pfb_sizes;

/* per-stage device times (CUDA events, ms) and counters of the last run */
typedef struct {
    float normalise_ms, trigger_ms, compact_ms, fingerprint_ms, dedup_ms, phrase_sort_ms, dict_build_ms;
    float parse_sa_ms, parse_bwt_ms, dict_sort_ms, groups_ms, emit_ms, runs_ms;
    uint32_t dedup_attempts, phrase_sort_rounds, parse_sa_rounds, dict_sort_rounds;
    uint64_t launches_own;  /* kernels written in this repository launched so far in this process */
    uint64_t launches_lib;  /* library (CUB) kernels launched so far                              */
    uint64_t dev_bytes_peak;
    /* k_radix_pass launches over >= 2^24 pairs so far in this process: summed CUDA-event time, count, pairs moved */
    double radix_pass_ms; uint64_t radix_passes, radix_pairs;
    float run_ms;           /* device time of the last pfb_run_resident (CUDA events on the pipeline stream) */
    /* multi-GPU job: time spent in exchanges (incl. waiting for peers) / in the dictionary merge, bytes received */
    float shard_exchange_ms, shard_merge_ms;
    uint64_t shard_bytes_moved;
    /* FASTA files whose records were split on the device (pfb_add_fasta on plain files) and the device time of that */
    float ingest_ms; uint64_t ingest_files;
    /* last pfb_build_bwt_to_files: key-range buckets the build was cut into, peak device bytes while it ran */
    uint64_t stream_buckets, stream_peak_bytes;
} pfb_stats;

/* which part of the job's outputs this context holds.  Without a communicator: everything. */
typedef struct {
    int rank, world;
    uint64_t text_start;            /* offset of this rank's records in S                     */
    uint64_t bwt_off, bwt_len;      /* BWT (and SA) positions [bwt_off, bwt_off + bwt_len)    */
    uint64_t ssa_off, ssa_cnt;      /* .ssa pairs                                             */
    uint64_t esa_off, esa_cnt;      /* .esa pairs                                             */
} pfb_shard_info;

/* ---- lifetime ------------------------------------------------------------ */
/* bits = 32 replaces pfbwt-f, 64 replaces pfbwt-f64 / merge_pfp (gsa/gsacak.h:40-58). */
pfb_ctx *pfb_create(int device, int bits);
void pfb_destroy(pfb_ctx *);
const char *pfb_last_error(const pfb_ctx *);
/* message for a failed pfb_create (no context to ask) */
const char *pfb_create_error(void);

/* PfParser(PfParserParams), include/pfparser.hpp:80-84 (incl. check_w :371-376) */
int pfb_set_params(pfb_ctx *, const pfb_parse_params *);

/* ---- parse step (pfbwtf::PfParser) ------------------------------------------ */
/* PfParser::add_fasta, include/pfparser.hpp:299-369.  path "-" = stdin; plain or gzip. */
int pfb_add_fasta(pfb_ctx *, const char *path_or_dash);
/* same, from host buffers: nrec records, concatenated bytes in seq, lengths in rec_len */
int pfb_add_records(pfb_ctx *, const uint8_t *seq, const uint64_t *rec_len, uint64_t nrec);
/* a piece of ONE record (mid-record cut): `len` bytes continuing the record the previous chunk left open, or starting a new
 * one; ends_record != 0 closes it (the w filler 'A's of include/pfparser.hpp:333-337 follow).  Chunks of a record added to
 * one context concatenate; in a multi-GPU job rank g may end with an open record that rank g+1 continues - the
 * w-overlapping chunk split that generalises the junction re-scan of PfParser::operator+= (include/pfparser.hpp:226-250). */
int pfb_add_chunk(pfb_ctx *, const uint8_t *seq, uint64_t len, int ends_record);
/* bytes [begin, end) of a plain FASTA file (one rank's share of a file split over the ranks of a multi-GPU job).  begin and
 * end are moved forward to the next line start; sequence bytes before the range's first header continue the record of the
 * previous range, and the last record stays open unless the next range starts with a header (or end reaches the file end). */
int pfb_add_fasta_range(pfb_ctx *, const char *path, uint64_t begin, uint64_t end);
/* PfParser::finalize, include/pfparser.hpp:484-492: runs the parse on the device */
int pfb_finalize(pfb_ctx *);
/* PfParser::bwt_of_parse, include/pfparser.hpp:379-467 */
int pfb_parse_bwt(pfb_ctx *);
/* PfParser::operator+=, include/pfparser.hpp:194-263 (call pfb_finalize(lhs) afterwards) */
int pfb_append(pfb_ctx *lhs, pfb_ctx *rhs);
/* save_parser + the bwt_of_parse writer: include/pfbwt_io.hpp:234-249, src/pfbwt-f.cpp:229-243.
 * with_parse_bwt: 0 = .dict .occ .parse .n (.docs); 1 = those + .bwlast .ilist (.bwsai with with_sai);
 * 2 = only .bwlast .ilist (.bwsai) - run_parser's third task, after the parse files were saved in its second */
int pfb_save_parse(pfb_ctx *, const char *prefix, int with_parse_bwt, int with_sai);
/* pfb_save_parse in the background: the files are written by a writer thread (own pinned staging, own copy streams) while
 * the caller goes on with pfb_parse_bwt / pfb_build_bwt / pfb_write_bwt on the same context.  pfb_wait_saves joins the
 * writers and returns their first error; everything that replaces the parse waits by itself. */
int pfb_save_parse_async(pfb_ctx *, const char *prefix, int with_parse_bwt, int with_sai);
int pfb_wait_saves(pfb_ctx *);
/* <prefix>.last (n_phrases bytes, PfParser::last_) and <prefix>.sai (n_phrases x bits/8, PfParser::sai_): the extensions
 * include/utils.h:22,24 defines; the reference never writes them (only tests/test_parser.cpp:126,139 reads them), so they
 * are written on request only */
int pfb_save_last_sai(pfb_ctx *, const char *prefix, int with_last, int with_sai);
/* load_parser (include/pfbwt_io.hpp:211-222) / PrefixFreeBWT ctor (include/pfbwt.hpp:64-81) */
int pfb_load_parse(pfb_ctx *, const char *prefix, int with_parse_bwt, int with_sai);

/* ---- BWT build (pfbwtf::PrefixFreeBWT) ----------------------------------------- */
/* PrefixFreeBWT::generate_bwt_lcp driven by run_pfbwt, include/pfbwt.hpp:96-194,
 * src/pfbwt-f.cpp:272-349.  sa / rssa = -s / -r. */
int pfb_build_bwt(pfb_ctx *, int sa, int rssa);
/* The -m mode (workspace larger than memory: include/file_wrappers.hpp:35-124 mmap sinks, include/pfbwt.hpp:208-209,
 * selected at src/pfbwt-f.cpp:359-365).  Build and write in one call: when the work arrays of a full build would not fit in
 * hbm_limit_bytes (0 = what the device has free; the environment variable PFB_HBM_LIMIT=48G overrides), the dictionary
 * suffixes are sorted and emitted one key range after the other and every finished slice of <prefix>.bwt/.sa/.ssa/.esa is
 * streamed to its offset of the file.  Afterwards pfb_get_sizes reports n, r, easy, hard; the arrays are not kept. */
int pfb_build_bwt_to_files(pfb_ctx *, const char *prefix, int sa, int rssa, uint64_t hbm_limit_bytes);
/* Document array (README.md:29-33 "Document Array and other SA-related data structures" - announced, not present in the
 * reference, so the format is ours): <prefix>.da = one uint32 per BWT position i, the index of the record (line of
 * <prefix>.docs, include/pfbwt_io.hpp:224-231) holding text position SA[i].  Needs store_docs and pfb_build_bwt with sa/rssa. */
int pfb_write_doc_array(pfb_ctx *, const char *prefix);
/* The marker-array join of src/mps_to_ma.cpp (write_marker_array, include/marker_array.hpp:137-174, over the rle_window_arr of
 * include/rle_window_array.hpp:15-51,118-130): <mps> holds runs of text positions with their variant markers
 * ([first, last, marker..., 2^64-1] as u64 words), the output holds, for every maximal run of BWT positions whose SA[i]
 * carry the same non-empty marker list, [first i, last i, marker..., 2^64-1].
 * pfb_write_marker_array joins with the SA pfb_build_bwt (sa or rssa) left on the device - no .sa stream; after
 * pfb_build_bwt_to_files (sa = 1) it joins with the <prefix>.sa file that build wrote (32- or 64-bit entries);
 * pfb_marker_join reads the SA as u64 entries from a file or from stdin ("-"), like the reference's tool (any context will do). */
int pfb_write_marker_array(pfb_ctx *, const char *mps_path, const char *out_path);
int pfb_marker_join(pfb_ctx *, const char *mps_path, const char *sa_path_or_dash, const char *out_path);
/* device bytes a full in-memory pfb_build_bwt needs at its peak (estimate), incl. what the context already holds */
uint64_t pfb_build_workspace_bytes(const pfb_ctx *, int sa, int rssa);
/* device bytes available to this context right now (free on the device + held idle by the library's allocator) */
int pfb_device_free_bytes(const pfb_ctx *, size_t *out);
/* the .bwt/.sa/.ssa/.esa writer of src/pfbwt-f.cpp:298-328; stdout_ext = "bwt" | "sa" | NULL (--stdout) */
int pfb_write_bwt(pfb_ctx *, const char *prefix, int sa, int rssa, const char *stdout_ext);

/* ---- getters (PfParser getters include/pfparser.hpp:469-544; copies into caller buffers) ---- */
int pfb_get_sizes(const pfb_ctx *, pfb_sizes *);
/* bases (incl. the w-byte padding after each record) added since the last pfb_clear: PfParser::get_pos() - 1 before finalize */
uint64_t pfb_get_pending_bases(const pfb_ctx *);

/* Known-answer hook for the trigger test of the parse (include/hash.hpp:12-21 wang_hash, include/pfparser.hpp:347
 * `hash % p == 0`): the DEVICE functions the trigger kernel uses are run on n caller-supplied k-mer values;
 * hash_out[i] = wang_hash(kmers[i]), divides_out[i] = 1 iff the kernel's division-free test says p divides it. */
int pfb_kat_trigger(pfb_ctx *, const uint64_t *kmers, size_t n, uint64_t p, uint64_t *hash_out, uint8_t *divides_out);
/* pfb_kat_trigger without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments and results as pfb_kat_trigger; the device scratch pfb_kat_trigger allocates per call (k-mers, hashes, flags)
 * is a static device array of PFB_KAT_STATIC_MAX k-mers (1088 KB held by every GPU that loads the library).  n above
 * PFB_KAT_STATIC_MAX fails with PFB_ERR_ARG.  All calls share the array: calls that overlap on one GPU race. */
#define PFB_KAT_STATIC_MAX 65536
int pfb_kat_trigger_static(pfb_ctx *, const uint64_t *kmers, size_t n, uint64_t p, uint64_t *hash_out, uint8_t *divides_out);
int pfb_get_stats(const pfb_ctx *, pfb_stats *);
int pfb_copy_dict(pfb_ctx *, uint8_t *out /* dict_bytes */);
int pfb_copy_occ(pfb_ctx *, void *out /* n_words * bits/8 */);
int pfb_copy_parse(pfb_ctx *, uint32_t *out /* n_phrases */);
int pfb_copy_last(pfb_ctx *, uint8_t *out /* n_phrases */);
int pfb_copy_sai(pfb_ctx *, void *out /* n_phrases * bits/8 */);
int pfb_copy_bwlast(pfb_ctx *, uint8_t *out /* n_phrases+1 */);
int pfb_copy_ilist(pfb_ctx *, void *out /* (n_phrases+1) * bits/8 */);
int pfb_copy_bwsai(pfb_ctx *, void *out /* (n_phrases+1) * bits/8 */);
int pfb_copy_bwt(pfb_ctx *, uint8_t *out /* n+1; a rank of a multi-GPU job: bwt_len */);
int pfb_copy_sa(pfb_ctx *, void *out /* (n+1) * bits/8 */);
int pfb_copy_ssa(pfb_ctx *, void *out /* 2*r * bits/8 */);
int pfb_copy_esa(pfb_ctx *, void *out /* 2*r * bits/8 */);
/* doc table of --print-docs (include/pfbwt_io.hpp:224-231): number of docs; start offsets copied when out != NULL */
int64_t pfb_get_docs(pfb_ctx *, uint64_t *starts_out);

/* ---- one job over several GPUs: one process (context) per GPU --------------------------------
 * Generalises merge_pfp (src/merge_pfp.cpp:97-170: per-file parsers merged with operator+=,
 * include/pfparser.hpp:194-263) to a single job: rank r of `world` adds the records of the r-th
 * contiguous piece of the text (rank order = text order; cuts at record boundaries with pfb_add_fasta /
 * pfb_add_records, anywhere - also inside a record - with pfb_add_chunk / pfb_add_fasta_range), then
 * every rank calls pfb_finalize / pfb_parse_bwt / pfb_build_bwt / pfb_write_bwt at the same time
 * (collective calls).  After pfb_finalize every rank holds the whole parse (getters as before);
 * after pfb_build_bwt a rank holds one contiguous slice of .bwt/.sa/.ssa/.esa (pfb_get_shard_info,
 * the pfb_copy_* getters copy the slice, pfb_write_bwt writes it at its offset of the common file).
 * Transport: NCCL over NVLink/NVSwitch (the id comes from rank 0 via the launcher), or a
 * caller-supplied host all-gather (tests: torch.distributed/gloo). */
typedef int (*pfb_allgatherv_fn)(void *user, const void *send, void *recv, const uint64_t *counts, int world);
int pfb_nccl_unique_id(void *out128);
int pfb_shard_init_nccl(pfb_ctx *, int rank, int world, const void *unique_id128);
int pfb_shard_init_host(pfb_ctx *, int rank, int world, pfb_allgatherv_fn fn, void *user);
int pfb_get_shard_info(const pfb_ctx *, pfb_shard_info *);

/* ---- measurement hooks (bench.py) -------------------------------------------------- */
/* stage the pending records into HBM (H2D + normalise) without parsing */
int pfb_upload_text(pfb_ctx *);
/* whole hot path on text already resident in HBM: parse -> parse BWT -> BWT(+SA); no host copies */
int pfb_run_resident(pfb_ctx *, int sa, int rssa);
/* release the device text and the staged input (records in HBM, host staging) once the parse exists; parsing again
 * needs pfb_clear and the input once more */
int pfb_release_text(pfb_ctx *);
/* forget all records, parses and outputs (keeps parameters and the CUDA stream) */
int pfb_clear(pfb_ctx *);

#ifdef __cplusplus
}
#endif
#endif
