/* pfp_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C CPU restatement of the reference hot path of alshai/pfbwt-f
 * (fasta -> prefix-free parse -> parse BWT -> dictionary suffix sort -> BWT/SA),
 * written from the algorithm's definition, each function citing the reference
 * file:line it follows (paths relative to /root/reference).
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement against
 *  (a) the reference's own golden vectors tests/data/{single_chrom,mult_chroms}.{bwt,sa}
 *      (committed under tests/golden/), and
 *  (b) byte-for-byte against the unmodified reference binaries built by
 *      oracle/Makefile into oracle/_ref/ for every intermediate and final file.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference
 * legs may call into this; libpfbwt_b200.so never does.
 */
#ifndef PFP_ORACLE_H
#define PFP_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PFO_DOLLAR 2       /* include/utils.h:8  */
#define PFO_ENDOFWORD 1    /* include/utils.h:9  */
#define PFO_ENDOFDICT 0    /* include/utils.h:10 */

typedef struct {
    int w, p;
    int non_acgt_to_a;
    /* normalised text S (upper-cased records, each followed by A^w) */
    uint8_t *S;      uint64_t n;
    /* parse */
    uint64_t P;      /* number of phrases                         */
    uint64_t Dw;     /* number of distinct phrases                */
    uint8_t *dict;   uint64_t dsize;   /* .dict image             */
    uint64_t *wstart;                  /* Dw+1 word starts in dict */
    uint64_t *occ;   /* Dw                                        */
    uint32_t *parse; /* P, 1-based ranks                          */
    uint8_t *last;   /* P                                         */
    uint64_t *sai;   /* P                                         */
    /* parse BWT (P+1 each) */
    uint8_t *bwlast; uint64_t *ilist; uint64_t *bwsai;
    /* final (n+1 each) */
    uint8_t *bwt;    uint64_t *sa;
    uint64_t r;      /* number of BWT runs */
    uint64_t easy, hard;
    /* run-length SA samples: pairs (i, SA[i]) */
    uint64_t *ssa; uint64_t nssa;  /* 2*nssa values */
    uint64_t *esa; uint64_t nesa;
    char err[256];
} pfo_t;

uint64_t pfo_wang_hash(uint64_t key);

/* records: concatenated raw sequence bytes; rec_len[i] lengths. Builds S. Returns 0 or -1 (err set). */
int pfo_set_records(pfo_t *o, const uint8_t *seq, const uint64_t *rec_len, uint64_t nrec);
/* read a plain (uncompressed) FASTA/FASTQ file with kseq semantics */
int pfo_read_fasta(pfo_t *o, const char *path);

int pfo_parse(pfo_t *o);          /* phrases, dict, occ, parse, last, sai        */
int pfo_parse_bwt(pfo_t *o);      /* bwlast, ilist, bwsai                        */
int pfo_build_bwt(pfo_t *o, int want_sa);  /* bwt (+sa), r, easy/hard, ssa/esa   */
/* independent check: SA/BWT of S.$ by plain prefix doubling (no PFP) */
int pfo_naive_sa(const uint8_t *S, uint64_t n, uint64_t *sa_out /* n+1 */);

/* restatement of PfParser::operator+= over a list of (dict, parse) pairs:
 * out gets P, Dw, dict, occ, parse, last, sai, n of the concatenation */
int pfo_merge(pfo_t *out, pfo_t *const *parts, int nparts);
/* load .dict/.parse written by anyone into a pfo_t (dict, wstart, parse, P, Dw) */
int pfo_load_parse(pfo_t *o, const char *prefix);

/* write the reference's file set; U = 4 or 8 */
int pfo_write_parse_files(const pfo_t *o, const char *prefix, int U, int with_sai);
int pfo_write_bwt_files(const pfo_t *o, const char *prefix, int U, int sa, int rssa);

/* mps_to_ma (src/mps_to_ma.cpp, include/marker_array.hpp:137-174): 0 ok, 1 = the reference's own refusal, < 0 bad input */
int pfo_marker_join(const char *mps_path, const char *sa_path, const char *out_path);
void pfo_init(pfo_t *o, int w, int p, int non_acgt_to_a);
void pfo_free(pfo_t *o);

#ifdef __cplusplus
}
#endif
#endif
