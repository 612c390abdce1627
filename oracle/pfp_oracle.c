/* pfp_oracle.c -- TEST INFRASTRUCTURE ONLY; see pfp_oracle.h.
 *
 * CPU restatement of the pfbwt-f hot path in plain C.  Written for clarity,
 * not speed: sorting is qsort with direct string comparison, suffix arrays use
 * naive prefix doubling.  Sized for inputs up to a few tens of Mbp.
 * Every function cites the reference location it restates (paths relative to
 * /root/reference).  Parity: PINNED (goldens + oracle/_ref binaries), see
 * tests/test_oracle.py.
 */
#define _GNU_SOURCE
#include "pfp_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ctype.h>

/* ---- tables: src/utils.c:139-180 (restated as functions) ---------------- */
static int nt4_ntoa(int c) { /* seq_nt4_ntoa_table, src/utils.c:139-161 */
    switch (c) {
    case 'A': case 'a': case 'N': case 'n': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case '-': return 3;
    default: return 5;
    }
}
static int nt4(int c) { /* seq_nt4_table, src/utils.c:163-180 */
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 4;
    }
}

/* include/hash.hpp:12-21 */
uint64_t pfo_wang_hash(uint64_t key) {
    key = (~key) + (key << 21);
    key = key ^ (key >> 24);
    key = (key + (key << 3)) + (key << 8);
    key = key ^ (key >> 14);
    key = (key + (key << 2)) + (key << 4);
    key = key ^ (key >> 28);
    key = key + (key << 31);
    return key;
}

void pfo_init(pfo_t *o, int w, int p, int non_acgt_to_a) {
    memset(o, 0, sizeof(*o));
    o->w = w; o->p = p; o->non_acgt_to_a = non_acgt_to_a;
}
void pfo_free(pfo_t *o) {
    free(o->S); free(o->dict); free(o->wstart); free(o->occ); free(o->parse);
    free(o->last); free(o->sai); free(o->bwlast); free(o->ilist); free(o->bwsai);
    free(o->bwt); free(o->sa); free(o->ssa); free(o->esa);
    int w = o->w, p = o->p, f = o->non_acgt_to_a;
    pfo_init(o, w, p, f);
}

/* include/pfparser.hpp:333-346 (char pipeline) + include/hash.hpp:29-31 (validation) */
int pfo_set_records(pfo_t *o, const uint8_t *seq, const uint64_t *rec_len, uint64_t nrec) {
    uint64_t tot = 0;
    for (uint64_t r = 0; r < nrec; ++r) tot += rec_len[r] + (uint64_t)o->w;
    free(o->S);
    o->S = (uint8_t *)malloc(tot + 1);
    uint64_t k = 0, src = 0;
    for (uint64_t r = 0; r < nrec; ++r) {
        for (uint64_t i = 0; i < rec_len[r] + (uint64_t)o->w; ++i) {
            int c = i < rec_len[r] ? toupper(seq[src + i]) : 'A';
            if (o->non_acgt_to_a && nt4(c) > 3) c = 'A';
            if (c >= 128 || nt4_ntoa(c) > 3) {
                snprintf(o->err, sizeof(o->err), "error, invalid character %d/%c -> %d", c, c, 5);
                return -1;
            }
            o->S[k++] = (uint8_t)c;
        }
        src += rec_len[r];
    }
    o->n = k;
    return 0;
}

/* include/kseq.h:178-216 (kseq_read) restated for plain files: header lines
 * start with '>' or '@'; sequence lines are concatenated; a line starting with
 * '+' begins a FASTQ quality block of the same length as the sequence. */
int pfo_read_fasta(pfo_t *o, const char *path) {
    FILE *fp = fopen(path, "rb");
    if (!fp) { snprintf(o->err, sizeof(o->err), "failed to open file!"); return -1; }
    fseek(fp, 0, SEEK_END); long fsz = ftell(fp); fseek(fp, 0, SEEK_SET);
    uint8_t *buf = (uint8_t *)malloc((size_t)fsz + 1);
    if (fread(buf, 1, (size_t)fsz, fp) != (size_t)fsz) { fclose(fp); free(buf); return -1; }
    fclose(fp);
    uint8_t *seq = (uint8_t *)malloc((size_t)fsz + 1);
    uint64_t *lens = NULL; uint64_t nrec = 0, cap = 0, k = 0;
    long i = 0;
    /* jump to first header */
    while (i < fsz && buf[i] != '>' && buf[i] != '@') ++i;
    while (i < fsz) {
        /* header line */
        while (i < fsz && buf[i] != '\n') ++i;
        if (i < fsz) ++i;
        uint64_t l = 0;
        int c = -1;
        while (i < fsz) {
            c = buf[i];
            if (c == '>' || c == '+' || c == '@') break;
            if (c == '\n') { ++i; continue; }
            long j = i;
            while (j < fsz && buf[j] != '\n') ++j;
            long e = j;
            if (e - i > 0 && l + (uint64_t)(e - i) > 1 && buf[e - 1] == '\r') --e; /* kseq.h:139 */
            memcpy(seq + k + l, buf + i, (size_t)(e - i));
            l += (uint64_t)(e - i);
            i = j < fsz ? j + 1 : j;
            c = -1;
        }
        if (nrec == cap) { cap = cap ? cap * 2 : 16; lens = (uint64_t *)realloc(lens, cap * sizeof(uint64_t)); }
        lens[nrec++] = l; k += l;
        if (i < fsz && buf[i] == '+') { /* FASTQ: skip '+' line and l quality chars */
            while (i < fsz && buf[i] != '\n') ++i;
            if (i < fsz) ++i;
            uint64_t q = 0;
            while (i < fsz && q < l) { if (buf[i] != '\n' && buf[i] != '\r') ++q; ++i; }
            while (i < fsz && buf[i] != '>' && buf[i] != '@') ++i;
        }
    }
    int rc = pfo_set_records(o, seq, lens, nrec);
    free(buf); free(seq); free(lens);
    return rc;
}

/* ---- phrase list helper -------------------------------------------------- */
typedef struct { const uint8_t *s; uint64_t len; uint64_t id; } phr_t;

static int phr_cmp(const void *a, const void *b) { /* strcmp order, include/pfparser.hpp:500-501 */
    const phr_t *x = (const phr_t *)a, *y = (const phr_t *)b;
    uint64_t m = x->len < y->len ? x->len : y->len;
    int c = memcmp(x->s, y->s, m);
    if (c) return c;
    if (x->len != y->len) return x->len < y->len ? -1 : 1;
    return 0;
}

/* Given P phrases in text order (ph[k].s/len) fill dict, wstart, occ, parse.
 * Restates process_phrase's dedup (include/pfparser.hpp:595-598), sort_dict
 * (:494-502), generate_ranks (:504-517), get_occs (:471-480) and the .dict
 * image of dict_to_file (include/pfbwt_io.hpp:71-82). */
static void finish_from_phrases(pfo_t *o, phr_t *ph, uint64_t P) {
    for (uint64_t k = 0; k < P; ++k) ph[k].id = k;
    phr_t *srt = (phr_t *)malloc(P * sizeof(phr_t));
    memcpy(srt, ph, P * sizeof(phr_t));
    qsort(srt, P, sizeof(phr_t), phr_cmp);
    o->P = P;
    o->parse = (uint32_t *)malloc(P * sizeof(uint32_t));
    uint64_t Dw = 0, dsize = 0;
    for (uint64_t k = 0; k < P; ++k)
        if (k == 0 || phr_cmp(&srt[k - 1], &srt[k]) != 0) { ++Dw; dsize += srt[k].len + 1; }
    dsize += 1;
    o->Dw = Dw; o->dsize = dsize;
    o->dict = (uint8_t *)malloc(dsize);
    o->wstart = (uint64_t *)malloc((Dw + 1) * sizeof(uint64_t));
    o->occ = (uint64_t *)calloc(Dw, sizeof(uint64_t));
    uint64_t r = 0, d = 0;
    for (uint64_t k = 0; k < P; ++k) {
        if (k == 0 || phr_cmp(&srt[k - 1], &srt[k]) != 0) {
            o->wstart[r] = d;
            memcpy(o->dict + d, srt[k].s, srt[k].len);
            d += srt[k].len;
            o->dict[d++] = PFO_ENDOFWORD;
            ++r;
        }
        o->occ[r - 1] += 1;
        o->parse[srt[k].id] = (uint32_t)r; /* 1-based rank */
    }
    o->wstart[Dw] = d;
    o->dict[d++] = PFO_ENDOFDICT;
    free(srt);
}

/* include/pfparser.hpp:299-369 (add_fasta) + :484-492 (finalize) + :595-601 */
int pfo_parse(pfo_t *o) {
    const uint64_t n = o->n, w = (uint64_t)o->w;
    if (o->w < 1 || o->w > 31) { snprintf(o->err, sizeof(o->err), "window size w must be < 32!"); return -1; }
    if (n == 0) { snprintf(o->err, sizeof(o->err), "empty input"); return -1; }
    /* T' = Dollar . S . Dollar^w    */
    uint8_t *T = (uint8_t *)malloc(n + w + 1);
    T[0] = PFO_DOLLAR; memcpy(T + 1, o->S, n); memset(T + 1 + n, PFO_DOLLAR, w);
    uint64_t mask = (1ULL << (2 * w)) - 1, kmer = 0;
    uint64_t cap = 1024, P = 0;
    uint64_t *ends = (uint64_t *)malloc(cap * sizeof(uint64_t));
    for (uint64_t q = 1; q <= n; ++q) {
        kmer = ((kmer << 2) | (uint64_t)nt4_ntoa(T[q])) & mask;   /* include/hash.hpp:32 */
        if (q > w && pfo_wang_hash(kmer) % (uint64_t)o->p == 0) {   /* include/pfparser.hpp:347 */
            if (P + 2 > cap) { cap *= 2; ends = (uint64_t *)realloc(ends, cap * sizeof(uint64_t)); }
            ends[P++] = q;
        }
    }
    ends[P++] = n + w;   /* finalize(): phrase closed by w Dollars, include/pfparser.hpp:485-489 */
    phr_t *ph = (phr_t *)malloc(P * sizeof(phr_t));
    o->last = (uint8_t *)malloc(P);
    o->sai = (uint64_t *)malloc(P * sizeof(uint64_t));
    for (uint64_t k = 0; k < P; ++k) {
        uint64_t s = k ? ends[k - 1] - w + 1 : 0;
        ph[k].s = T + s; ph[k].len = ends[k] - s + 1;
        o->last[k] = T[ends[k] - w];      /* include/pfparser.hpp:599 */
        o->sai[k] = ends[k];              /* include/pfparser.hpp:600 */
    }
    finish_from_phrases(o, ph, P);
    free(ph); free(ends); free(T);
    return 0;
}

/* ---- suffix array of an integer string by prefix doubling --------------- */
typedef struct { uint64_t r0, r1, idx; } dbl_t;
static int dbl_cmp(const void *a, const void *b) {
    const dbl_t *x = (const dbl_t *)a, *y = (const dbl_t *)b;
    if (x->r0 != y->r0) return x->r0 < y->r0 ? -1 : 1;
    if (x->r1 != y->r1) return x->r1 < y->r1 ? -1 : 1;
    return 0;
}
/* sym[0..m): suffix i < suffix j under plain lexicographic order where running
 * off the end is smaller than any symbol.  Replaces sacak_int (gsa/gsacak.c:2499). */
static void sa_doubling(const uint64_t *sym, uint64_t m, uint64_t *sa) {
    dbl_t *e = (dbl_t *)malloc(m * sizeof(dbl_t));
    uint64_t *rank = (uint64_t *)malloc(m * sizeof(uint64_t));
    for (uint64_t i = 0; i < m; ++i) rank[i] = sym[i] + 1;
    for (uint64_t h = 1;; h *= 2) {
        for (uint64_t i = 0; i < m; ++i) { e[i].r0 = rank[i]; e[i].r1 = i + h < m ? rank[i + h] : 0; e[i].idx = i; }
        qsort(e, m, sizeof(dbl_t), dbl_cmp);
        uint64_t r = 0; int all_unique = 1;
        for (uint64_t i = 0; i < m; ++i) {
            if (i == 0 || dbl_cmp(&e[i - 1], &e[i]) != 0) ++r; else all_unique = 0;
            rank[e[i].idx] = r;
        }
        if (all_unique || h >= m) break;
    }
    for (uint64_t i = 0; i < m; ++i) sa[i] = e[i].idx;
    free(e); free(rank);
}

/* include/pfparser.hpp:379-467 (bwt_of_parse) */
int pfo_parse_bwt(pfo_t *o) {
    const uint64_t P = o->P;
    if (P == 1) { snprintf(o->err, sizeof(o->err), "error: only one dict word total. Re-run with a smaller p modulus"); return -1; }
    uint64_t *R = (uint64_t *)malloc((P + 1) * sizeof(uint64_t));
    for (uint64_t i = 0; i < P; ++i) R[i] = o->parse[i];
    R[P] = 0;                                   /* EOS, :407-410 */
    uint64_t *SA = (uint64_t *)malloc((P + 1) * sizeof(uint64_t));
    sa_doubling(R, P + 1, SA);                  /* :425 */
    if (SA[0] != P) { snprintf(o->err, sizeof(o->err), "SA[0] != n"); return -1; }
    o->bwlast = (uint8_t *)malloc(P + 1);
    o->bwsai = (uint64_t *)malloc((P + 1) * sizeof(uint64_t));
    o->ilist = (uint64_t *)malloc((P + 1) * sizeof(uint64_t));
    uint64_t *bwtp = (uint64_t *)malloc((P + 1) * sizeof(uint64_t));
    for (uint64_t i = 0; i <= P; ++i) {         /* :429-451 */
        uint64_t s = SA[i];
        if (s == 0) { bwtp[i] = 0; o->bwlast[i] = 0; o->bwsai[i] = 0; }
        else {
            o->bwlast[i] = s == 1 ? o->last[P - 1] : o->last[s - 2];
            o->bwsai[i] = o->sai[s - 1];
            bwtp[i] = R[s - 1];
        }
    }
    uint64_t *F = (uint64_t *)calloc(o->Dw + 1, sizeof(uint64_t));   /* :452-456 */
    F[0] = 0; F[1] = 1;
    for (uint64_t i = 2; i < o->Dw + 1; ++i) F[i] = F[i - 1] + o->occ[i - 2];
    for (uint64_t i = 0; i <= P; ++i) o->ilist[F[bwtp[i]]++] = i;     /* :459-462 */
    free(R); free(SA); free(bwtp); free(F);
    return 0;
}

/* ---- dictionary suffixes -------------------------------------------------- */
typedef struct { uint64_t start, len, word; } dsuf_t;
static const uint8_t *g_dict;
static int dsuf_cmp(const void *a, const void *b) {
    /* order of gsacak on the dict restricted to one word: bytes up to (not incl.)
     * EndOfWord, EndOfWord smaller than everything (include/pfbwt.hpp:206-211) */
    const dsuf_t *x = (const dsuf_t *)a, *y = (const dsuf_t *)b;
    uint64_t m = x->len < y->len ? x->len : y->len;
    int c = memcmp(g_dict + x->start, g_dict + y->start, m);
    if (c) return c;
    if (x->len != y->len) return x->len < y->len ? -1 : 1;
    return 0;
}
typedef struct { uint64_t bwtp; uint8_t c; } hard_t;
static int hard_cmp(const void *a, const void *b) {   /* SuffixT::operator<, include/pfbwt.hpp:29-31 */
    const hard_t *x = (const hard_t *)a, *y = (const hard_t *)b;
    return x->bwtp < y->bwtp ? -1 : x->bwtp > y->bwtp;
}

/* include/pfbwt.hpp:96-194 (generate_bwt_lcp) + src/pfbwt-f.cpp:298-328 (writer lambda) */
int pfo_build_bwt(pfo_t *o, int any_sa) {
    const uint64_t w = (uint64_t)o->w, n = o->n, Dw = o->Dw;
    uint64_t ns = 0;
    for (uint64_t d = 0; d < Dw; ++d) {
        uint64_t len = o->wstart[d + 1] - 1 - o->wstart[d];
        if (len > w) ns += len - w;
    }
    dsuf_t *suf = (dsuf_t *)malloc((ns + 1) * sizeof(dsuf_t));
    uint64_t k = 0;
    for (uint64_t d = 0; d < Dw; ++d) {
        uint64_t s = o->wstart[d], len = o->wstart[d + 1] - 1 - s;
        for (uint64_t off = 0; off + w < len; ++off) {   /* suff_len > w, include/pfbwt.hpp:114 */
            suf[k].start = s + off; suf[k].len = len - off; suf[k].word = d; ++k;
        }
    }
    g_dict = o->dict;
    qsort(suf, ns, sizeof(dsuf_t), dsuf_cmp);
    /* cumulative occ: ilist slots of word d are ilist[1+cum[d] .. 1+cum[d+1]) (include/pfbwt.hpp:259-268) */
    uint64_t *cum = (uint64_t *)malloc((Dw + 1) * sizeof(uint64_t));
    cum[0] = 0; for (uint64_t d = 0; d < Dw; ++d) cum[d + 1] = cum[d] + o->occ[d];
    o->bwt = (uint8_t *)malloc(n + 1);
    o->sa = (uint64_t *)malloc((n + 1) * sizeof(uint64_t));
    uint64_t pos = 0; o->easy = o->hard = 0;
    hard_t *hv = NULL; uint64_t hcap = 0;
    for (uint64_t i = 0; i < ns;) {
        uint64_t j = i + 1;
        uint64_t sl = suf[i].len;
        if (suf[i].start == o->wstart[suf[i].word]) {       /* full word: EASY1, :116-128 */
            uint64_t d = suf[i].word;
            for (uint64_t t = cum[d]; t < cum[d + 1]; ++t) {
                uint64_t jj = o->ilist[1 + t];
                if (pos > n) goto overflow;
                o->bwt[pos] = o->bwlast[jj];
                o->sa[pos] = any_sa ? o->bwsai[jj] - sl : 0;
                ++pos; ++o->easy;
            }
            i = j; continue;
        }
        while (j < ns && dsuf_cmp(&suf[i], &suf[j]) == 0) ++j;   /* glcp[j] >= suff_len, :137 */
        int same = 1; uint64_t tot = 0;
        uint8_t c0 = suf[i].start - 1 ? o->dict[suf[i].start - 1] : 0;   /* :132 */
        for (uint64_t g = i; g < j; ++g) {
            uint8_t c = suf[g].start - 1 ? o->dict[suf[g].start - 1] : 0;
            if (c != c0) same = 0;
            tot += o->occ[suf[g].word];
        }
        if (pos + tot > n + 1) goto overflow;
        if ((!any_sa && same) || (any_sa && j - i == 1)) {      /* EASY2, :146-159 */
            for (uint64_t g = i; g < j; ++g) {
                uint64_t d = suf[g].word;
                for (uint64_t t = cum[d]; t < cum[d + 1]; ++t) {
                    uint64_t jj = o->ilist[1 + t];
                    o->bwt[pos] = c0;
                    o->sa[pos] = any_sa ? o->bwsai[jj] - sl : 0;
                    ++pos; ++o->easy;
                }
            }
        } else {                                                  /* HARD, :160-182 */
            if (tot > hcap) { hcap = tot * 2; hv = (hard_t *)realloc(hv, hcap * sizeof(hard_t)); }
            uint64_t m = 0;
            for (uint64_t g = i; g < j; ++g) {
                uint64_t d = suf[g].word;
                uint8_t c = suf[g].start - 1 ? o->dict[suf[g].start - 1] : 0;
                for (uint64_t t = cum[d]; t < cum[d + 1]; ++t) { hv[m].bwtp = o->ilist[1 + t]; hv[m].c = c; ++m; }
            }
            qsort(hv, m, sizeof(hard_t), hard_cmp);
            for (uint64_t t = 0; t < m; ++t) {
                o->bwt[pos] = hv[t].c;
                o->sa[pos] = any_sa ? o->bwsai[hv[t].bwtp] - sl : 0;
                ++pos; ++o->hard;
            }
        }
        i = j;
    }
    free(hv); free(suf); free(cum);
    if (pos != n + 1) { snprintf(o->err, sizeof(o->err), "emitted %llu positions, expected %llu", (unsigned long long)pos, (unsigned long long)(n + 1)); return -1; }
    if (any_sa) o->sa[0] = n;                                     /* src/pfbwt-f.cpp:301 */
    /* runs + samples, src/pfbwt-f.cpp:298-328 */
    o->ssa = (uint64_t *)malloc(2 * (n + 2) * sizeof(uint64_t));
    o->esa = (uint64_t *)malloc(2 * (n + 2) * sizeof(uint64_t));
    o->r = 0; o->nssa = o->nesa = 0;
    uint8_t pb = 0;
    for (uint64_t i = 0; i <= n; ++i) {
        if (o->bwt[i] != pb) {
            ++o->r;
            o->ssa[2 * o->nssa] = i; o->ssa[2 * o->nssa + 1] = o->sa[i]; ++o->nssa;
            if (i) { o->esa[2 * o->nesa] = i - 1; o->esa[2 * o->nesa + 1] = o->sa[i - 1]; ++o->nesa; }
        }
        pb = o->bwt[i];
    }
    o->esa[2 * o->nesa] = n; o->esa[2 * o->nesa + 1] = o->sa[n]; ++o->nesa;   /* :325-328 */
    return 0;
overflow:
    free(hv); free(suf); free(cum);
    snprintf(o->err, sizeof(o->err), "emission overflow");
    return -1;
}

/* independent check (not PFP): SA of S.$ */
int pfo_naive_sa(const uint8_t *S, uint64_t n, uint64_t *sa_out) {
    uint64_t *sym = (uint64_t *)malloc((n + 1) * sizeof(uint64_t));
    for (uint64_t i = 0; i < n; ++i) sym[i] = (uint64_t)S[i] + 1;
    sym[n] = 0;
    sa_doubling(sym, n + 1, sa_out);
    free(sym);
    return 0;
}

/* ---- files ---------------------------------------------------------------- */
static int wr_u(FILE *f, const uint64_t *v, uint64_t m, int U) {
    for (uint64_t i = 0; i < m; ++i) {
        if (U == 4) { uint32_t x = (uint32_t)v[i]; if (fwrite(&x, 4, 1, f) != 1) return -1; }
        else { if (fwrite(&v[i], 8, 1, f) != 1) return -1; }
    }
    return 0;
}
static FILE *open_ext(const char *prefix, const char *ext, const char *mode) {
    char path[4096]; snprintf(path, sizeof(path), "%s.%s", prefix, ext);
    return fopen(path, mode);
}

/* include/pfbwt_io.hpp:234-249 (save_parser) + src/pfbwt-f.cpp:229-243 */
int pfo_write_parse_files(const pfo_t *o, const char *prefix, int U, int with_sai) {
    FILE *f;
    if (!(f = open_ext(prefix, "dict", "wb"))) return -1; fwrite(o->dict, 1, o->dsize, f); fclose(f);
    if (!(f = open_ext(prefix, "occ", "wb"))) return -1; wr_u(f, o->occ, o->Dw, U); fclose(f);
    if (!(f = open_ext(prefix, "parse", "wb"))) return -1; fwrite(o->parse, 4, o->P, f); fclose(f);
    if (!(f = open_ext(prefix, "n", "w"))) return -1; fprintf(f, "%llu\n", (unsigned long long)o->n); fclose(f);
    if (o->bwlast) {
        if (!(f = open_ext(prefix, "bwlast", "wb"))) return -1; fwrite(o->bwlast, 1, o->P + 1, f); fclose(f);
        if (!(f = open_ext(prefix, "ilist", "wb"))) return -1; wr_u(f, o->ilist, o->P + 1, U); fclose(f);
        if (with_sai) { if (!(f = open_ext(prefix, "bwsai", "wb"))) return -1; wr_u(f, o->bwsai, o->P + 1, U); fclose(f); }
    }
    return 0;
}
/* src/pfbwt-f.cpp:298-328 */
int pfo_write_bwt_files(const pfo_t *o, const char *prefix, int U, int sa, int rssa) {
    FILE *f;
    if (!(f = open_ext(prefix, "bwt", "wb"))) return -1; fwrite(o->bwt, 1, o->n + 1, f); fclose(f);
    if (sa) { if (!(f = open_ext(prefix, "sa", "wb"))) return -1; wr_u(f, o->sa, o->n + 1, U); fclose(f); }
    if (rssa) {
        if (!(f = open_ext(prefix, "ssa", "wb"))) return -1; wr_u(f, o->ssa, 2 * o->nssa, U); fclose(f);
        if (!(f = open_ext(prefix, "esa", "wb"))) return -1; wr_u(f, o->esa, 2 * o->nesa, U); fclose(f);
    }
    return 0;
}

/* include/pfbwt_io.hpp:104-148,211-222 (load_dict, load_parse_ranks, load_parser) */
int pfo_load_parse(pfo_t *o, const char *prefix) {
    FILE *f = open_ext(prefix, "dict", "rb");
    if (!f) { snprintf(o->err, sizeof(o->err), "bad dict file"); return -1; }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    o->dict = (uint8_t *)malloc((size_t)sz); o->dsize = (uint64_t)sz;
    if (fread(o->dict, 1, (size_t)sz, f) != (size_t)sz) { fclose(f); return -1; }
    fclose(f);
    uint64_t Dw = 0; for (long i = 0; i < sz; ++i) if (o->dict[i] == PFO_ENDOFWORD) ++Dw;
    o->Dw = Dw; o->wstart = (uint64_t *)malloc((Dw + 1) * sizeof(uint64_t));
    uint64_t d = 0; o->wstart[0] = 0;
    for (long i = 0; i < sz; ++i) if (o->dict[i] == PFO_ENDOFWORD) o->wstart[++d] = (uint64_t)i + 1;
    f = open_ext(prefix, "parse", "rb");
    if (!f) { snprintf(o->err, sizeof(o->err), "bad parse file"); return -1; }
    fseek(f, 0, SEEK_END); sz = ftell(f); fseek(f, 0, SEEK_SET);
    o->P = (uint64_t)sz / 4; o->parse = (uint32_t *)malloc((size_t)sz);
    if (fread(o->parse, 4, o->P, f) != o->P) { fclose(f); return -1; }
    fclose(f);
    return 0;
}

/* PfParser::operator+= (include/pfparser.hpp:194-263) applied left to right over
 * parts[], each given as (dict, wstart, parse, P); phrases are re-materialised as
 * init_from_dict_ranks does (:549-567).  Result = parse of the concatenation. */
int pfo_merge(pfo_t *out, pfo_t *const *parts, int nparts) {
    const uint64_t w = (uint64_t)out->w;
    uint64_t totP = 0, totB = 0;
    for (int g = 0; g < nparts; ++g) {
        totP += parts[g]->P + w + 1;
        for (uint64_t k = 0; k < parts[g]->P; ++k) {
            uint64_t d = parts[g]->parse[k] - 1;
            totB += parts[g]->wstart[d + 1] - parts[g]->wstart[d] + 2 * w + 2;
        }
    }
    uint8_t *buf = (uint8_t *)malloc(totB + 16);
    phr_t *ph = (phr_t *)malloc(totP * sizeof(phr_t));
    uint8_t *last = (uint8_t *)malloc(totP);
    uint64_t *sai = (uint64_t *)malloc(totP * sizeof(uint64_t));
    uint64_t P = 0, b = 0, pos = 0;
#define PROCESS_PHRASE(ptr, len_) do { ph[P].s = (ptr); ph[P].len = (len_); \
        last[P] = (ptr)[(len_) - w - 1]; sai[P] = pos; ++P; } while (0)   /* :595-601 */
    for (int g = 0; g < nparts; ++g) {
        const pfo_t *r = parts[g];
#define RWORD(k) (r->dict + r->wstart[r->parse[k] - 1])
#define RLEN(k) (r->wstart[r->parse[k]] - 1 - r->wstart[r->parse[k] - 1])
        if (r->P == 0) continue;
        if (P == 0) {   /* operator=(rhs), :196 ; positions per init_from_dict_ranks :555-562 */
            for (uint64_t k = 0; k < r->P; ++k) {
                uint64_t len = RLEN(k);
                memcpy(buf + b, RWORD(k), len);
                pos += k ? len - w : len - 1;
                PROCESS_PHRASE(buf + b, len);
                b += len;
            }
            continue;
        }
        /* pop lhs last phrase, strip Dollars (:198-219) */
        --P;
        uint8_t *phrase = (uint8_t *)ph[P].s; uint64_t plen = ph[P].len;
        if (phrase[plen - 1] == PFO_DOLLAR) { plen -= w; pos -= w; }
        if (RWORD(0)[0] != PFO_DOLLAR) { snprintf(out->err, sizeof(out->err), "rhs parser malformed"); return -1; }
        /* the popped phrase is the most recent thing in buf: extend it in place */
        b = (uint64_t)(phrase - buf) + plen;
        uint64_t kmer = 0, mask = (1ULL << (2 * w)) - 1;   /* hasher primed with w 'A's == 0, :226-227 */
        uint64_t len0 = RLEN(0);
        const uint8_t *w0 = RWORD(0);
        for (uint64_t i = 0; i < w; ++i) {                 /* :232-245 */
            uint8_t c = w0[1 + i];
            buf[b++] = c; ++plen;
            kmer = ((kmer << 2) | (uint64_t)nt4_ntoa(c)) & mask;
            if (pfo_wang_hash(kmer) % (uint64_t)out->p == 0) {
                ++pos;
                PROCESS_PHRASE(phrase, plen);
                /* phrase.erase(0, size-w): next phrase starts with the last w chars */
                memmove(buf + b, phrase + plen - w, w);   /* non-overlapping: b == phrase+plen */
                phrase = buf + b; plen = w; b += w;
                --pos;
            }
            ++pos;
        }
        memcpy(buf + b, w0 + w + 1, len0 - w - 1);         /* :248-250 */
        b += len0 - w - 1; plen += len0 - w - 1;
        pos += len0 - w - 1;
        PROCESS_PHRASE(phrase, plen);
        for (uint64_t k = 1; k < r->P; ++k) {              /* :252-257 */
            uint64_t len = RLEN(k);
            memcpy(buf + b, RWORD(k), len);
            pos += len - w;
            PROCESS_PHRASE(buf + b, len);
            b += len;
        }
    }
    out->last = (uint8_t *)malloc(P); memcpy(out->last, last, P);
    out->sai = (uint64_t *)malloc(P * sizeof(uint64_t)); memcpy(out->sai, sai, P * sizeof(uint64_t));
    out->n = pos >= w ? pos - w : 0;                        /* get_n, :529-532 */
    finish_from_phrases(out, ph, P);
    free(buf); free(ph); free(last); free(sai);
    return 0;
}


/* ---- marker-array join: src/mps_to_ma.cpp:43-50 -> write_marker_array (include/marker_array.hpp:137-174) over an
 * rle_window_arr (include/rle_window_array.hpp).  The reference answers at(s) with rank/select on bit vectors of the
 * run starts and ends (:118-130): "the number of starts <= s is one more than the number of ends < s", i.e. s lies in
 * the run [first, last] found by a search over the (ascending) starts.  A run that starts or ends at 0 makes the
 * reference exit (:28-31, rs == re == 0 there); marker lists are compared by content (vec_eq, marker_array.hpp:148). */
static uint64_t *mj_in; static uint64_t mj_R; static uint64_t *mj_start, *mj_end, *mj_voff;
static long mj_at(uint64_t s) {
    if (!mj_R || s < mj_start[0]) return -1;
    uint64_t lo = 0, hi = mj_R;
    while (hi - lo > 1) { uint64_t mid = lo + (hi - lo) / 2; if (mj_start[mid] <= s) lo = mid; else hi = mid; }
    return s <= mj_end[lo] ? (long)lo : -1;
}
static int mj_same(long a, long b) {
    if (a == b) return 1;
    if (a < 0 || b < 0) return 0;
    uint64_t ka = mj_voff[a + 1] - mj_voff[a], kb = mj_voff[b + 1] - mj_voff[b];
    if (ka != kb) return 0;
    return memcmp(mj_in + mj_voff[a], mj_in + mj_voff[b], ka * 8) == 0;
}
int pfo_marker_join(const char *mps_path, const char *sa_path, const char *out_path) {
    FILE *f = fopen(mps_path, "rb");
    if (!f) return -1;
    fseek(f, 0, SEEK_END); long bytes = ftell(f); fseek(f, 0, SEEK_SET);
    uint64_t nw = (uint64_t)bytes / 8;
    mj_in = (uint64_t *)malloc((nw + 1) * 8);
    if (nw && fread(mj_in, 8, nw, f) != nw) { fclose(f); return -1; }
    fclose(f);
    mj_R = 0;
    for (uint64_t i = 0; i < nw; ++i) if (mj_in[i] == ~0ull) ++mj_R;
    mj_start = (uint64_t *)malloc((mj_R + 1) * 8); mj_end = (uint64_t *)malloc((mj_R + 1) * 8); mj_voff = (uint64_t *)malloc((mj_R + 2) * 8);
    uint64_t r = 0, b0 = 0;
    for (uint64_t i = 0; i < nw; ++i) {
        if (mj_in[i] != ~0ull) continue;
        if (i - b0 < 3) return -2;                                             /* two keys, at least one marker */
        mj_start[r] = mj_in[b0]; mj_end[r] = mj_in[b0 + 1]; mj_voff[r] = b0 + 2; /* values stay in place: voff indexes mj_in */
        if (mj_start[r] == 0 || mj_end[r] == 0) { fprintf(stderr, "warning equal run start at 0 or run end at 0\n"); return 1; }
        ++r; b0 = i + 1;
    }
    /* voff[r+1] - voff[r] must be the value count: keep explicit ends */
    uint64_t *vend = (uint64_t *)malloc((mj_R + 1) * 8);
    r = 0; b0 = 0;
    for (uint64_t i = 0; i < nw; ++i) if (mj_in[i] == ~0ull) { vend[r++] = i; b0 = i + 1; }
    /* compact the values so that voff is a prefix array */
    uint64_t *vals = (uint64_t *)malloc((nw + 1) * 8), *voff = (uint64_t *)malloc((mj_R + 2) * 8), t = 0;
    for (r = 0; r < mj_R; ++r) { voff[r] = t; for (uint64_t j = mj_voff[r]; j < vend[r]; ++j) vals[t++] = mj_in[j]; }
    voff[mj_R] = t;
    free(mj_in); free(mj_voff); free(vend);
    mj_in = vals; mj_voff = voff;
    FILE *sa = !strcmp(sa_path, "-") ? stdin : fopen(sa_path, "rb");
    FILE *o = fopen(out_path, "wb");
    if (!sa || !o) return -1;
    uint64_t s, i = 0, first = 0, last = 0; const uint64_t delim = ~0ull;
    long prev = -1;
    while (fread(&s, 8, 1, sa) == 1) {                                         /* marker_array.hpp:145-163 */
        long cur = mj_at(s);
        if (!mj_same(cur, prev)) {
            if (prev >= 0) {
                fwrite(&first, 8, 1, o); fwrite(&last, 8, 1, o);
                fwrite(mj_in + mj_voff[prev], 8, mj_voff[prev + 1] - mj_voff[prev], o);
                fwrite(&delim, 8, 1, o);
            }
            first = i;
        }
        last = i; prev = cur; ++i;
    }
    if (prev >= 0) {                                                           /* :164-171 */
        fwrite(&first, 8, 1, o); fwrite(&last, 8, 1, o);
        fwrite(mj_in + mj_voff[prev], 8, mj_voff[prev + 1] - mj_voff[prev], o);
        fwrite(&delim, 8, 1, o);
    }
    fclose(o);
    if (sa != stdin) fclose(sa);
    free(mj_in); free(mj_voff); free(mj_start); free(mj_end);
    return 0;
}
