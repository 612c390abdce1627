/* pfp_oracle_main.c -- TEST INFRASTRUCTURE ONLY.  CLI around pfp_oracle.c that
 * writes the same file set as the reference's pfbwt-f / pfbwt-f64 / merge_pfp
 * (src/pfbwt-f.cpp:209-245,272-349; src/merge_pfp.cpp:153-168).
 *
 *   pfp_oracle [-w W] [-p P] [-s] [-r] [--bits 32|64] [--non-acgt-to-a]
 *              [--parse-only] [-o prefix] in.fa
 *   pfp_oracle --merge [-w W] [-p P] [-s] [--bits 64] -o out pfx1 pfx2 ...
 */
#include "pfp_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char **argv) {
    if (argc == 5 && !strcmp(argv[1], "--marker-join")) {            /* pfp_oracle --marker-join <mps> <sa|-> <out> */
        int rc = pfo_marker_join(argv[2], argv[3], argv[4]);
        if (rc < 0) fprintf(stderr, "bad mps / sa input\n");
        return rc ? 1 : 0;
    }
    int w = 10, p = 100, sa = 0, rssa = 0, bits = 32, ntoa = 0, parse_only = 0, merge = 0, write_last = 0;
    const char *out = NULL; const char *in[1024]; int nin = 0;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "-w")) w = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-p")) p = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-s")) sa = 1;
        else if (!strcmp(argv[i], "-r")) rssa = 1;
        else if (!strcmp(argv[i], "--bits")) bits = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--non-acgt-to-a")) ntoa = 1;
        else if (!strcmp(argv[i], "--parse-only")) parse_only = 1;
        else if (!strcmp(argv[i], "--merge")) merge = 1;
        else if (!strcmp(argv[i], "--write-last")) write_last = 1;   /* <prefix>.last / .sai = last_ / sai_ (include/pfparser.hpp:599-600) */
        else if (!strcmp(argv[i], "-o")) out = argv[++i];
        else if (nin < 1024) in[nin++] = argv[i];
    }
    if (nin < 1) { fprintf(stderr, "usage: see source header\n"); return 2; }
    int U = bits == 64 ? 8 : 4;
    pfo_t o; pfo_init(&o, w, p, ntoa);
    if (merge) {
        pfo_t *parts[1024];
        for (int g = 0; g < nin; ++g) {
            parts[g] = (pfo_t *)malloc(sizeof(pfo_t)); pfo_init(parts[g], w, p, 0);
            if (pfo_load_parse(parts[g], in[g])) { fprintf(stderr, "%s\n", parts[g]->err); return 1; }
        }
        if (!out) out = "out";
        if (pfo_merge(&o, parts, nin)) { fprintf(stderr, "%s\n", o.err); return 1; }
        if (pfo_parse_bwt(&o)) { fprintf(stderr, "%s\n", o.err); return 1; }
        if (pfo_write_parse_files(&o, out, U, sa)) return 1;
        return 0;
    }
    if (!out) out = in[0];
    if (pfo_read_fasta(&o, in[0])) { fprintf(stderr, "%s\n", o.err); return 1; }
    if (pfo_parse(&o)) { fprintf(stderr, "%s\n", o.err); return 1; }
    if (write_last) {
        char fn[4096];
        FILE *f;
        snprintf(fn, sizeof(fn), "%s.last", out);
        if (!(f = fopen(fn, "wb"))) return 1;
        fwrite(o.last, 1, o.P, f); fclose(f);
        snprintf(fn, sizeof(fn), "%s.sai", out);
        if (!(f = fopen(fn, "wb"))) return 1;
        for (uint64_t k = 0; k < o.P; ++k) { if (U == 4) { uint32_t v = (uint32_t)o.sai[k]; fwrite(&v, 4, 1, f); } else fwrite(&o.sai[k], 8, 1, f); }
        fclose(f);
    }
    if (pfo_parse_bwt(&o)) { fprintf(stderr, "%s\n", o.err); return 1; }
    if (pfo_write_parse_files(&o, out, U, sa || rssa)) return 1;
    if (parse_only) return 0;
    if (pfo_build_bwt(&o, sa || rssa)) { fprintf(stderr, "%s\n", o.err); return 1; }
    if (pfo_write_bwt_files(&o, out, U, sa, rssa)) return 1;
    fprintf(stderr, "# easy cases: %llu, # hard cases: %llu\nn: %llu\nr: %llu\n",
            (unsigned long long)o.easy, (unsigned long long)o.hard, (unsigned long long)o.n, (unsigned long long)o.r);
    pfo_free(&o);
    return 0;
}
