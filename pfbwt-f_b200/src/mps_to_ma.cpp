// mps_to_ma command-line driver over libpfbwt_b200 (C ABI): joins the marker runs of <mps> with a suffix array read as
// u64 entries from a file or from stdin ("-") and writes the marker array, like the reference's tool
// (src/mps_to_ma.cpp:16-50: `mps_to_ma [-m] [-o out] <mps> <sa>`; the join itself: include/marker_array.hpp:137-174).
// -m (mmap'ed tables in the reference) is accepted: the tables live in device memory either way.
#include "pfbwt_b200.h"
#include <cstdio>
#include <cstdlib>
#include <getopt.h>
#include <string>

int main(int argc, char **argv) {
    std::string output = "out";
    int device = 0, c;
    static struct option lopts[] = {{"mmap", no_argument, NULL, 'm'},
                                    {"output", required_argument, NULL, 'o'},
                                    {"device", required_argument, NULL, 'D'},
                                    {0, 0, 0, 0}};
    while ((c = getopt_long(argc, argv, "o:mh", lopts, NULL)) != -1) {
        switch (c) {
        case 'm': break;
        case 'o': output = optarg; break;
        case 'D': device = atoi(optarg); break;
        case 'h': fprintf(stderr, "usage: ./mps_to_ma [-m] [-o <output file>] <marker positions (.mps)> <suffix array file, or - for stdin>\n"); return 0;
        case '?': fprintf(stderr, "Unknown option.\n"); exit(1);
        case ':': fprintf(stderr, "no argument specified for option\n"); exit(1);
        }
    }
    if (argc - optind < 2) { fprintf(stderr, "usage: ./mps_to_ma [-m] [-o <output file>] <marker positions (.mps)> <suffix array file, or - for stdin>\n"); return 1; }
    const char *mps = argv[optind], *sa = argv[optind + 1];
    pfb_ctx *ctx = pfb_create(device, 64);
    if (!ctx) { fprintf(stderr, "%s\n", pfb_create_error()); return 1; }
    if (pfb_marker_join(ctx, mps, sa, output.c_str())) { fprintf(stderr, "%s\n", pfb_last_error(ctx)); pfb_destroy(ctx); return 1; }
    pfb_destroy(ctx);
    return 0;
}
