// merge_pfp command-line driver over libpfbwt_b200 (C ABI): concatenates parses
// (loading <prefix>.dict/.parse when present, else parsing the FASTA <prefix>) in the
// order given and writes the merged parse.  Flags, outputs and log line as the
// reference (src/merge_pfp.cpp:26-72 flags, :115-170 merge_pfp).  -t is accepted; the
// merge itself runs on the GPU.
#include "pfbwt_b200.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <getopt.h>
#include <string>
#include <sys/stat.h>
#include <vector>

static bool exists(const std::string &p) { struct stat b; return stat(p.c_str(), &b) == 0; }

int main(int argc, char **argv) {
    std::vector<std::string> prefixes;
    std::string output = "out";
    size_t nthreads = 1;
    int w = 10, p = 100, store_docs = 0, parse_bwt = 0, sai = 0, device = 0;
    static struct option lopts[] = {{"docs", no_argument, &store_docs, 1},
                                    {"window-size", required_argument, NULL, 'w'},
                                    {"mod-val", required_argument, NULL, 'p'},
                                    {"output", required_argument, NULL, 'o'},
                                    {"threads", required_argument, NULL, 't'},
                                    {"parse-bwt", no_argument, &parse_bwt, 1},
                                    {"sai", no_argument, NULL, 's'},
                                    {"device", required_argument, NULL, 'D'},
                                    {0, 0, 0, 0}};
    int c;
    while ((c = getopt_long(argc, argv, "dw:p:o:t:s", lopts, NULL)) != -1) {
        switch (c) {
        case 'd': store_docs = 1; break;
        case 'w': w = atoi(optarg); break;
        case 'p': p = atoi(optarg); break;
        case 'o': output = optarg; break;
        case 't': nthreads = atoi(optarg); break;
        case 's': sai = 1; break;
        case 'D': device = atoi(optarg); break;
        case '?':
            fprintf(stderr, "Unknown option.\nusage: ./merge_pfp [--docs] -w <window size> -p <mod> -o <output prefix> -t <threads> <prefix 1> <prefix 2> ... \n");
            exit(1);
        }
    }
    for (int i = optind; i < argc; ++i) prefixes.push_back(argv[i]);
    fprintf(stderr, "%lu %lu - %lu\n", nthreads, prefixes.size(), nthreads ? prefixes.size() / nthreads : 0);
    pfb_parse_params pp;
    memset(&pp, 0, sizeof(pp));
    pp.w = (uint32_t)w; pp.p = (uint64_t)p; pp.get_sai = sai; pp.store_docs = store_docs;
    pfb_ctx *acc = pfb_create(device, 64);
    if (!acc) { fprintf(stderr, "%s\n", pfb_create_error()); return 1; }
    if (pfb_set_params(acc, &pp)) { fprintf(stderr, "%s\n", pfb_last_error(acc)); return 1; }
    // src/merge_pfp.cpp:122,156-160: with few files per thread the reference runs serially and logs to <out>.pfbwt.log;
    // the threaded mode logs to stderr.  The merge itself is device work either way.
    FILE *logf = stderr;
    if (!(nthreads && prefixes.size() / nthreads > 2)) {
        logf = fopen((output + ".pfbwt.log").c_str(), "w");
        if (logf == NULL) { fprintf(stderr, "error opening log\n"); exit(1); }
        fprintf(logf, "not using threads (%lu files, %lu threads specified)\n", prefixes.size(), nthreads);
        fprintf(stderr, "not using threads (%lu files, %lu threads specified)\n", prefixes.size(), nthreads);
    }
    for (auto &prefix : prefixes) {   // load_or_generate_parser_w_log, include/pfbwt_io.hpp:270-285
        if (exists(prefix + ".dict") && exists(prefix + ".parse")) {
            fprintf(logf, "loading %s, %s, and maybe %s from file\n", (prefix + ".dict").c_str(), (prefix + ".parse").c_str(), (prefix + ".docs").c_str());
            pfb_ctx *rhs = pfb_create(device, 64);
            if (!rhs) { fprintf(stderr, "%s\n", pfb_create_error()); return 1; }
            if (pfb_set_params(rhs, &pp) || pfb_load_parse(rhs, prefix.c_str(), 0, 0)) { fprintf(stderr, "%s\n", pfb_last_error(rhs)); return 1; }
            if (pfb_append(acc, rhs)) { fprintf(stderr, "%s\n", pfb_last_error(acc)); return 1; }
            pfb_destroy(rhs);
        } else {
            fprintf(logf, "generating parse for %s\n", prefix.c_str());
            if (exists(prefix)) {
                if (pfb_add_fasta(acc, prefix.c_str())) { fprintf(stderr, "%s\n", pfb_last_error(acc)); return 1; }
            } else {
                fprintf(logf, "ERROR: %s not found, cannot add it to parse!\n", prefix.c_str());
            }
        }
    }
    if (logf != stderr) fflush(logf);
    if (pfb_finalize(acc)) { fprintf(stderr, "%s\n", pfb_last_error(acc)); return 1; }
    if (parse_bwt && pfb_parse_bwt(acc)) { fprintf(stderr, "%s: Success\n", pfb_last_error(acc)); return 1; }
    if (pfb_save_parse(acc, output.c_str(), parse_bwt, sai)) { fprintf(stderr, "%s\n", pfb_last_error(acc)); return 1; }
    pfb_destroy(acc);
    return 0;
}
