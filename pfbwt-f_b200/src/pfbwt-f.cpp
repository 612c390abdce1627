// pfbwt-f / pfbwt-f64 command-line driver over libpfbwt_b200 (C ABI).  Same flags,
// file set and stderr lines as the reference driver (src/pfbwt-f.cpp:102-182 flags,
// :209-245 run_parser, :272-349 run_pfbwt, :351-367 main).  -DM64 builds pfbwt-f64.
#include "pfbwt_b200.h"
#include <chrono>
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <getopt.h>
#include <string>

#ifdef M64
static const int BITS = 64;
#else
static const int BITS = 32;
#endif

struct Args {
    std::string in_fname, output, stdout_ext, marker_mps;
    size_t w = 10, p = 100;
    int sa = 0, rssa = 0, mmap = 0, parse_only = 0, trim_non_acgt = 0, non_acgt_to_a = 0, pfbwt_only = 0, verbose = 0, print_docs = 0;
    int write_last = 0, write_sai = 0, doc_array = 0;
    int device = 0;
};

struct Timer {   // src/pfbwt-f.cpp:35-50
    std::chrono::time_point<std::chrono::system_clock> start;
    std::string msg;
    Timer(std::string m) : msg(m) { start = std::chrono::system_clock::now(); }
    ~Timer() {
        std::chrono::duration<double> d = std::chrono::system_clock::now() - start;
        fprintf(stderr, "%s%.2fs\n", msg.c_str(), d.count());
    }
};

static void die(const char *s) { perror(s); exit(1); }   // src/utils.c:13-17
static void die_ctx(pfb_ctx *c) {
    const std::string msg = pfb_last_error(c);
    pfb_wait_saves(c);                  // let the background writer finish what it was handed
    fprintf(stderr, "%s\n", msg.c_str());
    exit(1);
}

static void print_help() {
    fprintf(stderr, "%s. use the prefix-free parsing algorithm to build a BWT for %sgenomic data (B200 build).\n\n"
                    "usage\n    ./%s [options] <fasta file>\n\n",
            BITS == 64 ? "pfbwt-f64" : "pfbwt-f", BITS == 64 ? "BIG " : "", BITS == 64 ? "pfbwt-f64" : "pfbwt-f");
    fprintf(stderr,
"results\n    BWT of input saved to <fasta file>.bwt. Header lines are excluded.\n\noptions\n"
"    -o                  output prefix.\n"
"    -s                  Output full suffix array to <fasta file>.sa\n"
"    -r                  Output run-length sampled suffix arrray to <fasta file>.ssa (run-starts) and <fasta file>.esa (run-ends)\n"
"    -w <int>            window-size for parsing [default: 10]\n"
"    -p <int>            modulo for parsing [default: 100]\n"
"    -m                  workspace larger than the device memory: build and write the BWT one key range after the other\n"
"                        (PFB_HBM_LIMIT=<bytes|48G> caps the device memory used; chosen automatically when the work arrays\n"
"                        of a full build would not fit)\n"
"    --parse-only        only produce parse (dict, occ, ilist, last, bwlast)\n"
"    --pfbwt-only        build pfbwt from parse + parse-bwt. Requires -o to match parse files' prefix.\n"
"    --non-acgt-to-a     convert non-ACGT characters to A\n"
"    --print-docs        write <prefix>.docs\n"
"    --write-last        also write <prefix>.last (last char of every phrase, include/utils.h EXTLST)\n"
"    --doc-array         also write <prefix>.da: uint32 record index of SA[i] for every BWT position (implies --print-docs, needs -s or -r)\n"
"    --marker-array <f>  also write <prefix>.ma: the join of the marker runs in <f> (.mps) with the SA, as mps_to_ma writes it from\n"
"                        `--stdout sa` (needs -s or -r; the SA does not leave the device)\n"
"    --write-sai         also write <prefix>.sai (end position of every phrase, include/utils.h EXTSAI)\n"
"    --stdout <ext>      output file ending <ext> will be stdout instead (bwt, sa)\n"
"    --device <int>      CUDA device ordinal [default: 0]\n"
"    -h                  print this help message\n");
}

static Args parse_args(int argc, char **argv) {
    Args args;
    fputs("==== Command line:", stderr);
    for (int i = 0; i < argc; i++) fprintf(stderr, " %s", argv[i]);
    fputs("\n", stderr);
    static struct option lopts[] = {
        {"parse-only", no_argument, &args.parse_only, 1},
        {"pfbwt-only", no_argument, &args.pfbwt_only, 1},
        {"trim-non-acgt", no_argument, &args.trim_non_acgt, 1},
        {"non-acgt-to-a", no_argument, &args.non_acgt_to_a, 1},
        {"print-docs", no_argument, &args.print_docs, 1},
        {"write-last", no_argument, &args.write_last, 1},
        {"write-sai", no_argument, &args.write_sai, 1},
        {"doc-array", no_argument, &args.doc_array, 1},
        {"marker-array", required_argument, NULL, 'M'},
        {"stdout", required_argument, NULL, 'c'},
        {"verbose", no_argument, &args.verbose, 1},
        {"sa", no_argument, NULL, 's'},
        {"rssa", no_argument, NULL, 'r'},
        {"mmap", no_argument, NULL, 'm'},
        {"output", required_argument, NULL, 'o'},
        {"window-size", required_argument, NULL, 'w'},
        {"mod-val", required_argument, NULL, 'p'},
        {"device", required_argument, NULL, 'D'},
        {0, 0, 0, 0}};
    int c;
    while ((c = getopt_long(argc, argv, "w:p:o:hsrfm", lopts, NULL)) != -1) {
        switch (c) {
        case 'f': break;   // legacy
        case 's': args.sa = 1; break;
        case 'r': args.rssa = 1; break;
        case 'w': args.w = atoi(optarg); break;
        case 'm': args.mmap = 1; break;
        case 'p': args.p = atoi(optarg); break;
        case 'h': print_help(); exit(0);
        case 'o': args.output.assign(optarg); break;
        case 'c': args.stdout_ext = optarg; break;
        case 'D': args.device = atoi(optarg); break;
        case 'M': args.marker_mps = optarg; break;
        case '?': fprintf(stderr, "Unknown option. Use -h for help.\n"); exit(1);
        }
    }
    if (argc == optind + 1) args.in_fname.assign(argv[optind]);
    else { fprintf(stderr, "reading from stdin. Parsing might be a bit slow.\n"); args.in_fname.assign("-"); }
    if (args.non_acgt_to_a && args.trim_non_acgt) die("cannot have both --non-acgt-to-a and --trim-non-acgt options enabled at same time");
    if (args.in_fname == "-" && args.output == "") die("if reading from stdin, need a prefix for output files (-o, --output)");
    if (args.in_fname != "-" && args.output == "") args.output = args.in_fname;
    if (args.parse_only && args.pfbwt_only) die("cannot simulatneously do parse_only and pfbwt_only");
    if (!args.marker_mps.empty() && !args.sa && !args.rssa) { fprintf(stderr, "--marker-array needs the suffix array (-s or -r)\n"); exit(1); }
    if (args.doc_array) {
        args.print_docs = 1;
        if (!args.sa && !args.rssa) { fprintf(stderr, "--doc-array needs the suffix array (-s or -r)\n"); exit(1); }
    }
    return args;
}

// PFB_CLI_TIMES=1: finer wall-clock lines than the reference's TASK lines (device start-up, compute vs file writing)
struct SubTimer {
    std::chrono::time_point<std::chrono::system_clock> start;
    const char *what;
    SubTimer(const char *w) : what(w) { start = std::chrono::system_clock::now(); }
    ~SubTimer() {
        static const bool on = getenv("PFB_CLI_TIMES") != NULL;
        std::chrono::duration<double> d = std::chrono::system_clock::now() - start;
        if (on) fprintf(stderr, "TIME\t%s\t%.3fs\n", what, d.count());
    }
};

int main(int argc, char **argv) {
    const auto t_main = std::chrono::system_clock::now();
    Args args = parse_args(argc, argv);
    pfb_ctx *ctx;
    { SubTimer t("device start-up (pfb_create)"); ctx = pfb_create(args.device, BITS); }
    if (!ctx) { fprintf(stderr, "%s\n", pfb_create_error()); return 1; }
    pfb_parse_params pp;
    memset(&pp, 0, sizeof(pp));
    pp.w = (uint32_t)args.w; pp.p = args.p; pp.get_sai = args.sa || args.rssa; pp.store_docs = args.print_docs;
    pp.verbose = args.verbose; pp.trim_non_acgt = args.trim_non_acgt; pp.non_acgt_to_a = args.non_acgt_to_a;
    if (pfb_set_params(ctx, &pp)) die_ctx(ctx);
    const int any_sa = args.sa || args.rssa;
    if (!args.pfbwt_only) {   // run_parser, src/pfbwt-f.cpp:209-245
        fprintf(stderr, "running parser...\n");
        fprintf(stderr, "starting...\n");
        {
            Timer t("TASK\tparsing input\t");
            if (pfb_add_fasta(ctx, args.in_fname.c_str())) die_ctx(ctx);
        }
        {
            // src/pfbwt-f.cpp:221-228: the parse files are written in this task, before bwt_of_parse can fail
            Timer t("TASK\tfinalizing parse, writing dict, occs, and ranks\t");
            { SubTimer t2("parse on device"); if (pfb_finalize(ctx)) die_ctx(ctx); }
            SubTimer t3("handing the parse files to the writer thread");
            // the GB-sized files are written in the background while the device goes on; PFB_SYNC_WRITES=1 writes in place
            if (getenv("PFB_SYNC_WRITES") ? pfb_save_parse(ctx, args.output.c_str(), 0, 0) : pfb_save_parse_async(ctx, args.output.c_str(), 0, 0)) die_ctx(ctx);
            if ((args.write_last || args.write_sai) && pfb_save_last_sai(ctx, args.output.c_str(), args.write_last, args.write_sai)) die_ctx(ctx);
        }
        {
            Timer t("TASK\tranking and bwt-ing parse and processing last-chars\t");
            {
                SubTimer t2("parse BWT on device");
                if (pfb_parse_bwt(ctx)) {
                    fprintf(stderr, "%s: Success\n", pfb_last_error(ctx));
                    pfb_wait_saves(ctx);          // the parse files of the previous task are complete before the exit, as in the reference
                    exit(1);
                }
            }
            SubTimer t3("handing the parse BWT files to the writer thread");
            if (getenv("PFB_SYNC_WRITES") ? pfb_save_parse(ctx, args.output.c_str(), 2, any_sa) : pfb_save_parse_async(ctx, args.output.c_str(), 2, any_sa)) die_ctx(ctx);
        }
        if (args.trim_non_acgt) { FILE *f = fopen((args.output + ".ntab").c_str(), "wb"); if (f) fclose(f); }
        pfb_release_text(ctx);
        if (args.parse_only && pfb_wait_saves(ctx)) die_ctx(ctx);
    } else {
        if (pfb_load_parse(ctx, args.output.c_str(), 1, any_sa)) die_ctx(ctx);
    }
    if (!args.parse_only) {   // run_pfbwt, src/pfbwt-f.cpp:272-349
        fprintf(stderr, "generating BWT using pfbwt algorithm...\n");
        // src/pfbwt-f.cpp:359-365: -m selects the file-backed workspace.  Here: key-range buckets streamed to the files; also
        // taken when a full build would not fit the device memory (the reference would swap or die)
        bool bucketed = args.mmap || getenv("PFB_HBM_LIMIT") != NULL;
        if (!bucketed && args.stdout_ext.empty()) {
            size_t fr = 0;
            const uint64_t need = pfb_build_workspace_bytes(ctx, args.sa, args.rssa);
            if (pfb_device_free_bytes(ctx, &fr) == 0 && need > fr) bucketed = true;
        }
        if (bucketed && !args.stdout_ext.empty()) { fprintf(stderr, "--stdout is not available with -m\n"); exit(1); }
        if (bucketed && !args.marker_mps.empty() && !args.sa) { fprintf(stderr, "--marker-array with a streamed build (-m, or a build that does not fit the device) joins with <prefix>.sa: add -s\n"); exit(1); }
        fprintf(stderr, bucketed ? "workspace will be cut into key-range buckets streamed to the output files\n" : "workspace will be contained in HBM\n");
        {
            Timer t(any_sa ? "TASK\tgenerating final BWT w/ full and/or run-length SA\t" : "TASK\tgenerating final BWT w/o SA\t");
            if (bucketed) {
                if (pfb_build_bwt_to_files(ctx, args.output.c_str(), args.sa, args.rssa, 0)) die_ctx(ctx);
                pfb_stats s;
                pfb_get_stats(ctx, &s);
                fprintf(stderr, "buckets: %lu, peak device bytes: %lu\n", (unsigned long)s.stream_buckets, (unsigned long)s.stream_peak_bytes);
                if (!args.marker_mps.empty() && pfb_write_marker_array(ctx, args.marker_mps.c_str(), (args.output + ".ma").c_str())) die_ctx(ctx);
            } else {
                { SubTimer t2("BWT on device"); if (pfb_build_bwt(ctx, args.sa, args.rssa)) die_ctx(ctx); }
                SubTimer t3("writing BWT files");
                if (pfb_write_bwt(ctx, args.output.c_str(), args.sa, args.rssa, args.stdout_ext.empty() ? NULL : args.stdout_ext.c_str())) die_ctx(ctx);
                if (args.doc_array && pfb_write_doc_array(ctx, args.output.c_str())) die_ctx(ctx);
                if (!args.marker_mps.empty() && pfb_write_marker_array(ctx, args.marker_mps.c_str(), (args.output + ".ma").c_str())) die_ctx(ctx);
            }
        }
        pfb_sizes sz;
        pfb_get_sizes(ctx, &sz);
        fprintf(stderr, "# easy cases: %lu, # hard cases: %lu\n", (unsigned long)sz.easy, (unsigned long)sz.hard);
        fprintf(stderr, "n: %lu\n", (unsigned long)sz.n);
        fprintf(stderr, "r: %lu\n", (unsigned long)sz.r);
        fprintf(stderr, "n/r: %.3f\n", (double)sz.n / (double)sz.r);
    }
    { SubTimer t("waiting for the writer thread"); if (pfb_wait_saves(ctx)) die_ctx(ctx); }
    if (args.verbose) {
        pfb_stats s;
        pfb_get_stats(ctx, &s);
        fprintf(stderr, "GPU\tnormalise %.3f trigger %.3f compact %.3f fingerprint %.3f dedup %.3f phrase_sort %.3f dict_build %.3f "
                        "parse_sa %.3f parse_bwt %.3f dict_sort %.3f groups %.3f emit %.3f runs %.3f fasta_records %.3f ms\n",
                s.normalise_ms, s.trigger_ms, s.compact_ms, s.fingerprint_ms, s.dedup_ms, s.phrase_sort_ms, s.dict_build_ms,
                s.parse_sa_ms, s.parse_bwt_ms, s.dict_sort_ms, s.groups_ms, s.emit_ms, s.runs_ms, s.ingest_ms);
    }
    // every output file is closed by now: leave without unmapping tens of GB of device slabs and pinned buffers
    // (the driver reclaims them with the process; pfb_destroy costs 0.2-0.3 s at this point)
    if (getenv("PFB_CLI_TIMES")) {
        std::chrono::duration<double> d = std::chrono::system_clock::now() - t_main;
        fprintf(stderr, "TIME\tmain() until all files are closed\t%.3fs\n", d.count());
    }
    fflush(stdout);
    fflush(stderr);
    _exit(0);
}
