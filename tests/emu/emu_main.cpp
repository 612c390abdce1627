// emu_main.cpp -- DEVELOPER LOGIC CHECK, TEST INFRASTRUCTURE ONLY.
// Compiles the device pipeline (pfbwt-f_b200/csrc/pipeline.cuh, engine.hpp) with
// -DPFB_EMU so that every launch_for / scan / sort runs as a serial host loop.
// It lets the `-m "not gpu"` tests check the pipeline's index arithmetic against
// the oracle on a box without a GPU.  It is NOT a fallback: nothing here is
// linked into libpfbwt_b200.so, which fails loudly without a CUDA device.
//
//   pfb_emu [-w W] [-p P] [-s] [-r] [--bits 32|64] [--non-acgt-to-a] [--parse-only] [-o prefix] in.fa
//   pfb_emu --merge [-w W] [-p P] [-s] [--bits 64] -o out pfx1 pfx2 ...
//   pfb_emu --marker-join <mps> <sa file or -> <out>          (mps_to_ma)
//   pfb_emu ... -s --marker-array <mps> ... in.fa              (also <prefix>.ma from the SA in memory)
#define PFB_EMU 1
#include "engine.hpp"

namespace pfb {
size_t &dev_bytes_live() { static size_t v = 0; return v; }
size_t &dev_bytes_peak() { static size_t v = 0; return v; }
}

int main(int argc, char **argv) {
    int w = 10, p = 100, sa = 0, rssa = 0, bits = 32, ntoa = 0, parse_only = 0, merge = 0, docs = 0, write_last = 0, bucketed = 0;
    std::string out, marker_mps;
    std::vector<std::string> in;
    if (argc == 5 && std::string(argv[1]) == "--marker-join") {
        pfb::Engine J;
        if (J.marker_join(argv[2], argv[3], argv[4])) { fprintf(stderr, "%s\n", J.err.c_str()); return 1; }
        return 0;
    }
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-w") w = atoi(argv[++i]);
        else if (a == "-p") p = atoi(argv[++i]);
        else if (a == "-s") sa = 1;
        else if (a == "-r") rssa = 1;
        else if (a == "--bits") bits = atoi(argv[++i]);
        else if (a == "--non-acgt-to-a") ntoa = 1;
        else if (a == "--parse-only") parse_only = 1;
        else if (a == "--print-docs") docs = 1;
        else if (a == "--merge") merge = 1;
        else if (a == "--write-last") write_last = 1;
        else if (a == "-m") bucketed = 1;                 // key-range buckets streamed to the files (PFB_BUCKETS / PFB_HBM_LIMIT)
        else if (a == "-o") out = argv[++i];
        else if (a == "--marker-array") marker_mps = argv[++i];
        else in.push_back(a);
    }
    if (in.empty()) { fprintf(stderr, "usage: see source header\n"); return 2; }
    pfb::Engine E;
    E.prm.w = w; E.prm.p = (uint64_t)p; E.prm.bits = bits; E.prm.non_acgt_to_a = ntoa; E.prm.get_sai = sa || rssa;
    E.store_docs = docs;
    if (merge) {
        if (out.empty()) out = "out";
        for (auto &pfx : in) {
            pfb::Engine R;
            R.prm = E.prm; R.store_docs = docs;
            if (R.load_parse(pfx, false, false)) { fprintf(stderr, "%s\n", R.err.c_str()); return 1; }
            if (E.append_parse_of(R)) { fprintf(stderr, "%s\n", E.err.c_str()); return 1; }
        }
        if (E.finalize() || E.do_parse_bwt() || E.save_parse(out, true, sa)) { fprintf(stderr, "%s\n", E.err.c_str()); return 1; }
        return 0;
    }
    if (out.empty()) out = in[0];
    for (auto &f : in)                      // several inputs = one record list (PfParser::add_fasta called once per file)
        if (E.add_fasta(f)) { fprintf(stderr, "%s\n", E.err.c_str()); return 1; }
    fprintf(stderr, "ingest: %s\n", E.ingest_files == in.size() ? "device" : E.ingest_files ? "mixed" : "host");
    // the order of run_parser (src/pfbwt-f.cpp:221-236): parse files first, parse-BWT files after bwt_of_parse
    if (E.finalize() || E.save_parse(out, false, false) || (write_last && E.save_last_sai(out, true, true)) || E.do_parse_bwt() ||
        E.save_parse(out, true, sa || rssa, true)) { fprintf(stderr, "%s\n", E.err.c_str()); return 1; }
    if (parse_only) return 0;
    if (bucketed ? E.build_to_files(out, sa, rssa, 0) : (E.build(sa, rssa) || E.write_bwt_files(out, sa, rssa, ""))) { fprintf(stderr, "%s\n", E.err.c_str()); return 1; }
    if (bucketed) fprintf(stderr, "buckets: %lu\n", (unsigned long)E.stream_buckets);
    if (!marker_mps.empty() && E.write_marker_array(marker_mps, out + ".ma")) { fprintf(stderr, "%s\n", E.err.c_str()); return 1; }
    fprintf(stderr, "# easy cases: %lu, # hard cases: %lu\nn: %lu\nr: %lu\nrounds: phrase %u parse_sa %u dict %u\n",
            (unsigned long)E.bwt.easy, (unsigned long)E.bwt.hard, (unsigned long)E.bwt.n, (unsigned long)E.bwt.r,
            E.tm.phrase_sort_rounds, E.tm.parse_sa_rounds, E.tm.dict_sort_rounds);
    return 0;
}

// ---- multi-rank logic check (tests/test_shard_gloo.py): one call = one rank of a job whose exchanges go
// through the caller's host all-gather (torch.distributed / gloo).  Writes this rank's slices of
// <out>.bwt/.sa/.ssa/.esa at their offsets and, on rank 0, the merged parse files.
// range_end > range_begin: the rank takes that byte range of the (common) FASTA file instead of a file of its own
extern "C" int pfb_emu_shard_run(int rank, int world, pfb_allgatherv_fn fn, void *user, const char *fasta, int w, int p, int bits,
                                 int sa, int rssa, const char *out, unsigned long long *info /* 8 */,
                                 unsigned long long range_begin, unsigned long long range_end) {
    pfb::Engine E;
    E.prm.w = w; E.prm.p = (uint64_t)p; E.prm.bits = bits; E.prm.get_sai = sa || rssa;
    pfb::HostCallComm *hc = new pfb::HostCallComm();
    hc->rank = rank; hc->world = world; hc->fn = fn; hc->user = user;
    E.comm.reset(hc);
    // a pre-parsed part (<fasta>.dict and <fasta>.parse exist) is loaded instead of parsed: src/merge_pfp.cpp:100-113
    int rc;
    const std::string pfx(fasta);
    if (!(range_end > range_begin) && access((pfx + ".dict").c_str(), R_OK) == 0 && access((pfx + ".parse").c_str(), R_OK) == 0) {
        pfb::Engine R;
        R.prm = E.prm;
        rc = R.load_parse(pfx, false, false);
        if (rc) E.err = R.err; else rc = E.append_parse_of(R);
    } else rc = range_end > range_begin ? E.add_fasta_range(fasta, range_begin, range_end) : E.add_fasta(fasta);
    if (rc) fprintf(stderr, "[emu rank %d] %s\n", rank, E.err.c_str());
    // a failed load still has to take part in the collective finalize
    rc = E.finalize();
    if (!rc) rc = E.do_parse_bwt();
    if (!rc && rank == 0) rc = E.save_parse(out, true, sa || rssa);
    if (!rc) rc = E.build(sa, rssa);
    if (!rc) rc = E.write_bwt_files(out, sa, rssa, "");
    if (rc) { fprintf(stderr, "[emu rank %d] %s\n", rank, E.err.c_str()); return rc; }
    if (info) {
        info[0] = E.bwt.off; info[1] = E.bwt.len; info[2] = E.bwt.ssa_off; info[3] = E.bwt.ssa_cnt;
        info[4] = E.bwt.esa_off; info[5] = E.bwt.esa_cnt; info[6] = E.bwt.r; info[7] = E.parse.P;
    }
    return 0;
}
