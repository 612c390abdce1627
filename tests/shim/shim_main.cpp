// shim_main.cpp -- a driver written the way a user of the reference's C++ API writes one
// (parser -> save_parser -> bwt_of_parse -> PrefixFreeBWT::generate_bwt_lcp with a per-position
// writer callback; cf. reference src/pfbwt-f.cpp:212-349), compiled against include/pfbwtf_shim.hpp
// instead of the reference headers.  tests/test_shim.py compares every file it writes with the
// unmodified reference binary.
//   shim_demo [-s] [-r] [-w W] [-p P] -o out a.fa [b.fa ...]      parse each file, += them, build
#include "pfbwtf_shim.hpp"

int main(int argc, char **argv) {
    bool sa = false, rssa = false;
    size_t w = 10, p = 100;
    std::string out;
    std::vector<std::string> in;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-s") sa = true;
        else if (a == "-r") rssa = true;
        else if (a == "-w") w = (size_t)atoi(argv[++i]);
        else if (a == "-p") p = (size_t)atoi(argv[++i]);
        else if (a == "-o") out = argv[++i];
        else in.push_back(a);
    }
    if (in.empty() || out.empty()) { fprintf(stderr, "usage: shim_demo [-s] [-r] [-w W] [-p P] -o out in.fa...\n"); return 2; }
    pfbwtf::PfParserParams params(w, p, sa || rssa, false, false, false, false);
    using parser_t = pfbwtf::PfParser<>;
    parser_t parser(params);
    if (in.size() == 1) {
        parser.add_fasta(in[0]);
    } else {                                   // merge_pfp style: one parser per file, then +=
        for (auto &f : in) {
            parser_t part = parse_from_fasta(f, params);
            parser += part;
        }
    }
    parser.finalize();
    const size_t n = parser.get_n();
    save_parser(parser, out);
    parser.bwt_of_parse([&](const std::vector<char> &bwlast, const std::vector<parser_t::UIntType> &ilist,
                            const std::vector<parser_t::UIntType> &bwsai) { save_parse_bwt(bwlast, ilist, bwsai, out); });
    // round trip through the files, like --pfbwt-only
    pfbwtf::PrefixFreeBWTParams bp;
    bp.prefix = out; bp.w = w; bp.sa = sa; bp.rssa = rssa;
    using pfbwt_t = pfbwtf::PrefixFreeBWT<VecFileSource, VecFileSinkPrivate>;
    pfbwt_t builder(bp);
    FILE *bwt_fp = fopen((out + ".bwt").c_str(), "wb");
    FILE *sa_fp = sa ? fopen((out + ".sa").c_str(), "wb") : nullptr;
    FILE *ssa_fp = rssa ? fopen((out + ".ssa").c_str(), "wb") : nullptr;
    FILE *esa_fp = rssa ? fopen((out + ".esa").c_str(), "wb") : nullptr;
    size_t r = 0;
    pfbwt_t::UIntType i = 0, pi = 0, psa = 0;
    builder.generate_bwt_lcp([&](const pfbwtf::out_fn_arg a) {
        fputc(a.bwtc, bwt_fp);
        pfbwt_t::UIntType x = i ? a.sa : (pfbwt_t::UIntType)n;
        if (sa) fwrite(&x, sizeof(x), 1, sa_fp);
        if (a.bwtc != a.pbwtc) {
            ++r;
            if (rssa) {
                fwrite(&i, sizeof(i), 1, ssa_fp); fwrite(&x, sizeof(x), 1, ssa_fp);
                if (i) { pfbwt_t::UIntType y = pi ? psa : (pfbwt_t::UIntType)n; fwrite(&pi, sizeof(pi), 1, esa_fp); fwrite(&y, sizeof(y), 1, esa_fp); }
            }
        }
        pi = i; psa = a.sa; ++i;
    });
    if (rssa) { fwrite(&pi, sizeof(pi), 1, esa_fp); fwrite(&psa, sizeof(psa), 1, esa_fp); }
    fclose(bwt_fp);
    if (sa_fp) fclose(sa_fp);
    if (ssa_fp) { fclose(ssa_fp); fclose(esa_fp); }
    // getters the reference's tests use (tests/test_parser.cpp:108-234)
    parser_t again = load_parser(out, params);
    if (again.get_parse_ranks() != parser.get_parse_ranks() || again.get_sorted_phrases().size() != parser.get_sorted_phrases().size()) {
        fprintf(stderr, "load_parser(save_parser(p)) != p\n");
        return 1;
    }
    fprintf(stderr, "n: %lu\nr: %lu\nphrases: %lu words: %lu easy: %lu hard: %lu\n", (unsigned long)n, (unsigned long)r,
            (unsigned long)parser.get_parse_size(), (unsigned long)parser.get_freqs().size(), (unsigned long)builder.easy_cases,
            (unsigned long)builder.hard_cases);
    return 0;
}
