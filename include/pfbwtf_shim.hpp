// pfbwtf_shim.hpp -- source-level drop-in for the reference's header-level C++ API on the hot path,
// implemented over the C ABI of libpfbwt_b200.so (include/pfbwt_b200.h).  Code written against
//   pfbwtf::PfParserParams / PfParser<>            (reference include/pfparser.hpp:30-44, 72-603)
//   pfbwtf::PrefixFreeBWTParams / PrefixFreeBWT    (reference include/pfbwt.hpp:36-52, 56-271)
//   save_parser / load_parser / parse_from_fasta / save_parse_bwt   (reference include/pfbwt_io.hpp:211-262,
//                                                                      src/pfbwt-f.cpp:229-243)
// compiles against this header unchanged: same names, same signatures, same meaning of every
// argument, same error behaviour (message on stderr + exit(1), like die() in src/utils.c:13-17).
// What differs is where the work happens: every method forwards to the device pipeline; the
// per-position callback of generate_bwt_lcp is fed from bulk device->host copies.
//
// Define M64 to 1 before including for the pfbwt-f64 flavour (gsa/gsacak.h:40-58).
// Not reproduced: out_fn_arg::dif (the per-position Difficulty is always EASY1; only the totals
// are available, see PrefixFreeBWT::easy_cases/hard_cases), get_ntab() (always empty, as in the
// reference where the code that fills it is commented out, include/pfparser.hpp:338-359).
#ifndef PFBWTF_SHIM_HPP
#define PFBWTF_SHIM_HPP
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <utility>
#include <vector>
#include <unistd.h>
#include "pfbwt_b200.h"

#ifndef M64
#define M64 0
#endif
#if M64
typedef int64_t int_t;
typedef uint64_t uint_t;
#else
typedef int32_t int_t;
typedef uint32_t uint_t;
#endif
typedef uint32_t int_text;

// container tags accepted (and ignored) as PrefixFreeBWT's template arguments: the workspace lives in HBM
template <typename T, typename...> class VecFileSource {};
template <typename T, typename...> class VecFileSinkPrivate {};
template <typename T, typename...> class MMapFileSource {};
template <typename T, typename...> class MMapFileSink {};

namespace pfbwtf {

struct WangHash {};   // the hasher is fixed on the device (reference include/hash.hpp:23-43)

namespace detail {
inline void die(pfb_ctx *c, int rc) {
    if (rc == 0) return;
    fprintf(stderr, "%s\n", c ? pfb_last_error(c) : pfb_create_error());
    exit(1);
}
inline pfb_ctx *make_ctx() {
    int dev = 0;
    if (const char *e = getenv("PFB_DEVICE")) dev = atoi(e);
    pfb_ctx *c = pfb_create(dev, M64 ? 64 : 32);
    if (!c) die(nullptr, -1);
    return c;
}
}  // namespace detail

struct PfParserParams {
    constexpr PfParserParams() {}
    constexpr PfParserParams(size_t wsize, size_t pmod, bool s, bool d, bool v, bool trim, bool ntoa)
        : w(wsize), p(pmod), get_sai(s), store_docs(d), verbose(v), trim_non_acgt(trim), non_acgt_to_a(ntoa) {}
    size_t w = 10;
    size_t p = 100;
    bool get_sai = false;
    bool store_docs = false;
    bool verbose = false;
    bool trim_non_acgt = false;
    bool non_acgt_to_a = false;
};

template <typename T>
struct Freq {
    Freq() {}
    Freq(T x) : n(x) {}
    Freq(T x, T y) : n(x), r(y) {}
    T n = 0;
    T r = 0;
    bool operator==(const Freq<T> &rhs) const { return rhs.n == n && rhs.r == r; }
    bool operator!=(const Freq<T> &rhs) const { return !(*this == rhs); }
};
struct ntab_entry { size_t pos = 0; size_t l = 0; void clear() { pos = 0; l = 0; } };
template <typename UIntType> using FreqMap = std::map<std::string, Freq<UIntType>>;

template <typename Hasher = WangHash>
struct PfParser {
    using UIntType = uint_t;
    using IntType = int_text;

    PfParser() : PfParser(PfParserParams()) {}
    PfParser(PfParserParams p) : params_(p) { init(); }
    // from a serialized dictionary + parse ranks (reference include/pfparser.hpp:86-132)
    PfParser(PfParserParams params, const std::vector<std::string> &sorted_phrases, const std::vector<IntType> &parse_ranks,
             const std::vector<UIntType> &doc_starts = std::vector<UIntType>(),
             const std::vector<std::string> &doc_names = std::vector<std::string>())
        : params_(params), doc_starts_(doc_starts), doc_names_(doc_names) {
        init();
        // hand the two images to the library through its own file formats (Appendix A of SURVEY.md)
        char tmpl[] = "/tmp/pfbshimXXXXXX";
        if (!mkdtemp(tmpl)) { perror("mkdtemp"); exit(1); }
        std::string pfx = std::string(tmpl) + "/p";
        FILE *f = fopen((pfx + ".dict").c_str(), "wb");
        for (const auto &s : sorted_phrases) { fwrite(s.data(), 1, s.size(), f); fputc(1, f); }
        fputc(0, f);
        fclose(f);
        f = fopen((pfx + ".parse").c_str(), "wb");
        fwrite(parse_ranks.data(), sizeof(IntType), parse_ranks.size(), f);
        fclose(f);
        int rc = pfb_load_parse(ctx_, pfx.c_str(), 0, 0);
        unlink((pfx + ".dict").c_str()); unlink((pfx + ".parse").c_str()); rmdir(tmpl);
        detail::die(ctx_, rc);
        loaded_ = true;
    }
    PfParser(const PfParser &rhs) : params_(rhs.params_), doc_starts_(rhs.doc_starts_), doc_names_(rhs.doc_names_) {
        init();
        if (rhs.has_content()) {
            if (!rhs.finalized()) { fprintf(stderr, "copying a parser that holds unparsed records is not supported: finalize() it first\n"); exit(1); }
            detail::die(ctx_, pfb_append(ctx_, rhs.ctx_));
            finalize();
        }
    }
    PfParser(PfParser &&rhs) noexcept { steal(rhs); }
    PfParser &operator=(PfParser rhs) { if (ctx_) pfb_destroy(ctx_); steal(rhs); return *this; }
    ~PfParser() { if (ctx_) pfb_destroy(ctx_); }

    // reference include/pfparser.hpp:299-369: returns the parser's position (bases so far + 1)
    size_t add_fasta(std::string fasta_fname) {
        detail::die(ctx_, pfb_add_fasta(ctx_, fasta_fname.c_str()));
        dirty();
        return (size_t)pfb_get_pending_bases(ctx_) + 1;
    }
    // reference :484-492
    void finalize() { detail::die(ctx_, pfb_finalize(ctx_)); dirty(); finalized_ = true; }
    // reference :194-263.  As there, the result has to be finalize()d by the caller.
    PfParser &operator+=(const PfParser &rhs) {
        if (rhs.params_.w != params_.w) { fprintf(stderr, "invalid w\n"); exit(1); }
        if (rhs.params_.p != params_.p) { fprintf(stderr, "invalid p\n"); exit(1); }
        const_cast<PfParser &>(rhs).ensure_parsed();
        detail::die(ctx_, pfb_append(ctx_, rhs.ctx_));
        doc_starts_.clear();                                                // the library keeps the rebased offsets (pfb_get_docs)
        for (auto &nm : rhs.get_doc_names()) doc_names_.push_back(nm);
        dirty();
        finalized_ = false; loaded_ = false;
        return *this;
    }
    // reference :265-297
    bool operator==(const PfParser &rhs) {
        ensure_parsed(); const_cast<PfParser &>(rhs).ensure_parsed();
        if (get_n() != rhs.get_n() || get_parse_size() != rhs.get_parse_size()) return false;
        if (get_parse_ranks() != rhs.get_parse_ranks()) return false;
        if (dict_image() != rhs.dict_image() || get_occs() != rhs.get_occs()) return false;
        if (get_last() != rhs.get_last()) return false;
        if (params_.get_sai && get_sai() != rhs.get_sai()) return false;
        return true;
    }
    // reference :379-467: out_fn(bwlast, ilist, bwsai)
    template <typename OutFn>
    void bwt_of_parse(OutFn out_fn) {
        ensure_parsed();
        detail::die(ctx_, pfb_parse_bwt(ctx_));
        pfb_sizes s = sizes();
        std::vector<char> bwlast(s.n_phrases + 1);
        std::vector<UIntType> ilist(s.n_phrases + 1), bwsai;
        detail::die(ctx_, pfb_copy_bwlast(ctx_, (uint8_t *)bwlast.data()));
        detail::die(ctx_, pfb_copy_ilist(ctx_, ilist.data()));
        if (params_.get_sai) { bwsai.resize(s.n_phrases + 1); detail::die(ctx_, pfb_copy_bwsai(ctx_, bwsai.data())); }
        out_fn(bwlast, ilist, bwsai);
    }
    template <typename OutFn>
    void bwt_of_parse(OutFn out_fn, bool /*get_sai*/) { bwt_of_parse(out_fn); }

    // ---- getters (reference :469-544); host copies are made on first use and kept until the parser changes ----
    size_t get_parse_size() const { return (size_t)sizes().n_phrases; }
    size_t get_n() const { return (size_t)(finalized() ? sizes().n : pfb_get_pending_bases(ctx_)); }
    size_t get_pos() const { return get_n() + (finalized() ? params_.w : 1); }
    const std::vector<UIntType> get_occs() const {
        std::vector<UIntType> v(sizes().n_words);
        detail::die(ctx_, pfb_copy_occ(ctx_, v.data()));
        return v;
    }
    const std::vector<UIntType> &get_sai() const {
        if (sai_.empty() && params_.get_sai) { sai_.resize(sizes().n_phrases); detail::die(ctx_, pfb_copy_sai(ctx_, sai_.data())); }
        return sai_;
    }
    const std::vector<char> &get_last() const {
        if (last_.empty()) { last_.resize(sizes().n_phrases); detail::die(ctx_, pfb_copy_last(ctx_, (uint8_t *)last_.data())); }
        return last_;
    }
    const std::vector<int_text> &get_parse_ranks() const {
        if (ranks_.empty()) { ranks_.resize(sizes().n_phrases); detail::die(ctx_, pfb_copy_parse(ctx_, ranks_.data())); }
        return ranks_;
    }
    // NUL-terminated C strings in dictionary order, like the reference's pointers into its map keys
    const std::vector<const char *> &get_sorted_phrases() const {
        if (phrases_.empty()) {
            dict_ = dict_image();
            const char *start = dict_.data();
            for (size_t i = 0; i + 1 < dict_.size(); ++i)
                if (dict_[i] == 1) { dict_[i] = 0; phrases_.push_back(start); start = dict_.data() + i + 1; }
        }
        return phrases_;
    }
    const std::vector<const char *> &get_parse() const {
        if (parse_.empty()) { const auto &ph = get_sorted_phrases(); for (auto r : get_parse_ranks()) parse_.push_back(ph[r - 1]); }
        return parse_;
    }
    const FreqMap<UIntType> &get_freqs() const {
        if (freqs_.empty()) {
            const auto &ph = get_sorted_phrases();
            const auto occ = get_occs();
            for (size_t i = 0; i < ph.size(); ++i) freqs_[ph[i]] = Freq<UIntType>(occ[i], (UIntType)(i + 1));
        }
        return freqs_;
    }
    const std::vector<ntab_entry> &get_ntab() const { return ntab_; }
    const std::vector<UIntType> &get_doc_starts() const {
        if (doc_starts_.empty()) {
            int64_t k = pfb_get_docs(ctx_, nullptr);
            if (k > 0) { std::vector<uint64_t> t((size_t)k); pfb_get_docs(ctx_, t.data()); doc_starts_.assign(t.begin(), t.end()); }
        }
        return doc_starts_;
    }
    const std::vector<std::string> &get_doc_names() const { return doc_names_; }
    const PfParserParams get_params() const { return params_; }

    pfb_ctx *native_handle() const { return ctx_; }      // for the bulk (non-callback) entry points of the C ABI
    bool finalized() const { return finalized_ || loaded_; }
    void finalize_loaded() { loaded_ = true; dirty(); }    // the context was filled by pfb_load_parse

   private:
    void init() {
        if (params_.w >= 32) { fprintf(stderr, "window size w must be < 32!\n"); exit(1); }     // reference :371-376
        ctx_ = detail::make_ctx();
        pfb_parse_params pp;
        pp.w = (uint32_t)params_.w; pp.p = params_.p; pp.get_sai = params_.get_sai; pp.store_docs = params_.store_docs;
        pp.verbose = params_.verbose; pp.trim_non_acgt = params_.trim_non_acgt; pp.non_acgt_to_a = params_.non_acgt_to_a;
        detail::die(ctx_, pfb_set_params(ctx_, &pp));
    }
    void steal(PfParser &rhs) {
        ctx_ = rhs.ctx_; rhs.ctx_ = nullptr;
        params_ = rhs.params_; finalized_ = rhs.finalized_; loaded_ = rhs.loaded_;
        doc_starts_ = std::move(rhs.doc_starts_); doc_names_ = std::move(rhs.doc_names_);
        dirty();
    }
    bool has_content() const { return loaded_ || finalized_ || pfb_get_pending_bases(ctx_) > 0; }
    void ensure_parsed() { if (!finalized()) finalize(); }
    void dirty() { sai_.clear(); last_.clear(); ranks_.clear(); phrases_.clear(); parse_.clear(); freqs_.clear(); dict_.clear(); }
    pfb_sizes sizes() const { pfb_sizes s; pfb_get_sizes(ctx_, &s); return s; }
    std::vector<char> dict_image() const {
        std::vector<char> d(sizes().dict_bytes);
        detail::die(ctx_, pfb_copy_dict(ctx_, (uint8_t *)d.data()));
        return d;
    }

    pfb_ctx *ctx_ = nullptr;
    PfParserParams params_;
    bool finalized_ = false, loaded_ = false;
    mutable std::vector<UIntType> sai_, doc_starts_;
    mutable std::vector<char> last_, dict_;
    mutable std::vector<int_text> ranks_;
    mutable std::vector<const char *> phrases_, parse_;
    mutable FreqMap<UIntType> freqs_;
    std::vector<ntab_entry> ntab_;
    std::vector<std::string> doc_names_;
};

// ---- BWT build (reference include/pfbwt.hpp) ---------------------------------------------------------
enum class RunType { OTHER, START, END };
enum class Difficulty { EASY1, EASY2, HARD };

struct out_fn_arg {
    out_fn_arg(uint_t p, uint_t s, uint8_t pc, uint8_t c, Difficulty d = Difficulty::EASY1) : pos(p), sa(s), pbwtc(pc), bwtc(c), dif(d) {}
    uint_t pos;
    uint_t sa;
    uint8_t pbwtc;
    uint8_t bwtc;
    Difficulty dif;
};

struct PrefixFreeBWTParams {
    std::string prefix;
    size_t w;
    bool sa = false;
    bool rssa = false;
    bool verb = false;
};

template <template <typename, typename...> class ReadConType, template <typename, typename...> class WriteConType>
class PrefixFreeBWT {
   public:
    using UIntType = uint_t;
    using IntType = int_t;

    // reference include/pfbwt.hpp:64-81: loads <prefix>.dict/.bwlast/.ilist(/.bwsai)
    PrefixFreeBWT(PrefixFreeBWTParams args) : args_(args) {
        ctx_ = detail::make_ctx();
        pfb_parse_params pp;
        memset(&pp, 0, sizeof(pp));
        pp.w = (uint32_t)args.w; pp.p = 100; pp.get_sai = args.sa || args.rssa; pp.verbose = args.verb;
        detail::die(ctx_, pfb_set_params(ctx_, &pp));
        detail::die(ctx_, pfb_load_parse(ctx_, args.prefix.c_str(), 1, args.sa || args.rssa));
        if (args.verb) fprintf(stderr, "loaded files\n");
    }
    // build from a parser that is still in HBM (no files in between)
    template <typename H>
    PrefixFreeBWT(PfParser<H> &parser, PrefixFreeBWTParams args) : args_(args), ctx_(parser.native_handle()), borrowed_(true) {}
    PrefixFreeBWT(const PrefixFreeBWT &) = delete;
    ~PrefixFreeBWT() { if (ctx_ && !borrowed_) pfb_destroy(ctx_); }

    // reference :96-194: out_fn(out_fn_arg) once per BWT position, in BWT order.  pos/sa are filled only
    // with SA output (UPDATE_SA vs UPDATE_BWT, :87-94); pbwtc is the previous position's BWT byte (0 first).
    template <typename Fn>
    void generate_bwt_lcp(Fn out_fn) {
        const bool any_sa = args_.sa || args_.rssa;
        detail::die(ctx_, pfb_build_bwt(ctx_, args_.sa, args_.rssa));
        pfb_sizes s;
        pfb_get_sizes(ctx_, &s);
        easy_cases = (size_t)s.easy; hard_cases = (size_t)s.hard;
        std::vector<uint8_t> bwt(s.n + 1);
        detail::die(ctx_, pfb_copy_bwt(ctx_, bwt.data()));
        std::vector<UIntType> sa;
        if (any_sa) { sa.resize(s.n + 1); detail::die(ctx_, pfb_copy_sa(ctx_, sa.data())); }
        uint8_t pbwtc = 0;
        for (size_t i = 0; i <= s.n; ++i) {
            if (any_sa) out_fn(out_fn_arg((UIntType)i, sa[i], pbwtc, bwt[i]));
            else out_fn(out_fn_arg(0, 0, pbwtc, bwt[i]));
            pbwtc = bwt[i];
        }
        if (args_.verb) fprintf(stderr, "# easy cases: %lu, # hard cases: %lu\n", (unsigned long)easy_cases, (unsigned long)hard_cases);
    }
    pfb_ctx *native_handle() const { return ctx_; }
    size_t easy_cases = 0, hard_cases = 0;      // reference :188

   private:
    PrefixFreeBWTParams args_;
    pfb_ctx *ctx_ = nullptr;
    bool borrowed_ = false;
};

}  // namespace pfbwtf

// ---- free functions of include/pfbwt_io.hpp used on the path -------------------------------------------
/* reference include/pfbwt_io.hpp:234-249 */
inline void save_parser(const pfbwtf::PfParser<> &parser, std::string prefix) {
    pfbwtf::detail::die(parser.native_handle(), pfb_save_parse(parser.native_handle(), prefix.c_str(), 0, 0));
}
/* reference include/pfbwt_io.hpp:211-222 */
inline pfbwtf::PfParser<> load_parser(std::string prefix, pfbwtf::PfParserParams p) {
    pfbwtf::PfParser<> parser(p);
    pfbwtf::detail::die(parser.native_handle(), pfb_load_parse(parser.native_handle(), prefix.c_str(), 0, 0));
    parser.finalize_loaded();
    return parser;
}
/* reference include/pfbwt_io.hpp:251-262 */
inline pfbwtf::PfParser<> parse_from_fasta(std::string fasta_fname, pfbwtf::PfParserParams p) {
    pfbwtf::PfParser<> parser(p);
    parser.add_fasta(fasta_fname);
    parser.finalize();
    return parser;
}
/* reference src/pfbwt-f.cpp:229-243: the three parse-BWT files */
template <typename UIntType>
inline void save_parse_bwt(const std::vector<char> &bwlast, const std::vector<UIntType> &ilist, const std::vector<UIntType> &bwsai,
                           std::string prefix) {
    auto dump = [&](const std::string &ext, const void *p, size_t bytes) {
        FILE *f = fopen((prefix + "." + ext).c_str(), "wb");
        if (!f) { perror("open() file failed"); exit(1); }
        if (bytes && fwrite(p, 1, bytes, f) != bytes) { perror("fwrite() file failed"); exit(1); }
        fclose(f);
    };
    dump("bwlast", bwlast.data(), bwlast.size());
    dump("ilist", ilist.data(), ilist.size() * sizeof(UIntType));
    if (!bwsai.empty()) dump("bwsai", bwsai.data(), bwsai.size() * sizeof(UIntType));
}
#endif
