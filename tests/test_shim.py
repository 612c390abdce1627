"""Source-level drop-in (include/pfbwtf_shim.hpp): a driver written against the reference's C++ class API
(PfParser / PrefixFreeBWT / save_parser / load_parser, per-position output callback) compiled against the
shim must write the same files as the unmodified reference binary."""
import os
import pytest
from pfp_util import ROOT, ref_bin, write_fasta, pangenome_records, random_bases, run, exts_for, assert_same_files

BIN = os.path.join(ROOT, "bin")


def test_shim_header_names_the_reference_api():
    """CPU check: the shim declares every class / function of SURVEY.md 8(b)'s C++ API row"""
    src = open(os.path.join(ROOT, "include", "pfbwtf_shim.hpp")).read()
    for name in ["struct PfParserParams", "struct PfParser", "size_t add_fasta(std::string", "void finalize()", "void bwt_of_parse(",
                 "operator+=", "operator==", "get_n()", "get_pos()", "get_parse_size()", "get_occs()", "get_sai()", "get_last()",
                 "get_freqs()", "get_parse()", "get_parse_ranks()", "get_sorted_phrases()", "get_doc_starts()", "get_doc_names()",
                 "get_params()", "get_ntab()", "struct PrefixFreeBWTParams", "class PrefixFreeBWT", "void generate_bwt_lcp(",
                 "struct out_fn_arg", "save_parser(", "load_parser(", "parse_from_fasta(", "save_parse_bwt("]:
        assert name in src, name


@pytest.mark.gpu
@pytest.mark.parametrize("bits,flags,nfiles", [(32, ["-s", "-r"], 1), (64, ["-s"], 3), (32, [], 2)])
def test_shim_driver_matches_reference(bits, flags, nfiles, datadir):
    if ref_bin("pfbwt-f") is None:
        pytest.skip("reference binaries not built")
    recs = pangenome_records(120_000, 6, 100, 41) + [("tail", random_bases(50_000, 42))]
    k = (len(recs) + nfiles - 1) // nfiles
    fas = []
    for g in range(nfiles):
        fa = os.path.join(datadir, f"shim{bits}_{nfiles}_{g}.fa")
        write_fasta(fa, recs[g * k:(g + 1) * k], line_width=60)
        fas.append(fa)
    whole = os.path.join(datadir, f"shim{bits}_{nfiles}_whole.fa")
    write_fasta(whole, recs, line_width=60)
    a, b = os.path.join(datadir, f"shim{bits}_{nfiles}.ref"), os.path.join(datadir, f"shim{bits}_{nfiles}.out")
    run([ref_bin("pfbwt-f64" if bits == 64 else "pfbwt-f")] + flags + ["-o", a, whole])
    run([os.path.join(BIN, "shim_demo64" if bits == 64 else "shim_demo")] + flags + ["-o", b] + fas)
    assert_same_files(a, b, exts_for("-s" in flags, "-r" in flags))
