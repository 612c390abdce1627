"""CPU tests (-m "not gpu") of the HOST LOGIC of the device pipeline: tests/emu/pfb_emu compiles
pfbwt-f_b200/csrc/{pipeline.cuh,engine.hpp} with every launch/scan/sort executed as a serial loop
(developer logic check, test infrastructure only - never linked into the library) and must write
files identical to the oracle.  The real parity tests are the `-m gpu` ones."""
import os
import subprocess
import pytest
import numpy as np
from pfp_util import ROOT, write_fasta, random_bases, pangenome_records, nrun_records, run, exts_for, assert_same_files

EMU = os.path.join(ROOT, "tests", "emu", "pfb_emu")


@pytest.fixture(scope="module")
def emu():
    subprocess.run(["make", "-C", ROOT, "emu"], check=True, stdout=subprocess.DEVNULL)
    return EMU


CASES = [
    ("e_rand", lambda: [("r", random_bases(150_000, 2))], 32, ["-s", "-r"]),
    ("e_pan", lambda: pangenome_records(60_000, 6, 60, 8), 64, ["-s", "-r"]),
    ("e_pan_nosa", lambda: pangenome_records(60_000, 6, 60, 8), 32, []),
    ("e_nrun", lambda: nrun_records(2, 50_000, 38), 64, ["-s", "--non-acgt-to-a"]),
    ("e_w3p5", lambda: [("a", random_bases(20_000, 3)), ("b", random_bases(0, 4)), ("c", random_bases(7, 5))], 32,
     ["-s", "-r", "-w", "3", "-p", "5"]),
    # S[0..e1] is also a whole phrase elsewhere: the reference emits EndOfWord (0x01) bytes into the BWT
    # for the full word grouped behind the first word's suffix (include/pfbwt.hpp:136-141)
    ("e_w4p7", lambda: [("a", random_bases(50_000, 3)), ("b", random_bases(1, 4)), ("c", random_bases(30_000, 5))], 32,
     ["-s", "-r", "-w", "4", "-p", "7"]),
    ("e_w4p7_nosa", lambda: [("a", random_bases(50_000, 3)), ("b", random_bases(1, 4)), ("c", random_bases(30_000, 5))], 32,
     ["-w", "4", "-p", "7"]),
]


@pytest.mark.parametrize("name,gen,bits,flags", CASES, ids=[c[0] for c in CASES])
def test_emu_matches_oracle(name, gen, bits, flags, emu, ref_bins, datadir):
    fa = os.path.join(datadir, name + ".fa")
    write_fasta(fa, gen(), line_width=50)
    a, b = os.path.join(datadir, name + ".port"), os.path.join(datadir, name + ".emu")
    run([ref_bins["pfp_oracle"], "--bits", str(bits)] + flags + ["-o", a, fa])
    run([emu, "--bits", str(bits)] + flags + ["-o", b, fa])
    assert_same_files(a, b, exts_for("-s" in flags, "-r" in flags))


def test_emu_merge(emu, ref_bins, datadir):
    recs = pangenome_records(50_000, 6, 50, 77)
    parts = []
    for g in range(3):
        p = os.path.join(datadir, f"epart{g}.fa")
        write_fasta(p, recs[2 * g:2 * g + 2], line_width=60)
        run([emu, "--bits", "64", "--parse-only", "-s", p])
        parts.append(p)
    allfa = os.path.join(datadir, "eall.fa")
    write_fasta(allfa, recs, line_width=60)
    mono = os.path.join(datadir, "emono")
    run([ref_bins["pfp_oracle"], "--bits", "64", "--parse-only", "-s", "-o", mono, allfa])
    merged = os.path.join(datadir, "emerged")
    run([emu, "--merge", "--bits", "64", "-s", "-o", merged] + parts)
    assert_same_files(mono, merged, ["dict", "occ", "parse", "n", "bwlast", "ilist", "bwsai"])


@pytest.mark.parametrize("ext_rounds", ["0", "1", "100"])
def test_emu_tie_refinement_paths(ext_rounds, emu, ref_bins, datadir):
    """the dictionary suffix sort refines ties by symbol-extension rounds and/or prefix doubling;
    the PFB_EXT_ROUNDS test knob forces doubling only (0), mixed (1) and extension only (100)."""
    fa = os.path.join(datadir, "e_paths.fa")
    recs = pangenome_records(40_000, 5, 40, 9) + nrun_records(1, 20_000, 11)
    write_fasta(fa, recs, line_width=70)
    a, b = os.path.join(datadir, "e_paths.port"), os.path.join(datadir, "e_paths.emu" + ext_rounds)
    flags = ["-s", "-r", "--non-acgt-to-a"]
    run([ref_bins["pfp_oracle"], "--bits", "64"] + flags + ["-o", a, fa])
    env = dict(os.environ, PFB_EXT_ROUNDS=ext_rounds)
    subprocess.run([emu, "--bits", "64"] + flags + ["-o", b, fa], check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    assert_same_files(a, b, exts_for(True, True))


@pytest.mark.parametrize("pivot,key_syms", [("1", "19"), ("1", "3"), ("0", "3")])
def test_emu_pivot_rounds(pivot, key_syms, emu, ref_bins, datadir):
    """long ties (variant phrases of a pangenome) are refined by pivot rounds: every member of a tie group is
    compared with the group's first member and the group is ordered in one sort (build.cuh).  PFB_PIVOT=1
    forces them from the first round, PFB_KEY_SYMS=3 makes nearly every suffix start out tied; the N-run
    record exercises the comparison cap (poly-A) and the hand-over to prefix doubling."""
    fa = os.path.join(datadir, "e_pivot.fa")
    recs = pangenome_records(30_000, 8, 120, 19) + nrun_records(1, 30_000, 12, big_run=9000)
    write_fasta(fa, recs, line_width=70)
    a, b = os.path.join(datadir, "e_pivot.port"), os.path.join(datadir, f"e_pivot.emu{pivot}{key_syms}")
    flags = ["-s", "-r", "--non-acgt-to-a"]
    if not os.path.exists(a + ".bwt"):
        run([ref_bins["pfp_oracle"], "--bits", "64"] + flags + ["-o", a, fa])
    env = dict(os.environ, PFB_PIVOT=pivot, PFB_KEY_SYMS=key_syms)
    r = subprocess.run([emu, "--bits", "64"] + flags + ["-o", b, fa], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    assert_same_files(a, b, exts_for(True, True))


def fuzz_case(seed):
    """random job: window / modulus, record mix (random, variants of one base, tandem repeats, poly-A, N runs, tiny records), flags"""
    rng = np.random.default_rng(seed)
    w = int(rng.integers(3, 13))
    p = int(rng.choice([5, 7, 11, 20, 50, 100, 128]))
    recs = []
    base = random_bases(int(rng.integers(2_000, 25_000)), seed + 1)
    for r in range(int(rng.integers(1, 7))):
        kind = int(rng.integers(0, 6))
        if kind == 0:
            s = random_bases(int(rng.integers(1, 20_000)), seed * 31 + r)
        elif kind == 1:                                     # variant of the shared base: long ties in the dictionary
            s = base.copy()
            k = int(rng.integers(0, 30))
            pos = rng.integers(0, len(s), k)
            s[pos] = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, k)]
        elif kind == 2:                                     # tandem repeat
            unit = random_bases(int(rng.integers(1, 40)), seed * 17 + r)
            s = np.tile(unit, int(rng.integers(2, 400)))
        elif kind == 3:                                     # one symbol
            s = np.full(int(rng.integers(1, 5_000)), ord("ACGT"[int(rng.integers(0, 4))]), dtype=np.uint8)
        elif kind == 4:                                     # N runs and lower case inside random text
            s = random_bases(int(rng.integers(500, 15_000)), seed * 13 + r).copy()
            a = int(rng.integers(0, len(s) // 2))
            s[a:a + int(rng.integers(1, len(s) // 2))] = ord("N")
            b = int(rng.integers(0, len(s) // 2))
            s[b:b + 50] |= 0x20
        else:
            s = random_bases(int(rng.integers(0, 12)), seed * 7 + r)     # shorter than / around the window
        recs.append((f"r{r}", s))
    flags = ["-w", str(w), "-p", str(p)]
    if rng.random() < 0.7:
        flags.append("-s")
    if rng.random() < 0.6:
        flags.append("-r")
    if rng.random() < 0.5:
        flags.append("--non-acgt-to-a")
    return recs, (32 if rng.random() < 0.5 else 64), flags


FUZZ_KNOBS = [{}, {"PFB_PIVOT": "1", "PFB_KEY_SYMS": "3"}, {"PFB_PIVOT": "0", "PFB_KEY_SYMS": "3"}, {"PFB_EXT_ROUNDS": "0"},
              {"PFB_EXT_ROUNDS": "100", "PFB_PIVOT": "0"}, {"PFB_HARD_LIMITS": "2,4"}, {"PFB_KEY_COUNT_BITS": "0"},
              {"PFB_KEY_COUNT_BITS": "2", "PFB_PIVOT": "1", "PFB_KEY_SYMS": "5"},
              {"PFB_SMALL_GROUP": "0", "PFB_PIVOT": "1", "PFB_KEY_SYMS": "4"}, {"PFB_SMALL_GROUP": "3", "PFB_KEY_SYMS": "3"},
              {"PFB_SMALL_GROUP": "2", "PFB_PIVOT": "1", "PFB_KEY_SYMS": "2"},
              # 64-bit dictionary positions (the build of a dictionary of 2^32 bytes or more) forced at small sizes, alone,
              # through the in-memory key ranges such a dictionary is built in, and through the prefix-doubling hand-over
              {"PFB_DICT64": "1"}, {"PFB_DICT64": "1", "PFB_MEM_BUCKETS": "3", "PFB_PIVOT": "1", "PFB_KEY_SYMS": "3"},
              {"PFB_DICT64": "1", "PFB_EXT_ROUNDS": "0"}, {"PFB_MEM_BUCKETS": "4", "PFB_KEY_COUNT_BITS": "0"}]


@pytest.mark.parametrize("block", range(5))
def test_emu_fuzz_against_restatement(block, emu, ref_bins, datadir):
    """50 random jobs: the host logic of the device pipeline (emulator) writes what the C restatement of the reference writes;
    a job the restatement refuses (e.g. a single dictionary word) must be refused too"""
    for seed in range(block * 10, block * 10 + 10):
        recs, bits, flags = fuzz_case(7000 + seed)
        fa = os.path.join(datadir, f"fz{seed}.fa")
        write_fasta(fa, recs, line_width=int(50 + seed))
        a, b = os.path.join(datadir, f"fz{seed}.port"), os.path.join(datadir, f"fz{seed}.emu")
        ra = subprocess.run([ref_bins["pfp_oracle"], "--bits", str(bits)] + flags + ["-o", a, fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        knobs = FUZZ_KNOBS[seed % len(FUZZ_KNOBS)]          # the tie-refinement / counting paths rotate with the seed
        rb = subprocess.run([emu, "--bits", str(bits)] + flags + ["-o", b, fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                            env=dict(os.environ, **knobs))
        assert (ra.returncode == 0) == (rb.returncode == 0), (seed, flags, knobs, ra.stderr[-300:], rb.stderr[-300:])
        if ra.returncode == 0:
            try:
                assert_same_files(a, b, exts_for("-s" in flags, "-r" in flags))
            except AssertionError as e:
                raise AssertionError(f"seed {seed} flags {flags} bits {bits} knobs {knobs}: {e}")


def test_emu_write_last_sai_and_merge_docs(emu, ref_bins, datadir):
    """host logic of two small rows: <o>.last / <o>.sai (= last_ / sai_ of the restatement) and --docs through a merge of
    pre-parsed prefixes (load_parse reads <prefix>.docs, operator+= offsets the starts; include/pfbwt_io.hpp:216-218)"""
    recs = pangenome_records(30_000, 4, 30, 5)
    fa = os.path.join(datadir, "wl.fa")
    write_fasta(fa, recs, line_width=50)
    a, b = os.path.join(datadir, "wl.port"), os.path.join(datadir, "wl.emu")
    run([ref_bins["pfp_oracle"], "--bits", "64", "--write-last", "--parse-only", "-o", a, fa])
    run([emu, "--bits", "64", "--write-last", "--parse-only", "-o", b, fa])
    assert_same_files(a, b, ["last", "sai", "parse"])
    # docs: monolithic --print-docs == merge of two pre-parsed halves with their .docs
    mono = os.path.join(datadir, "wl_mono")
    run([emu, "--bits", "64", "--print-docs", "--parse-only", "-o", mono, fa])
    parts = []
    for g in range(2):
        p = os.path.join(datadir, f"wl_part{g}.fa")
        write_fasta(p, recs[2 * g:2 * g + 2], line_width=50)
        run([emu, "--bits", "64", "--print-docs", "--parse-only", p])
        parts.append(p)
    merged = os.path.join(datadir, "wl_merged")
    run([emu, "--merge", "--print-docs", "--bits", "64", "-o", merged] + parts)
    assert_same_files(mono, merged, ["dict", "parse", "docs"])
    if ref_bins["pfbwt-f64"]:
        refd = os.path.join(datadir, "wl_refdocs")
        run([ref_bins["pfbwt-f64"], "--print-docs", "--parse-only", "-o", refd, fa])
        assert_same_files(refd, merged, ["docs"])


@pytest.mark.parametrize("buckets,flags,bits", [(1, ["-s", "-r"], 32), (3, ["-s", "-r"], 64), (7, ["-r"], 32), (5, [], 32), (64, ["-s", "-r"], 32)])
def test_emu_bucketed_build(buckets, flags, bits, emu, ref_bins, datadir):
    """-m: the BWT is built one key range after the other and every slice streamed to its offset of .bwt/.sa/.ssa/.esa
    (run boundaries and the deferred run end across bucket borders, empty buckets) - same bytes as the one-piece build"""
    fa = os.path.join(datadir, "bucket.fa")
    if not os.path.exists(fa):
        write_fasta(fa, pangenome_records(40_000, 5, 40, 3) + nrun_records(1, 20_000, 9, big_run=3000), line_width=60)
    a = os.path.join(datadir, f"bucket{bits}{''.join(flags)}.port")
    b = os.path.join(datadir, f"bucket{bits}{''.join(flags)}_{buckets}.emu")
    run([ref_bins["pfp_oracle"], "--bits", str(bits)] + flags + ["-o", a, fa])
    r = subprocess.run([emu, "--bits", str(bits), "-m"] + flags + ["-o", b, fa], env=dict(os.environ, PFB_BUCKETS=str(buckets)),
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-1500:]
    assert f"buckets: {buckets}" in r.stderr.decode()
    assert_same_files(a, b, exts_for("-s" in flags, "-r" in flags))


def test_emu_load_rejects_mismatched_files(emu, datadir):
    """load_parser with a stale file set: a rank outside 1..Dw or an .occ that is not the histogram of .parse must end in
    'bad parse file', not in out-of-range reads"""
    fa = os.path.join(datadir, "stale.fa")
    write_fasta(fa, pangenome_records(20_000, 3, 20, 4), line_width=50)
    run([emu, "--bits", "64", "--parse-only", fa])
    import shutil
    for what in ("occ", "parse"):
        pfx = os.path.join(datadir, "stale_" + what)
        for e in ("dict", "occ", "parse", "n"):
            shutil.copy(fa + "." + e, pfx + "." + e)
        raw = bytearray(open(pfx + "." + what, "rb").read())
        if what == "occ":
            raw[0] ^= 1                       # a count that no longer matches the parse
        else:
            raw[4:8] = (0xFFFFFF).to_bytes(4, "little")       # a rank far past the dictionary
        open(pfx + "." + what, "wb").write(bytes(raw))
        r = run([emu, "--merge", "--bits", "64", "-o", os.path.join(datadir, "stale_out"), pfx], check=False)
        assert r.returncode != 0 and b"bad parse file" in r.stderr, (what, r.stderr[-300:])


WIDE_KNOBS = [{"PFB_DICT64": "1"}, {"PFB_DICT64": "1", "PFB_MEM_BUCKETS": "5"}, {"PFB_DICT64": "1", "PFB_MEM_BUCKETS": "64"},
              {"PFB_DICT64": "1", "PFB_KEY_COUNT_BITS": "0", "PFB_MEM_BUCKETS": "2"}, {"PFB_DICT64": "1", "PFB_BUCKETS": "3", "-m": "1"},
              {"PFB_DICT64": "1", "PFB_PIVOT": "1", "PFB_KEY_SYMS": "3", "PFB_SMALL_GROUP": "2"}, {"PFB_DICT64": "1", "PFB_EXT_ROUNDS": "0"}]


@pytest.mark.parametrize("knobs", WIDE_KNOBS, ids=["-".join(f"{k[4:] if k.startswith('PFB_') else k}{v}" for k, v in kn.items()) for kn in WIDE_KNOBS])
@pytest.mark.parametrize("flags,bits", [(["-s", "-r"], 64), ([], 32)])
def test_emu_wide_dictionary_positions(knobs, flags, bits, emu, ref_bins, datadir):
    """dictionaries of 2^32 bytes or more (pfbwt-f64 indexes the dictionary with 64-bit uint_t, gsa/gsacak.h:44-58): the
    build then runs on 64-bit dictionary positions - (u64, u64) sort pairs, two-array word lookup, ilist bases from the
    classify pass also without SA output - and one key range after the other, the slices put together in memory
    (Engine::build_in_key_ranges) or streamed (-m).  PFB_DICT64=1 / PFB_MEM_BUCKETS force all of it at a size the oracle runs."""
    fa = os.path.join(datadir, "wide.fa")
    if not os.path.exists(fa):
        write_fasta(fa, pangenome_records(40_000, 5, 40, 23) + nrun_records(1, 20_000, 29, big_run=3000) + [("r", random_bases(30_000, 31))], line_width=60)
    fl = flags + ["--non-acgt-to-a"]
    a = os.path.join(datadir, f"wide{bits}{''.join(flags)}.port")
    b = os.path.join(datadir, f"wide{bits}{''.join(flags)}_{'_'.join(knobs.values())}.emu")
    if not os.path.exists(a + ".bwt"):
        run([ref_bins["pfp_oracle"], "--bits", str(bits)] + fl + ["-o", a, fa])
    env = dict(os.environ, PFB_TRACE="1", **{k: v for k, v in knobs.items() if k.startswith("PFB_")})
    r = subprocess.run([emu, "--bits", str(bits)] + (["-m"] if "-m" in knobs else []) + fl + ["-o", b, fa], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-1500:]
    ranges = int(knobs.get("PFB_MEM_BUCKETS", knobs.get("PFB_BUCKETS", "1")))
    assert r.stderr.decode().count("K7 word tables (64-bit dictionary positions)") == ranges      # the path under test ran, once per key range
    assert_same_files(a, b, exts_for("-s" in flags, "-r" in flags))
