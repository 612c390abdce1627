"""N>1 path of the single-job multi-GPU build on CPU: world_size-2/3 jobs whose
exchanges run over torch.distributed/gloo.  The device pipeline is the serial logic emulator
(tests/emu), so what is checked is the host logic of the sharded path -- piece boundaries and
halos, the straddling phrase, the dictionary merge, rank renaming, key-range splitters, slice
offsets, run boundaries across slices -- against the unmodified reference binary on the
concatenated input (merge_pfp semantics: src/merge_pfp.cpp:153-155)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from pfp_util import ROOT, ref_bin, run, write_fasta, pangenome_records, random_bases, nrun_records, exts_for, assert_same_files

EMU_LIB = os.path.join(ROOT, "tests", "emu", "libpfb_emu.so")
WORKER = os.path.join(ROOT, "tests", "shard_worker.py")

pytestmark = pytest.mark.skipif(not os.path.exists(EMU_LIB) or ref_bin("pfbwt-f") is None,
                                reason="emulator library or reference binaries not built")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_job(mode, tmp, pieces, w, p, bits, sa, rssa, non_acgt_to_a=False, expect_fail=False, byte_cuts=None, line_width=70, preparse=None):
    """pieces: list (one per rank) of record lists.  Returns (sharded prefix, reference prefix).
    byte_cuts = (world, [fractions]) instead splits ONE file holding all records by byte ranges (cuts fall anywhere,
    also inside records and header lines: pfb_add_fasta_range).  preparse = a command prefix (e.g. the emulator or bin/pfbwt-f64):
    every rank's part is parsed to <part>.dict/.parse/... first and the job LOADS the parts (merge_pfp over ranks)."""
    world = len(pieces)
    allrecs = [r for pc in pieces for r in pc]
    whole = os.path.join(tmp, "whole.fa")
    write_fasta(whole, allrecs, line_width=line_width)
    fas = []
    if byte_cuts is not None:
        world, fr = byte_cuts
        size = os.path.getsize(whole)
        cuts = [0] + [int(size * f) for f in fr] + [size]
        assert len(cuts) == world + 1
        fas = [f"{whole}@{cuts[g]}:{cuts[g + 1]}" for g in range(world)]
    else:
        for g, pc in enumerate(pieces):
            fa = os.path.join(tmp, f"part{g}.fa")
            write_fasta(fa, pc, line_width=line_width)
            fas.append(fa)
            if preparse:
                run(preparse + ["-w", str(w), "-p", str(p), "--parse-only", fa])
                assert os.path.exists(fa + ".dict") and os.path.exists(fa + ".parse")
    refp = os.path.join(tmp, "ref")
    cmd = [ref_bin("pfbwt-f64" if bits == 64 else "pfbwt-f"), "-w", str(w), "-p", str(p), "-o", refp]
    if sa:
        cmd.append("-s")
    if rssa:
        cmd.append("-r")
    if not expect_fail:
        run(cmd + [whole])
    out = os.path.join(tmp, "shard")
    port = free_port()
    procs = [subprocess.Popen([sys.executable, WORKER, mode, str(g), str(world), str(port), out, str(w), str(p), str(bits),
                               str(int(sa)), str(int(rssa)), fas[g]], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
             for g in range(world)]
    errs = []
    for g, pr in enumerate(procs):
        try:
            _, e = pr.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        if pr.returncode != 0:
            errs.append(f"rank {g} rc={pr.returncode}: {e.decode(errors='replace')[-1500:]}")
    if expect_fail:
        assert len(errs) == world, f"every rank must fail together, got {len(errs)} of {world}: " + "\n".join(errs)
        return out, "\n".join(errs)
    assert not errs, "\n".join(errs)
    return out, refp


def split_records(recs, world):
    """contiguous, balanced, never empty while len(recs) >= world (the rule of pfbwt-f_b200/multi.py)"""
    n = len(recs)
    return [recs[n * g // world:n * (g + 1) // world] for g in range(world)]


@pytest.mark.parametrize("world,sa,rssa,bits", [(2, False, False, 32), (2, True, True, 32), (3, True, True, 64)])
def test_pangenome_job(tmp_path, world, sa, rssa, bits):
    recs = pangenome_records(6000, 6, 12, seed=11)
    out, refp = run_job("emu", str(tmp_path), split_records(recs, world), 10, 100, bits, sa, rssa)
    assert_same_files(out, refp, exts_for(sa, rssa))


def test_random_job_small_modulus(tmp_path):
    # p = 7: short phrases, many triggers next to the cuts, first trigger often inside the first w bytes
    recs = [(f"r{i}", random_bases(3000 + 17 * i, 100 + i)) for i in range(4)]
    out, refp = run_job("emu", str(tmp_path), split_records(recs, 2), 4, 7, 32, True, True)
    assert_same_files(out, refp, exts_for(True, True))


def test_rank_without_trigger(tmp_path):
    # the middle rank's records hold no trigger at all (p large, tiny records): it owns no phrase and the
    # phrase that starts on rank 0 runs across it into rank 2
    big = [(f"b{i}", random_bases(20000, 7 + i)) for i in range(2)]
    tiny = [("t0", np.frombuffer(b"ACGTTGCAAC", dtype=np.uint8))]
    pieces = [[big[0]], tiny, [big[1]]]
    whole_has_trigger = True
    out, refp = run_job("emu", str(tmp_path), pieces, 10, 2000, 32, True, False)
    assert whole_has_trigger
    assert_same_files(out, refp, exts_for(True, False))


def test_failure_is_collective(tmp_path):
    # rank 1's records hold a byte outside the hasher's alphabet (include/hash.hpp:31): the error is found on
    # rank 1 only, every rank must return it together instead of waiting in an exchange
    good = [(f"g{i}", random_bases(5000, 40 + i)) for i in range(2)]
    bad = ("bad", np.frombuffer(b"ACGTACGTRRACGT" * 50, dtype=np.uint8))
    out, msg = run_job("emu", str(tmp_path), [[good[0]], [bad], [good[1]]], 10, 100, 32, True, False, expect_fail=True)
    assert "invalid character" in msg


def test_nrun_job(tmp_path):
    # N runs of 10^4 symbols (kept as N: a stopper symbol, one window round would consume one symbol each): the ranks hold
    # only their key ranges, so these ties are closed by run-aware pivot rounds, not by prefix doubling
    recs = nrun_records(4, 40_000, 5, big_run=12_000)
    out, refp = run_job("emu", str(tmp_path), split_records(recs, 2), 10, 100, 64, True, True)
    assert_same_files(out, refp, exts_for(True, True))


@pytest.mark.parametrize("seed", [9041, 9047, 9058, 9066])
def test_random_jobs_over_ranks(tmp_path, seed):
    """random jobs of tests/test_emu_logic.py::fuzz_case (window, modulus, record mix incl. tandem repeats, one-symbol and tiny
    records) cut into 2-3 contiguous pieces: the ranks together write what the reference writes for the whole record list
    (a one-off run over 90 seeds found no difference)"""
    from test_emu_logic import fuzz_case
    recs, bits, flags = fuzz_case(seed)
    while len(recs) < 2:
        seed += 1000
        recs, bits, flags = fuzz_case(seed)
    world = 2 if len(recs) < 3 else 3
    cuts = [len(recs) * g // world for g in range(world + 1)]
    pieces = [recs[cuts[g]:cuts[g + 1]] for g in range(world)]
    w, p = int(flags[flags.index("-w") + 1]), int(flags[flags.index("-p") + 1])
    sa, rssa = "-s" in flags, "-r" in flags
    try:
        out, refp = run_job("emu", str(tmp_path), pieces, w, p, bits, sa, rssa)
    except RuntimeError:
        pytest.skip("the reference refuses this job")
    assert_same_files(out, refp, exts_for(sa, rssa))


# ---- mid-record cuts: ONE FASTA file split by byte ranges over the ranks (pfb_add_fasta_range) -----------------------
@pytest.mark.parametrize("world,fracs,sa,rssa,bits", [
    (2, [0.5], True, True, 32),                  # one record cut in the middle
    (3, [0.31, 0.64], True, True, 64),
    (4, [0.2, 0.5, 0.77], False, False, 32),
])
def test_single_record_split_by_bytes(tmp_path, world, fracs, sa, rssa, bits):
    recs = [("one", random_bases(60_000, 321))]
    out, refp = run_job("emu", str(tmp_path), [recs], 10, 100, bits, sa, rssa, byte_cuts=(world, fracs))
    assert_same_files(out, refp, exts_for(sa, rssa))


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_multi_record_split_by_bytes(tmp_path, seed):
    """several records of uneven length, cuts at pseudo-random byte offsets: inside records, inside header lines, right at
    a header, at a line end; small modulus so that triggers sit next to the cuts"""
    rng = np.random.default_rng(seed)
    recs = [(f"rec{i}_{'x' * int(rng.integers(0, 40))}", random_bases(int(rng.integers(30, 9000)), 1000 * seed + i)) for i in range(int(rng.integers(2, 7)))]
    world = int(rng.integers(2, 4))
    fracs = sorted(float(x) for x in rng.uniform(0.05, 0.95, world - 1))
    w, p = (4, 7) if seed % 2 else (10, 50)
    out, refp = run_job("emu", str(tmp_path), [recs], w, p, 32, True, True, byte_cuts=(world, fracs), line_width=int(rng.integers(20, 90)))
    assert_same_files(out, refp, exts_for(True, True))


def test_split_cut_exactly_at_header_and_in_header(tmp_path):
    recs = [("alpha", random_bases(5000, 77)), ("beta_with_a_long_name_to_cut_inside", random_bases(7000, 78)), ("gamma", random_bases(6000, 79))]
    whole = os.path.join(str(tmp_path), "probe.fa")
    write_fasta(whole, recs, line_width=60)
    data = open(whole, "rb").read()
    size = len(data)
    h2 = data.index(b">beta")
    h3 = data.index(b">gamma")
    for cuts in ([h2, h3], [h2 + 7, h3 - 1], [h2 - 1, h2 + 1]):
        sub = tmp_path / f"c{cuts[0]}_{cuts[1]}"
        sub.mkdir()
        fr = [(c + 0.25) / size for c in cuts]
        out, refp = run_job("emu", str(sub), [recs], 10, 100, 32, True, False, byte_cuts=(3, fr), line_width=60)
        assert_same_files(out, refp, exts_for(True, False))


@pytest.mark.parametrize("world,bits", [(2, 64), (3, 32)])
def test_preparsed_parts_job(tmp_path, world, bits):
    """merge_pfp over ranks: every rank LOADS its pre-parsed part (<part>.dict/.parse from --parse-only; the rule of
    src/merge_pfp.cpp:100-113) instead of parsing a FASTA, the merge and the BWT build are the same collective job"""
    emu = os.path.join(ROOT, "tests", "emu", "pfb_emu")
    recs = pangenome_records(5000, 6, 10, seed=17) + [("t", random_bases(900, 3))]
    out, refp = run_job("emu", str(tmp_path), split_records(recs, world), 10, 100, bits, True, True, preparse=[emu, "--bits", str(bits)])
    assert_same_files(out, refp, exts_for(True, True))


def test_skewed_dictionary_owners(tmp_path):
    """dictionary merge by owner ranks with a skewed key distribution: tandem repeats make most words start with the same 21
    symbols (one owner gets nearly everything, others nothing) and the same words occur on every rank (deduplicated by their
    owner); poly-A records add words that are prefixes of each other"""
    rng = np.random.default_rng(77)
    unit = random_bases(37, 5)
    recs = []
    for i in range(8):
        s = np.tile(unit, 300).copy()
        pos = rng.integers(0, len(s), 25)
        s[pos] = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 25)]
        recs.append((f"t{i}", s))
        recs.append((f"a{i}", np.full(200 + 13 * i, ord("A"), dtype=np.uint8)))
    out, refp = run_job("emu", str(tmp_path), split_records(recs, 4), 6, 11, 64, True, True)
    assert_same_files(out, refp, exts_for(True, True))


def test_world8_job(tmp_path):
    """the rank count of the north-star run: 8 ranks (pieces of unequal size, one of them a single tiny record)"""
    recs = pangenome_records(3000, 9, 10, seed=31) + [("tiny", random_bases(15, 2))] + [(f"r{i}", random_bases(2500 + 31 * i, 40 + i)) for i in range(6)]
    out, refp = run_job("emu", str(tmp_path), split_records(recs, 8), 10, 100, 64, True, True)
    assert_same_files(out, refp, exts_for(True, True))
