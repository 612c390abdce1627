"""Single job spread over several ranks on the GPU, through the C ABI.
`cuda` mode: the ranks are processes sharing cuda:0 and exchange through the host all-gather callback
(gloo) -- runs on the one-GPU test box.  `nccl` mode: one GPU per rank, exchanges over NCCL (skipped
when the box has fewer GPUs than ranks).  Outputs are compared with the unmodified reference binary on
the concatenated input: bit-exact .dict/.occ/.parse/.ilist/.bwlast/.bwsai/.bwt/.sa/.ssa/.esa."""
import pytest

from pfp_util import ref_bin, pangenome_records, random_bases, nrun_records, exts_for, assert_same_files
from test_shard_gloo import run_job, split_records

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(ref_bin("pfbwt-f") is None, reason="reference binaries not built")]


def ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world,sa,rssa,bits", [(2, False, False, 32), (2, True, True, 32), (3, True, True, 64)])
def test_pangenome_job_shared_gpu(tmp_path, world, sa, rssa, bits):
    recs = pangenome_records(300_000, 6, 300, seed=21)
    out, refp = run_job("cuda", str(tmp_path), split_records(recs, world), 10, 100, bits, sa, rssa)
    assert_same_files(out, refp, exts_for(sa, rssa))


def test_random_job_shared_gpu(tmp_path):
    recs = [(f"r{i}", random_bases(1_000_000 + 13 * i, 300 + i)) for i in range(4)]
    out, refp = run_job("cuda", str(tmp_path), split_records(recs, 2), 10, 100, 32, False, True)
    assert_same_files(out, refp, exts_for(False, True))


def test_small_modulus_job_shared_gpu(tmp_path):
    recs = [(f"r{i}", random_bases(40_000 + 17 * i, 500 + i)) for i in range(6)]
    out, refp = run_job("cuda", str(tmp_path), split_records(recs, 3), 4, 7, 32, True, True)
    assert_same_files(out, refp, exts_for(True, True))


def test_nrun_job_shared_gpu(tmp_path):
    # lower-case + N runs (kept as N) of 3*10^4 symbols: ties inside the runs are closed by run-aware pivot rounds
    recs = nrun_records(4, 200_000, 5, big_run=30_000)
    out, refp = run_job("cuda", str(tmp_path), split_records(recs, 2), 10, 100, 64, True, True)
    assert_same_files(out, refp, exts_for(True, True))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_pangenome_job_nccl(tmp_path, world):
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    recs = pangenome_records(400_000, 8, 400, seed=31)
    out, refp = run_job("nccl", str(tmp_path), split_records(recs, world), 10, 100, 64, True, True)
    assert_same_files(out, refp, exts_for(True, True))


@pytest.mark.parametrize("world", [1, 2])
def test_multi_launcher(tmp_path, world):
    """pfbwt-f_b200/multi.py under torchrun: files dealt to the ranks, NCCL merge, slices written at their offsets"""
    import os, subprocess, sys
    from pfp_util import ROOT, write_fasta, run
    from test_shard_gloo import free_port
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    recs = pangenome_records(250_000, 6, 250, seed=77)
    fas = []
    for g in range(3):
        fa = os.path.join(str(tmp_path), f"in{g}.fa")
        write_fasta(fa, recs[2 * g:2 * g + 2], line_width=80)
        fas.append(fa)
    whole = os.path.join(str(tmp_path), "whole.fa")
    write_fasta(whole, recs, line_width=80)
    refp, out = os.path.join(str(tmp_path), "ref"), os.path.join(str(tmp_path), "out")
    run([ref_bin("pfbwt-f64"), "-s", "-r", "-o", refp, whole])
    if world == 2:
        fas = [fas[0], fas[1]] + [fas[2]]          # 3 files over 2 ranks: 2 + 1
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "pfbwt-f_b200", "multi.py"), "-s", "-r", "--save-parse", "-o", out] + fas
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert r.returncode == 0, r.stderr.decode(errors="replace")[-3000:]
    assert_same_files(out, refp, exts_for(True, True))


# ---- mid-record cuts (pfb_add_chunk / pfb_add_fasta_range) --------------------------------------------------------------
@pytest.mark.parametrize("world,fracs,bits", [(2, [0.47], 32), (3, [0.3, 0.71], 64)])
def test_one_record_split_by_bytes_shared_gpu(tmp_path, world, fracs, bits):
    """ONE record (the shape of BASELINE.json configs[1]) cut inside the record: each rank takes a byte range of the file"""
    recs = [("one", random_bases(3_000_000, 901))]
    out, refp = run_job("cuda", str(tmp_path), [recs], 10, 100, bits, True, True, byte_cuts=(world, fracs))
    assert_same_files(out, refp, exts_for(True, True))


def test_records_split_by_bytes_shared_gpu(tmp_path):
    recs = pangenome_records(200_000, 5, 200, seed=41) + [("tiny", random_bases(7, 1)), ("last", random_bases(90_000, 2))]
    out, refp = run_job("cuda", str(tmp_path), [recs], 10, 100, 64, True, True, byte_cuts=(3, [0.21, 0.66]), line_width=61)
    assert_same_files(out, refp, exts_for(True, True))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_one_record_split_by_bytes_nccl(tmp_path, world):
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    recs = [("one", random_bases(4_000_000, 902))]
    fr = [(g + 1) / world for g in range(world - 1)]
    out, refp = run_job("nccl", str(tmp_path), [recs], 10, 100, 32, True, True, byte_cuts=(world, fr))
    assert_same_files(out, refp, exts_for(True, True))


def test_chunks_on_one_context_equal_records(tmp_path):
    """chunks of a record added to ONE context concatenate: same parse and BWT as pfb_add_records"""
    import numpy as np
    from pfbwt_b200_loader import binding
    B = binding()
    a, b = random_bases(500_000, 5), random_bases(300_000, 6)
    one = B.Context(device=0, bits=32, w=10, p=100, get_sai=True)
    one.add_records(np.concatenate([a, b]), np.array([len(a), len(b)], dtype=np.uint64))
    one.finalize(); one.bwt_of_parse(); one.generate_bwt(True, True)
    two = B.Context(device=0, bits=32, w=10, p=100, get_sai=True)
    two.add_chunk(np.ascontiguousarray(a[:123_457]), ends_record=False)
    with pytest.raises(B.PfbError):
        two.add_records(b, np.array([len(b)], dtype=np.uint64))            # a record is still open
    two.add_chunk(np.ascontiguousarray(a[123_457:]), ends_record=True)
    two.add_chunk(np.ascontiguousarray(b[:10]), ends_record=False)
    with pytest.raises(B.PfbError):
        two.finalize()                                                     # the last record must be closed
    two.add_chunk(np.ascontiguousarray(b[10:]), ends_record=True)
    two.finalize(); two.bwt_of_parse(); two.generate_bwt(True, True)
    for f in ("dict", "parse", "occ", "bwt", "sa", "ssa", "esa"):
        assert np.array_equal(getattr(one, f)(), getattr(two, f)()), f
    one.close(); two.close()


def test_multi_launcher_splits_one_file(tmp_path):
    """pfbwt-f_b200/multi.py with ONE input file and two ranks: the file is cut into byte ranges (mid-record), NCCL job"""
    import os, subprocess, sys
    from pfp_util import ROOT, write_fasta, run
    from test_shard_gloo import free_port
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    recs = [("only", random_bases(2_500_000, 55))]
    whole = os.path.join(str(tmp_path), "whole.fa")
    write_fasta(whole, recs, line_width=80)
    refp, out = os.path.join(str(tmp_path), "ref"), os.path.join(str(tmp_path), "out")
    run([ref_bin("pfbwt-f64"), "-s", "-r", "-o", refp, whole])
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "pfbwt-f_b200", "multi.py"), "-s", "-r", "--save-parse", "-o", out, whole]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert r.returncode == 0, r.stderr.decode(errors="replace")[-3000:]
    assert_same_files(out, refp, exts_for(True, True))


# ---- merge_pfp over ranks: every rank loads its pre-parsed parts ---------------------------------------------------------------
@pytest.mark.parametrize("mode,world", [("cuda", 2), ("nccl", 2)])
def test_preparsed_parts_job(tmp_path, mode, world):
    import os
    from pfp_util import ROOT
    if mode == "nccl" and ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    recs = pangenome_records(200_000, 6, 200, seed=57) + [("t", random_bases(50_000, 3))]
    out, refp = run_job(mode, str(tmp_path), split_records(recs, world), 10, 100, 64, True, True,
                        preparse=[os.path.join(ROOT, "bin", "pfbwt-f64")])
    assert_same_files(out, refp, exts_for(True, True))


def test_multi_launcher_preparsed_parts(tmp_path):
    """pfbwt-f_b200/multi.py with pre-parsed prefixes as inputs (= merge_pfp + pfbwt-f64 --pfbwt-only in one job); a FASTA
    and a prefix may be mixed like merge_pfp's inputs"""
    import os, subprocess, sys
    from pfp_util import ROOT, write_fasta, run
    from test_shard_gloo import free_port
    world = 2 if ngpu() >= 2 else 1
    recs = pangenome_records(150_000, 6, 150, seed=59)
    fas = []
    for g in range(3):
        fa = os.path.join(str(tmp_path), f"in{g}.fa")
        write_fasta(fa, recs[2 * g:2 * g + 2], line_width=80)
        fas.append(fa)
        if g != 1:
            run([os.path.join(ROOT, "bin", "pfbwt-f64"), "--parse-only", fa])
    whole = os.path.join(str(tmp_path), "whole.fa")
    write_fasta(whole, recs, line_width=80)
    refp, out = os.path.join(str(tmp_path), "ref"), os.path.join(str(tmp_path), "out")
    run([ref_bin("pfbwt-f64"), "-s", "-r", "-o", refp, whole])
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "pfbwt-f_b200", "multi.py"), "-s", "-r", "--save-parse", "-o", out] + fas
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert r.returncode == 0, r.stderr.decode(errors="replace")[-3000:]
    assert_same_files(out, refp, exts_for(True, True))
