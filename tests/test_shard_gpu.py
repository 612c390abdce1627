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


@pytest.mark.parametrize("world", [2, 4, 8])
def test_pangenome_job_nccl(tmp_path, world):
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    recs = pangenome_records(400_000, 8, 400, seed=31)
    out, refp = run_job("nccl", str(tmp_path), split_records(recs, world), 10, 100, 64, True, True)
    assert_same_files(out, refp, exts_for(True, True))
