import os
import sys
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def ref_bins():
    """Prebuilt oracle binaries: the unmodified reference (oracle/_ref/pfbwt-f*, merge_pfp)
    and the CPU restatement (oracle/_ref/pfp_oracle).  Built by `make -C oracle`; on the GPU box the
    prebuilt files travel with the snapshot."""
    d = os.path.join(ROOT, "oracle", "_ref")
    need = ["pfp_oracle"]
    if not all(os.path.exists(os.path.join(d, n)) for n in need):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "all"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    out = {n: os.path.join(d, n) for n in ["pfbwt-f", "pfbwt-f64", "merge_pfp", "pfp_oracle"]}
    return {k: (v if os.path.exists(v) else None) for k, v in out.items()}


@pytest.fixture(scope="session")
def datadir(tmp_path_factory):
    return str(tmp_path_factory.mktemp("pfbdata"))
