"""-m gpu, needs >= 2 GPUs (skipped otherwise): the NCCL exchange of the sample-sharded path, one process per GPU
under torch.distributed (rendezvous on 127.0.0.1), against the oracle on the whole cohort."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    import basevarc_b200 as B
    return B.load_library().bvc_device_count()


@pytest.mark.parametrize("world", [2])
def test_nccl_sharded_parity(world, tmp_path):
    if _n_gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tests", "shard_worker.py"), "nccl"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "shard_worker ok" in r.stdout
