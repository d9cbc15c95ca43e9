"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): NCCL block decomposition, both the neighbour-list loop
(persistent halo lists) and the per-step loop, against the single-rank run.  See tests/mgpu_check.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_ranks", [2, 4, 8])
def test_decomposed_run_matches_single_rank(n_ranks):
    import torch
    if torch.cuda.device_count() < n_ranks:
        pytest.skip("needs %d GPUs" % n_ranks)
    port = 29600 + n_ranks
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n_ranks),
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mgpu_check.py")],
                       capture_output=True, text=True, timeout=900)
    assert "MGPU_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
