"""Multi-GPU correctness on hardware: 2 ranks under torchrun, real NCCL, the drop-in module's additive
API (init_distributed / set_edge_shard) and the C ABI (m3r_gn_shard), both exchange paths and an uneven partition.
Every rank's poses must be BIT-IDENTICAL to the unsharded single-GPU call.  Skips with fewer than 2 devices
(run it with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`)."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(900)
def test_two_rank_sharded_solve_bit_identical_to_unsharded():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tests", "_multi_worker.py")]
    res = subprocess.run(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=840)
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("M3R_MULTI ")]
    assert res.returncode == 0 and len(lines) == 2, res.stdout[-4000:]
    for ln in lines:
        r = json.loads(ln[len("M3R_MULTI "):])
        print(json.dumps(r))
        assert r["unsharded_info"][:2] == [10, 0]
        assert r["auto_exchange"] == "allgather"
        for name in ("pybind_equal_allgather", "pybind_equal_allreduce", "pybind_uneven_allreduce", "cabi_equal_allgather",
                     "cabi_uneven_allreduce", "cabi_equal_peer", "cabi_equal_peer_again", "cabi_equal_peer_3", "cabi_equal_peer_4"):
            c = r["cases"][name]
            assert c["bit_identical_to_unsharded"] and c["identical_across_ranks"], (name, c)
            assert c["info"][:2] == [10, 0], (name, c)
        ss = r["cases"]["sharded_store"]
        assert ss["global_pairs"] == 90 and ss["local_pairs"] == 45 and ss["identical_across_ranks"], ss
        assert ss["max_abs_diff_vs_single_device_graph"] < 1e-5 and ss["moved"] > 1e-3, ss
        assert r["cases"]["allgather_on_uneven_rejected"] is True
        assert r["cases"]["peer_on_uneven_rejected"] is True
        assert r["cases"]["comm_allreduce_f64_exact"] is True
