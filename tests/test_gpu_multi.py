"""GPU, >= 2 devices: the edge-sharded Gauss-Newton driver over NCCL and over the peer-memory
exchange kernel (frodo_gn_peer_allreduce), launched with torchrun on 127.0.0.1.  Skipped on a
single-GPU box (the CPU-side host logic is covered by tests/test_dist_cpu.py with gloo)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2,
                    reason="needs two GPUs")
def test_sharded_gauss_newton_two_gpus():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29631",
           os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.count("MULTI_GPU_OK") == 2, \
        r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2,
                    reason="needs two GPUs")
def test_two_devices_in_one_process():
    """Kernel attributes (dynamic shared memory, cooperative launch plan) are set per DEVICE: a
    process that drives a second GPU must get the same results there as on the first."""
    import mast3r_slam_backends as be
    from frodo_slam_b200 import synth
    from helpers import GP, MP, backend_gn, graph_inputs
    g = graph_inputs("calib32", 48, 64)
    d = synth.matching_pairs(1, 64, 80, seed=3)
    rays, pts, p0 = synth.prep_for_iter_proj(d["X11"], d["X21"])
    outs = []
    for index in (0, 1):
        dev = torch.device("cuda", index)
        T, dx = backend_gn(be, "rays", g, dev, 1, 2)
        p, c = be.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), MP["max_iter"], MP["lambda_init"],
                            MP["convergence_thresh"])
        pn, sc = be.refine_matches(d["D11"].half().to(dev),
                                   d["D21"].half().reshape(1, 64 * 80, -1).to(dev), p.long(),
                                   MP["radius"], MP["dilation_max"])
        torch.cuda.synchronize(dev)
        outs.append((T, dx, p.cpu(), pn.cpu(), sc.cpu()))
    import numpy as np
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert torch.equal(outs[0][2], outs[1][2]) and torch.equal(outs[0][3], outs[1][3])
    assert torch.equal(outs[0][4], outs[1][4])
