"""Tensor parallelism over peer memory: single-rank mechanics on one GPU, multi-rank via torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-6)).item()


def test_fused_collectives_single_rank():
    """tp = 1: the kernels push to their own inbox, publish and consume their own flags."""
    from jax_llama_b200.ops import kernels as K
    from jax_llama_b200.parallel.comm import PeerComm

    dev = torch.device("cuda", 0)
    comm = PeerComm(1, 0, dev)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(4, 512, generator=g).to(torch.bfloat16).to(dev)
    w = (torch.randn(1000, 512, generator=g) * 0.05).to(torch.bfloat16).to(dev)
    res = torch.randn(4, 1000, generator=g).to(torch.bfloat16).to(dev)
    want = K.linear(x.cpu(), w.cpu(), residual=res.cpu(), mode=K.RESIDUAL)
    for it in range(3):                       # epochs advance; both inbox parities get used
        comm.begin_sequence()
        got = comm.gemv_allreduce(x, w, res)
        got_b = comm.gemv_allreduce(x, w, res)
        comm.end_sequence()
        assert rel(got, want) < 2e-2 and torch.equal(got, got_b)
    assert int(comm.seq.item()) == 6
    src = torch.arange(64, dtype=torch.int32, device=dev)
    out = torch.zeros(64, dtype=torch.int32, device=dev)
    comm.begin_sequence()
    comm.allgather_small(src, out)
    comm.end_sequence()
    assert torch.equal(src, out)
    comm.close()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_tensor_parallel_engine_multi_gpu(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29400 + world), os.path.join(ROOT, "tests", "tp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=540, env=env, cwd=ROOT)
    assert r.returncode == 0 and f"TP_WORKER_OK world={world}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
