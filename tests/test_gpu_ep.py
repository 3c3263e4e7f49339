"""GPU, 2 ranks: expert parallelism reproduces the single-GPU layer (routing and drops bit-exact, numerics within
the bf16 tolerance).  Needs two GPUs on the box; skipped otherwise (the host logic is covered on CPU with gloo)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ep_matches_single_gpu():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "ep_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "EP_CHECK_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_two_devices_in_one_process():
    """Kernel attributes (dynamic shared memory, occupancy) are per device: the layer must run on cuda:1 after it ran on
    cuda:0 in the same process, with the same results."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import sys
    sys.path.insert(0, os.path.join(ROOT, "multimodal-moe_b200"))
    import b200moe
    outs = []
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        torch.manual_seed(4)
        m = b200moe.ContextMoE(512, 2048, 16, 2, 1.25, context_vocab_sizes=(6, 10)).to(dev)
        g = torch.Generator().manual_seed(5)
        x = torch.randn(8, 400, 512, generator=g).to(torch.bfloat16).to(dev).requires_grad_(True)
        c = torch.stack([torch.randint(0, v, (8,), generator=g) for v in (6, 10)], 1).to(dev)
        with torch.cuda.device(dev):
            y = m(x, c)
            (y.float().sum() + m.aux_loss).backward()
            torch.cuda.synchronize()
        outs.append((y.detach().cpu(), x.grad.cpu(), m.w1.grad.cpu()))
    for other in outs[1:]:
        for a, b in zip(outs[0], other):
            assert torch.equal(a, b)
