"""Peer-memory bandwidth of the expert-parallel row kernels, all rows remote (torchrun, 2 ranks):
torch copy_ into / out of the peer buffer as the reference, then ep_p2p_dispatch / combine / combine_bwd / unpermute_bwd."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "multimodal-moe_b200"))
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
from b200moe import _lib

rank, lr = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
W = dist.get_world_size()
_lib.load()
R, d, k = int(sys.argv[1]) if len(sys.argv) > 1 else 16384, int(sys.argv[2]) if len(sys.argv) > 2 else 1024, 1
nb = R * d * 2
mem = symm.empty(4 * nb, dtype=torch.uint8, device=dev)
hdl = symm.rendezvous(mem, dist.group.WORLD)
peers = torch.tensor(list(hdl.buffer_ptrs), dtype=torch.int64, device=dev)
peer = (rank + 1) % W
pbuf = hdl.get_buffer(peer, (4, R, d), torch.bfloat16)
lbuf = mem.view(torch.bfloat16).view(4, R, d)
x = torch.randn(R, d, device=dev).to(torch.bfloat16)
i32 = torch.int32
E, EL = W, 1
tok = torch.arange(R, dtype=i32, device=dev)
send_off = torch.tensor([0] * (peer + 1) + [R] * (E - peer), dtype=i32, device=dev)      # every row goes to expert `peer`
meta = torch.tensor([R, R // 128, 0, R, 0, 0, 0, 0], dtype=i32, device=dev)
dest_off = torch.zeros(E, dtype=i32, device=dev)
dest_of_row = torch.empty(R, dtype=i32, device=dev)
roa = torch.arange(R, dtype=i32, device=dev).view(R, 1)
gate = torch.ones(R, 1, device=dev)
gate_row = torch.ones(R, device=dev)
dgate = torch.empty(R, device=dev)
y = torch.empty(R, d, dtype=torch.bfloat16, device=dev)
st = _lib.stream_ptr()


def timed(name, fn, nbytes, iters=20):
    for _ in range(3):
        fn()
    hdl.barrier(0)
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        hdl.barrier(0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    if rank == 0:
        print(f"{name:28s} {ts[len(ts)//2]:8.1f} us  {nbytes / ts[len(ts)//2] / 1e3:7.0f} GB/s", flush=True)


timed("torch copy_ local->peer", lambda: pbuf[0].copy_(x), nb)
timed("torch copy_ peer->local", lambda: y.copy_(pbuf[0]), nb)
timed("torch copy_ local->local", lambda: y.copy_(x), nb)
timed("ep_p2p_dispatch", lambda: _lib.call("b200moe_ep_p2p_dispatch", x.data_ptr(), tok.data_ptr(), send_off.data_ptr(),
      meta.data_ptr(), dest_off.data_ptr(), peers.data_ptr(), 0, R, R, E, EL, d, dest_of_row.data_ptr(), st), nb)
timed("ep_p2p_combine k=1", lambda: _lib.call("b200moe_ep_p2p_combine", peers.data_ptr(), 0, R, dest_of_row.data_ptr(),
      roa.data_ptr(), gate.data_ptr(), R, 1, d, y.data_ptr(), st), nb)
timed("ep_p2p_combine_bwd", lambda: _lib.call("b200moe_ep_p2p_combine_bwd", x.data_ptr(), peers.data_ptr(), 0, nb, R,
      dest_of_row.data_ptr(), tok.data_ptr(), gate_row.data_ptr(), meta.data_ptr(), R, d, dgate.data_ptr(), st), 2 * nb)
timed("ep_p2p_unpermute_bwd", lambda: _lib.call("b200moe_ep_p2p_unpermute_bwd", peers.data_ptr(), 0, R,
      dest_of_row.data_ptr(), roa.data_ptr(), None, None, R, 1, d, E, y.data_ptr(), st), nb)
hdl.barrier(0)
torch.cuda.synchronize()
dist.destroy_process_group()
