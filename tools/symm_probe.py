"""Probe: does torch symmetric memory (peer-mapped buffers + signal pads) come up on this box?  torchrun -n 2"""
import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1 << 20, dtype=torch.uint8, device=f"cuda:{lr}")
    h = symm.rendezvous(t, dist.group.WORLD)
    print(rank, "OK buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "signal", [hex(p) for p in h.signal_pad_ptrs],
          "multicast", hex(h.multicast_ptr) if h.multicast_ptr else None, flush=True)
    t.fill_(rank + 1)
    h.barrier()
    peer = h.get_buffer((rank + 1) % 2, (16,), torch.uint8)
    print(rank, "peer sees", peer[:4].tolist(), flush=True)
    h.barrier()
except Exception as e:  # noqa: BLE001
    print(rank, "FAIL", repr(e)[:500], flush=True)
print(rank, "p2p access", torch.cuda.can_device_access_peer(lr, 1 - lr), flush=True)
dist.destroy_process_group()
