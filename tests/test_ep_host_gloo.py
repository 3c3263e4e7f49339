"""CPU / gloo, world_size 2: the expert-parallel host logic (b200moe.ep.route_tables / split_sizes).

Each rank routes its own tokens; the all-gathered histograms must give every rank (i) the drops of a single-rank
run over the concatenated, rank-major batch and (ii) consistent all-to-all split sizes, checked by
actually moving rows with all_to_all_single and comparing with the oracle's expert-major layout.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import moe_ref as R


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, T, E, k, cf, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from b200moe import ep
        g = torch.Generator().manual_seed(100 + rank)
        logits = torch.randn(T, E, generator=g)
        logits[:, 0] += 1.5                                   # skew: expert 0 overflows its capacity
        _, idx, _ = R.route(logits, k)
        hist = torch.stack([torch.bincount(idx[:, s], minlength=E) for s in range(k)]).to(torch.int32)
        mine = torch.cat([hist.reshape(-1), torch.tensor([T], dtype=torch.int32)])
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        allh = torch.stack(allh)
        hist_all, tokens_all = allh[:, :-1].reshape(world, k, E), allh[:, -1]
        goff, cap, kept = ep.route_tables(hist_all, tokens_all, rank, cf, k)
        send, recv, recv_counts, seg_expert = ep.split_sizes(kept, rank, world)

        # local view of the global plan, through the oracle with this rank's global offsets
        pos, keep = R.assign_slots(idx, E, cap, global_offset=goff.long())
        offs, kept_l, soa, tor = R.permutation(idx, keep, E, pad_to=1)
        assert kept_l.tolist() == kept[rank].tolist()
        assert sum(send) == int(kept_l.sum())

        # move rows: payload = (global token id, expert) so the receiver can check what it got
        payload = torch.zeros(int(offs[-1]), 2, dtype=torch.int64)
        rows_e = torch.repeat_interleave(torch.arange(E), kept_l)
        payload[:, 0] = tor + rank * T
        payload[:, 1] = rows_e
        got = torch.empty(sum(recv), 2, dtype=torch.int64)
        dist.all_to_all_single(got, payload, output_split_sizes=recv, input_split_sizes=send)
        EL = E // world
        assert ((got[:, 1] >= rank * EL) & (got[:, 1] < (rank + 1) * EL)).all()      # only my experts arrive
        seg_rows = torch.repeat_interleave(seg_expert.long() + rank * EL, recv_counts.long())
        assert torch.equal(seg_rows, got[:, 1])                                       # in the announced segment order
        gathered = [None] * world
        dist.all_gather_object(gathered, (idx, keep, got))
        if rank == 0:
            # single-rank truth over the concatenated batch
            idx_all = torch.cat([g_[0] for g_ in gathered])
            cap1 = R.capacity_of(cf, T * world, k, E)
            assert cap1 == cap
            _, keep1 = R.assign_slots(idx_all, E, cap1)
            assert torch.equal(keep1, torch.cat([g_[1] for g_ in gathered]))          # drops identical to W=1
            # every kept (token, expert) pair arrived exactly once at the owner of the expert
            want = {(int(t), int(e)) for t, s in keep1.nonzero().tolist() for e in [idx_all[t, s]]}
            have = [(int(a), int(b)) for g_ in gathered for a, b in g_[2].tolist()]
            assert len(have) == len(want) and set(have) == want
            out.put("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("T,E,k,cf", [(257, 8, 2, 1.0), (64, 4, 1, 0.5), (300, 16, 2, None)])
def test_ep_tables_world2(T, E, k, cf):
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, T, E, k, cf, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get() == "ok"
