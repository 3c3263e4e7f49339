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


def _worker(rank, world, port, T, E, k, cf, chunks, out):
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
        C, ELc, order = ep.chunk_layout(E, world, chunks)
        send, recv, recv_counts, seg_expert = ep.split_sizes(kept, rank, world, C)

        # local view of the global plan, through the oracle with this rank's global offsets
        pos, keep = R.assign_slots(idx, E, cap, global_offset=goff.long())
        offs, kept_l, soa, tor = R.permutation(idx, keep, E, pad_to=1)
        assert kept_l.tolist() == kept[rank].tolist()
        assert sum(sum(s_) for s_ in send) == int(kept_l.sum())

        # send buffer in the chunked layout (group, destination rank, expert); payload = (global token id, expert)
        by_pos = sorted(range(E), key=lambda e: int(order[e]))
        rows = []
        for e in by_pos:
            seg = tor[int(offs[e]):int(offs[e + 1])]
            rows.append(torch.stack([seg + rank * T, torch.full_like(seg, e)], 1))
        payload = torch.cat(rows) if rows else torch.zeros(0, 2, dtype=torch.int64)
        EL = E // world
        got_all, start = [], 0
        for c in range(C):
            n_out = sum(send[c])
            got = torch.empty(sum(recv[c]), 2, dtype=torch.int64)
            dist.all_to_all_single(got, payload[start:start + n_out].contiguous(), output_split_sizes=recv[c],
                                   input_split_sizes=send[c])
            start += n_out
            lo = rank * EL + c * ELc
            assert ((got[:, 1] >= lo) & (got[:, 1] < lo + ELc)).all()                # only this group's experts arrive
            seg_rows = torch.repeat_interleave(seg_expert.long() + lo, recv_counts[c].long())
            assert torch.equal(seg_rows, got[:, 1])                                  # in the announced segment order
            got_all.append(got)
        got = torch.cat(got_all)
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


@pytest.mark.parametrize("T,E,k,cf,chunks", [(257, 8, 2, 1.0, 2), (64, 4, 1, 0.5, 1), (300, 16, 2, None, 4)])
def test_ep_tables_world2(T, E, k, cf, chunks):
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, T, E, k, cf, chunks, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get() == "ok"


def test_split_sizes_numpy_twin():
    """The numpy split-size helper of the hot path agrees with the torch one the gloo test pins."""
    from b200moe.ep import split_sizes, split_sizes_np
    g = torch.Generator().manual_seed(4)
    for W, E, C in ((2, 8, 1), (4, 16, 2), (8, 64, 4), (1, 4, 2)):
        kept = torch.randint(0, 500, (W, E), generator=g)
        for rank in range(W):
            send, recv, _, _ = split_sizes(kept, rank, W, C)
            send_np, recv_np = split_sizes_np(kept.numpy(), rank, W, C)
            assert send == send_np and recv == recv_np


def _p2p_worker(rank, world, port, T, E, k, cf, out):
    """Peer-memory dispatch on CPU: the 'stores into the expert rank's buffer' are emulated by gathering
    (destination rank, destination row, payload) triples; the layout must be a bijection onto the unpadded part of
    every expert segment, in global priority order (rank-major inside an expert)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from b200moe import ep, ep_p2p
        g = torch.Generator().manual_seed(200 + rank)
        logits = torch.randn(T, E, generator=g)
        logits[:, 1] += 1.0
        _, idx, _ = R.route(logits, k)
        hist = torch.stack([torch.bincount(idx[:, s], minlength=E) for s in range(k)]).to(torch.int32)
        mine = torch.cat([hist.reshape(-1), torch.tensor([T], dtype=torch.int32)])
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        allh = torch.stack(allh)
        goff, cap, kept = ep.route_tables(allh[:, :-1].reshape(world, k, E), allh[:, -1], rank, cf, k)
        dest_off, offsets = ep_p2p.p2p_layout(kept, rank)
        EL = E // world
        _, keep = R.assign_slots(idx, E, cap, global_offset=goff.long())
        offs, kept_l, soa, tor = R.permutation(idx, keep, E, pad_to=1)          # send order: experts back to back
        triples = []
        for e in range(E):
            seg = tor[int(offs[e]):int(offs[e + 1])]
            for i, t in enumerate(seg.tolist()):
                triples.append((e // EL, int(dest_off[e]) + i, rank * T + t, e))
        gathered = [None] * world
        dist.all_gather_object(gathered, (triples, offsets.tolist(), kept.tolist()))
        if rank == 0:
            for g_ in gathered:
                assert g_[1] == gathered[0][1] and g_[2] == gathered[0][2]      # every rank derived the same layout
            offsets_all, kept_all = torch.tensor(gathered[0][1]), torch.tensor(gathered[0][2])
            tot = kept_all.sum(0)
            for w in range(world):
                rows = sorted((row, tok, e) for g_ in gathered for (dw, row, tok, e) in g_[0] if dw == w)
                want_rows = [r_ for j in range(EL) for r_ in range(int(offsets_all[w, j]),
                                                                   int(offsets_all[w, j]) + int(tot[w * EL + j]))]
                assert [r_[0] for r_ in rows] == want_rows                       # bijection onto the unpadded rows
                for row, tok, e in rows:
                    j = e - w * EL
                    assert int(offsets_all[w, j]) <= row < int(offsets_all[w, j + 1])
                for j in range(EL):                                              # rank-major inside an expert
                    toks = [tok // T for row, tok, e in rows if e == w * EL + j]
                    assert toks == sorted(toks)
                assert int(offsets_all[w, EL]) % 128 == 0
            out.put("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("T,E,k,cf", [(257, 8, 2, 1.0), (64, 4, 1, 0.5), (300, 16, 2, 1.25)])
def test_p2p_layout_world2(T, E, k, cf):
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_p2p_worker, args=(r, 2, port, T, E, k, cf, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get() == "ok"
