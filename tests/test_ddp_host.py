"""CPU (gloo, world_size 2): host-side logic of the data-parallel training exchange -- the flat arena, its
bucket tiling in backward-completion order, and the bucketed gradient all-reduce + BatchNorm-sum all-reduce."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _model():
    import music_style_transfer_ldm_b200 as M
    m = M.LDM(latent_dim=32, pretrained_path=None)
    m._freeze_encoder()
    return m


def test_arena_and_buckets_cover_every_trainable_parameter():
    from music_style_transfer_ldm_b200 import train as T
    m = _model()
    before = {k: v.clone() for k, v in m.state_dict().items()}
    order = T._backward_order(m)
    assert {id(p) for p in order} == {id(p) for p in m.parameters() if p.requires_grad}
    arena = T._Arena(order, "cpu")
    assert all(torch.equal(before[k], v) for k, v in m.state_dict().items())          # values and keys unchanged
    assert all(p.data_ptr() % 16 == 0 and p.grad.data_ptr() % 16 == 0 for p in order)
    assert sum(p.numel() for p in order) == 9_769_697                                   # trainable params with frozen encoder
    arena.g.fill_(1.0)
    assert all(float(p.grad.sum()) == p.numel() for p in order)                         # grads are views of the arena
    for n in (1, 2, 3):
        spans = T.bucket_spans(m, arena, n)
        assert len(spans) == n and spans[0][0] == 0 and spans[-1][1] == arena.size
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        T.GradReducer(arena.g, spans)                                                    # validates the tiling
    # decoder parameters complete first, the style encoder's first layer last
    assert order[0] is m.decoder.decoder[6].weight and order[-1] is m.style_encoder.enc1.bias
    with pytest.raises(Exception):
        T.GradReducer(arena.g, [(0, 10), (20, arena.size)])


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from music_style_transfer_ldm_b200 import train as T
        torch.manual_seed(0)
        m = _model()
        arena = T._Arena(T._backward_order(m), "cpu")
        red = T.GradReducer(arena.g, T.bucket_spans(m, arena, 3), None, world)
        g = torch.Generator().manual_seed(100 + rank)
        arena.g.copy_(torch.randn(arena.size, generator=g))
        mine = arena.g.clone()
        for i in range(3):
            red.reduce_bucket(i)
        red.wait()
        sums = torch.tensor([1.0 + rank, 10.0 * (rank + 1)], dtype=torch.float64)
        red.all_reduce_small(sums)
        q.put((rank, mine[:1000].tolist(), arena.g[:1000].tolist(), float(arena.g.double().sum()), sums.tolist()))
    finally:
        dist.destroy_process_group()


def test_bucketed_allreduce_two_ranks():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=180) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, m0, r0, s0, b0), (_, m1, r1, s1, b1) = res
    m0, m1, r0, r1 = (torch.tensor(v) for v in (m0, m1, r0, r1))
    assert torch.allclose(r0, m0 + m1) and torch.equal(r0, r1)         # every rank holds the SUM (Adam scales by 1/world)
    assert s0 == s1
    assert b0 == b1 == [3.0, 30.0]
