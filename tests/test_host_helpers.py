"""CPU: small host-side pieces of the training step -- the BatchNorm statistic reduce hook of bn_sync_group (the trainer
routes it to its peer-memory exchange), the step-scoped zero pool, the pack recorder plumbing."""
import torch


def test_bn_sync_group_reduce_hook_and_scope():
    from music_style_transfer_ldm_b200 import model as Mo
    calls = []

    def reduce(t):
        calls.append(t.numel())
        t.mul_(2.0)

    s = torch.ones(8, dtype=torch.float64)
    assert Mo._sync_bn_stats(s.clone()).sum() == 8              # outside a scope: untouched, no collective
    with Mo.bn_sync_group(None, 2, reduce=reduce):
        assert Mo._bn_world() == 2
        out = Mo._sync_bn_stats(s.clone())
        assert out.sum() == 16 and calls == [8]
        with Mo.bn_sync_group(None, 1):                          # nested scope (the autograd face) switches it off again
            assert Mo._sync_bn_stats(s.clone()).sum() == 8
        assert Mo._bn_world() == 2
    assert Mo._bn_world() == 1 and Mo._BN_GROUP["reduce"] is None


def test_zero_pool_hands_out_aligned_zeroed_slices_and_falls_back():
    from music_style_transfer_ldm_b200 import ops
    pool = ops.ZeroPool("cpu", doubles=256)
    old, ops.ZERO_POOL = ops.ZERO_POOL, pool
    try:
        pool.begin()
        a, b = ops.zeros_f64(100, "cpu"), ops.zeros_f64(64, "cpu")
        assert a.data_ptr() % 256 == pool.buf.data_ptr() % 256 and (b.data_ptr() - a.data_ptr()) == 128 * 8
        a.fill_(3.0)
        c = ops.zeros_f64(200, "cpu")                            # does not fit any more: a fresh tensor
        assert c.numel() == 200 and float(c.sum()) == 0.0 and c.data_ptr() != pool.buf.data_ptr()
        pool.begin()                                             # next step: same addresses, zero again
        a2 = ops.zeros_f64(100, "cpu")
        assert a2.data_ptr() == a.data_ptr() and float(a2.sum()) == 0.0
    finally:
        ops.ZERO_POOL = old
    assert ops.zeros_f64(4, "cpu").numel() == 4                  # no pool active


def test_device_prefetch_cpu_path_keeps_order_and_structure():
    """train.device_prefetch on a CPU device: a plain pass-through (the CUDA path is exercised by tests/test_gpu_train.py)"""
    import torch
    from music_style_transfer_ldm_b200.train import device_prefetch
    loader = [((torch.full((2,), float(i)), "label"), (torch.full((3,), float(-i)), "label")) for i in range(4)]
    got = list(device_prefetch(loader, "cpu", pick=lambda el: (el[0][0], el[1][0])))
    assert [float(c[0]) for c, _ in got] == [0.0, 1.0, 2.0, 3.0] and [float(s[0]) for _, s in got] == [0.0, -1.0, -2.0, -3.0]
    single = list(device_prefetch([(torch.ones(2) * i, 0) for i in range(3)], torch.device("cpu"), pick=lambda el: el[0]))
    assert [float(t[0]) for t in single] == [0.0, 1.0, 2.0] and list(device_prefetch([], "cpu")) == []
