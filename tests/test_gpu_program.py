"""-m gpu: the persistent layer-program kernel (csrc/program.cu) must reproduce the launch-per-layer path: recorded
tcgen05 convolutions + attention cores executed by one cooperative kernel vs the same calls launched one by one."""
import pytest
import torch

from gpu_util import assert_close, gen

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _chain(ops, x, w1, b1, w2, b2, wq, kv, temb, skip):
    h = ops.conv(x, w1, b1, ksize=3, stride=1, act=ops.ACT_RELU)                              # 32 -> 64 @16x16
    h2 = ops.conv(h, w2, b2, ksize=3, stride=2, act=ops.ACT_RELU, add_nc=temb)                # 64 -> 128 @8x8 (+t_emb)
    q = ops.conv(h2, wq, None, ksize=1)                                                       # projection
    b, hh, ww, e = q.shape
    o = ops.attention(q.view(b, hh * ww, e), kv[:, :, :e], kv[:, :, e:], 4)
    return ops.conv(o.view(b, hh, ww, e), wq, None, ksize=1, add=skip), h


def test_program_matches_launches():
    from music_style_transfer_ldm_b200 import ops
    g = gen(11)
    n = 6
    bf = torch.bfloat16
    x = torch.randn((n, 16, 16, 32), generator=g).to(bf).to(DEV)
    w1 = ops.pack_weight((torch.randn((64, 32, 3, 3), generator=g) / 17).to(DEV), bf)
    w2 = ops.pack_weight((torch.randn((128, 64, 3, 3), generator=g) / 24).to(DEV), bf)
    wq = ops.pack_weight((torch.randn((128, 128, 1, 1), generator=g) / 11).to(DEV), bf)
    b1, b2 = torch.randn(64, generator=g).to(DEV), torch.randn(128, generator=g).to(DEV)
    kv = torch.randn((n, 64, 256), generator=g).to(bf).to(DEV)
    temb = torch.randn((n, 128), generator=g).to(DEV)
    skip = torch.randn((n, 8, 8, 128), generator=g).to(bf).to(DEV)
    ref, ref_h = _chain(ops, x, w1, b1, w2, b2, wq, kv, temb, skip)
    torch.cuda.synchronize()
    with ops.Program(DEV, 16) as prog:
        got, got_h = _chain(ops, x, w1, b1, w2, b2, wq, kv, temb, skip)
    assert prog.nops == 5
    got.zero_(); got_h.zero_()
    for _ in range(3):                      # the barrier state must be reusable across runs
        prog.run()
    torch.cuda.synchronize()
    assert_close(got_h.float(), ref_h.float(), 1e-6, "program: first conv")
    assert_close(got.float(), ref.float(), 1e-2, "program: chain output")
    # a sub-range: re-run only the last two ops
    got.zero_()
    prog.run(3, 2)
    torch.cuda.synchronize()
    assert_close(got.float(), ref.float(), 1e-2, "program: op sub-range")


def test_program_rejects_non_tc_ops():
    from music_style_transfer_ldm_b200 import ops
    x = torch.randn((2, 8, 8, 1), device=DEV)
    w = ops.pack_weight(torch.randn((16, 1, 3, 3), device=DEV), torch.float32)
    with pytest.raises(ops.MstkError):
        with ops.Program(DEV, 4):
            ops.conv(x, w, None, ksize=3, stride=2)


@pytest.mark.parametrize("batch,steps", [(2, 6), (64, 12)])
def test_sampling_loop_program_vs_graph(batch, steps):
    import music_style_transfer_ldm_b200 as M
    from oracle import ldm_oracle as O
    m = M.LDM(latent_dim=32, pretrained_path=None)
    m.load_state_dict(O.make_state_dict(1234), strict=True)
    m.to(DEV).eval().set_precision("bf16")
    g = gen(5)
    z = torch.randn((batch, 32, 16, 16), generator=g).to(DEV)
    st = {"s5": torch.rand((batch, 256, 4, 4), generator=g).to(DEV), "s6": torch.rand((batch, 512, 2, 2), generator=g).to(DEV)}
    m.use_programs = False
    ref, _ = m.content_style_ddim_sample(z, st, timesteps=steps, eta=0.0)
    m.use_programs = True
    for _ in range(2):                      # second call replays the cached program with refreshed inputs
        got, logs = m.content_style_ddim_sample(z, st, timesteps=steps, eta=0.0)
        torch.cuda.synchronize()
        assert_close(got, ref, 2e-3, "program sampling loop vs CUDA-graph loop")
    assert len(logs["timesteps"]) == steps - 1
