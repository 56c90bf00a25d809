"""The packed spectrogram format and the pair dataset (music_style_transfer_ldm_b200/dataset.py) against the reference's
models/dataset.py:28-53, 206-305 -- SURVEY 8(f) rank 3.

CPU: PNG folders written here with PIL (128 x 130, the reference's chunk size) -> packed store -> the same elements,
the same CSV pairing semantics and the same ``generate_pairings`` file; LIVE against the unmodified reference classes
where /root/reference exists.  GPU (-m gpu): the one-kernel batch path against the per-item format access.
"""
import csv
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import refshim  # noqa: E402

LABELS = {"guitar": 7, "harp": 5, "piano": 9}


@pytest.fixture(scope="module")
def png_root(tmp_path_factory):
    from PIL import Image
    root = tmp_path_factory.mktemp("spectrograms")
    rng = np.random.RandomState(3)
    for label, n in LABELS.items():
        os.makedirs(root / label)
        for i in range(n):
            Image.fromarray(rng.randint(0, 256, size=(128, 130), dtype=np.uint8)).save(root / label / f"chunk_{i:03d}.png")
    return str(root)


def _ds(root, csv_path):
    from music_style_transfer_ldm_b200.dataset import SpectrogramPairDataset
    return SpectrogramPairDataset(root, csv_path)


def test_pack_save_load_and_items(png_root, tmp_path):
    from PIL import Image
    from music_style_transfer_ldm_b200.dataset import PackedSpectrograms
    st = PackedSpectrograms.from_folder(png_root)
    assert st.labels == sorted(LABELS) and len(st) == sum(LABELS.values()) and (st.height, st.width) == (128, 130)
    ref = np.asarray(Image.open(os.path.join(png_root, "harp", "chunk_002.png")), dtype=np.float32)[:128, :128] / 255.0
    it = st.item("harp", 2)
    assert it.shape == (1, 128, 128) and it.dtype == torch.float32 and np.array_equal(it[0].numpy(), ref)
    st.save(str(tmp_path / "packed.pt"))
    st2 = PackedSpectrograms.load(str(tmp_path / "packed.pt"))
    assert all(torch.equal(st.images[k], st2.images[k]) for k in st.labels)
    with pytest.raises(IndexError):
        st.flat_index("harp", LABELS["harp"])


def test_pairings_and_elements(png_root, tmp_path):
    from music_style_transfer_ldm_b200.dataset import SpectrogramPairDataset
    csv_path = str(tmp_path / "pairs.csv")
    SpectrogramPairDataset.generate_pairings(png_root, csv_path, num_pairs=40)
    rows = list(csv.reader(open(csv_path)))
    assert len(rows) == 40 and all(r[0] != r[2] and 0 <= int(r[1]) < LABELS[r[0]] and 0 <= int(r[3]) < LABELS[r[2]] for r in rows)
    ds = _ds(png_root, csv_path)
    assert len(ds) == 40
    (a, la), (b, lb) = ds[7]
    assert (la, lb) == (rows[7][0], rows[7][2]) and a.shape == b.shape == (1, 128, 128)
    assert torch.equal(a, ds.store.item(rows[7][0], int(rows[7][1])))


@pytest.mark.skipif(not refshim.available(), reason="the reference tree exists only in the build container")
def test_live_against_reference_dataset(png_root, tmp_path):
    refshim.load()
    import importlib
    ref_ds = importlib.import_module("dataset")              # /root/reference/models/dataset.py (on sys.path via refshim)
    ours_csv, ref_csv = str(tmp_path / "ours.csv"), str(tmp_path / "ref.csv")
    from music_style_transfer_ldm_b200.dataset import SpectrogramDataset, SpectrogramPairDataset
    SpectrogramPairDataset.generate_pairings(png_root, ours_csv, num_pairs=60)
    ref_ds.SpectrogramPairDataset.generate_pairings(png_root, ref_csv, num_pairs=60)
    assert open(ours_csv).read() == open(ref_csv).read()
    r, m = ref_ds.SpectrogramPairDataset(png_root, ref_csv), _ds(png_root, ours_csv)
    assert len(r) == len(m)
    for i in (0, 13, 59):
        (ra, rla), (rb, rlb) = r[i]
        (ma, mla), (mb, mlb) = m[i]
        assert (rla, rlb) == (mla, mlb) and torch.equal(ra, ma) and torch.equal(rb, mb)
    cfg = {"processed_spectograms_dataset_folderpath": png_root}
    rs, ms = ref_ds.SpectrogramDataset(cfg), SpectrogramDataset(cfg)
    assert len(rs) == len(ms)
    for i in (0, 8, len(rs) - 1):
        assert torch.equal(rs[i][0], ms[i][0]) and rs[i][1] == ms[i][1]


@pytest.mark.gpu
def test_batch_loader_one_kernel_per_side(png_root, tmp_path):
    from music_style_transfer_ldm_b200 import ops
    from music_style_transfer_ldm_b200.dataset import SpectrogramDataset, SpectrogramPairDataset
    csv_path = str(tmp_path / "pairs.csv")
    SpectrogramPairDataset.generate_pairings(png_root, csv_path, num_pairs=50)
    ds = _ds(png_root, csv_path)
    ld = ds.loader(batch_size=16, shuffle=False)
    assert len(ld) == 4
    n0, seen = ops.LAUNCHES, 0
    for bi, ((x, l1), (s, l2)) in enumerate(ld):
        assert x.is_cuda and x.dtype == torch.float32 and x.shape[1:] == (1, 128, 128)
        for j in range(x.shape[0]):
            (a, la), (b, lb) = ds[bi * 16 + j]
            assert torch.equal(x[j].cpu(), a) and torch.equal(s[j].cpu(), b) and (l1[j], l2[j]) == (la, lb)
        seen += x.shape[0]
    assert seen == 50 and ops.LAUNCHES - n0 == 2 * len(ld)
    g = torch.Generator().manual_seed(1)
    shuffled = torch.cat([x.cpu() for (x, _), _ in ds.loader(16, shuffle=True, generator=g)])
    assert shuffled.shape[0] == 50 and not torch.equal(shuffled[:16], torch.stack([ds[i][0][0] for i in range(16)]))
    sd_ = SpectrogramDataset({"processed_spectograms_dataset_folderpath": ds.store})
    xb, cb = next(iter(sd_.loader(8)))
    assert xb.shape == (8, 1, 128, 128) and torch.equal(xb[3].cpu(), sd_[3][0]) and int(cb[3]) == sd_[3][1]


@pytest.mark.gpu
def test_pair_loader_feeds_train_epoch(png_root, tmp_path, sd):
    """§8(f) rank 3 end to end: the GPU-resident pair loader (elements already on the device) through LDMTrainer.train_epoch
    (train.py:209-243; ops.device_prefetch passes device tensors through) against a hand-written loop of train_step over the
    same batches -- four epoch averages, the ragged last batch of 2 included"""
    import numpy as np
    from music_style_transfer_ldm_b200.config import config
    from music_style_transfer_ldm_b200.dataset import SpectrogramPairDataset
    from test_gpu_train import _trainer
    csv_path = str(tmp_path / "pairs.csv")
    SpectrogramPairDataset.generate_pairings(png_root, csv_path, num_pairs=10)
    ds = _ds(png_root, csv_path)
    _, tr1 = _trainer(sd, "fp32", 1e-4, 3.0)
    tr1.train_loader = ds.loader(batch_size=4, shuffle=False)
    torch.manual_seed(3); torch.cuda.manual_seed(3)
    avg = tr1.train_epoch(0)
    _, tr2 = _trainer(sd, "fp32", 1e-4, 3.0)
    torch.manual_seed(3); torch.cuda.manual_seed(3)
    tot, n = np.zeros(4), 0
    for (c, _), (s, _) in ds.loader(batch_size=4, shuffle=False):
        out = tr2.train_step(c, s)
        tot += [out["total_loss"], out["compression_loss"], out["denoisinsg_loss"], out["style_loss"]]
        n += 1
    assert n == 3 and tr1.step_count == 3
    np.testing.assert_allclose(avg, tot / n * config["training_iteration_noise"], rtol=1e-4)
