"""Device-side batch assembly: oracle restatement vs its literal-loop form (CPU) and the CUDA kernels vs the
oracle, bit-exact (GPU)."""
import numpy as np
import pytest
import torch

from oracle import filler_port as fp
from ubresnet_b200 import fill, synthetic


def _roi_planes(batch, rows, cols, seed=3):
    """column-major larcv-style planes: ADC image (f32) + ROI-type label (u8: track -> kROIMuminus/kROIProton, shower ->
    kROIEminus/kROIGamma) from the synthetic generator"""
    rng = np.random.default_rng(seed)
    imgs, labs = [], []
    for b in range(batch):
        img, lab, _ = synthetic.make_crop(seed * 100 + b, rows, cols)
        roi = np.zeros_like(lab, dtype=np.uint8)
        roi[lab == 1] = rng.choice([6, 9], size=int((lab == 1).sum()))
        roi[lab == 2] = rng.choice([3, 4], size=int((lab == 2).sum()))
        imgs.append(np.ascontiguousarray(img.T).reshape(-1))       # column-major: index = col*rows + row
        labs.append(np.ascontiguousarray(roi.T).reshape(-1))
    return np.stack(imgs), np.stack(labs)


def test_port_vectorised_equals_literal_loops():
    rng = np.random.default_rng(1)
    rows, cols = 24, 40
    px = rng.normal(size=rows * cols).astype(np.float32)
    for mirror in (False, True):
        assert (fp.fill_image(px, rows, cols, mirror) == fp.fill_image_loops(px, rows, cols, mirror)).all()
    lab = np.zeros((rows, cols), dtype=np.int64)
    lab[3:5, 4:30] = 6
    lab[10:14, 20:24] = 4
    lab[23, 39] = 9
    lab[0, 0] = 3
    assert (fp.seg_weights(lab) == fp.seg_weights_loops(lab)).all()
    assert (fp.seg_weights(np.zeros((8, 8), dtype=np.int64)) == 1.0).all()            # empty image: weight image painted 1


def test_roitype_table_matches_training_config():
    # ClassTypeList [0,1,2], ClassTypeDef [0,0,0,2,2,2,1,1,1,1] (training/train_ubresnet2018_wlarcv1.py:166-167)
    lut = fp.roitype_to_class([0, 1, 2], [0, 0, 0, 2, 2, 2, 1, 1, 1, 1])
    assert list(lut) == [1, 2, 3, 3, 3, 3, 2, 2, 2, 2]
    assert fill.roitype_to_class([0, 1, 2], [0, 0, 0, 2, 2, 2, 1, 1, 1, 1]) == list(lut)
    with pytest.raises(ValueError):
        fill.roitype_to_class([0, 1, 2], [0, 0, 0])


def test_synthetic_weights_are_segweighttrackshower():
    """the generator's pixel weights (ubresnet_b200/synthetic.py) equal the restated SegWeightTrackShower on its labels"""
    for seed in (5, 6):
        _, lab, wgt = synthetic.make_crop(seed, 128, 128)
        roi = np.where(lab == 1, 6, np.where(lab == 2, 3, 0))
        assert (fp.seg_weights(roi) == wgt).all()


@pytest.mark.gpu
@pytest.mark.parametrize("rows,cols,batch", [(64, 96, 3), (512, 512, 4), (40, 72, 2)])
def test_device_batch_filler_bit_exact(rows, cols, batch):
    imgs, labs = _roi_planes(batch, rows, cols)
    mirror = [bool(b % 2) for b in range(batch)]
    lut = fp.roitype_to_class([0, 1, 2], [0, 0, 0, 2, 2, 2, 1, 1, 1, 1])
    filler = fill.DeviceBatchFiller("cuda:0")
    img, lab, wgt = filler.fill(torch.from_numpy(imgs), torch.from_numpy(labs), rows, cols, mirror=mirror, check=True)
    for b in range(batch):
        ref_img = fp.fill_image(imgs[b], rows, cols, mirror[b])
        ref_lab = fp.fill_label(labs[b], rows, cols, lut, mirror[b], shift=-1)
        ref_w = fp.seg_weights(ref_lab, shower=(2,), track=(1,))
        assert (img[b, 0].cpu().numpy() == ref_img).all()
        assert (lab[b].cpu().numpy() == ref_lab).all()
        assert (wgt[b].cpu().numpy() == ref_w).all()
    # fp16 / uint8 wire formats widen exactly
    h = torch.from_numpy(imgs).half()
    img16, _, _ = filler.fill(h, torch.from_numpy(labs), rows, cols)
    assert (img16[0, 0].cpu().numpy() == fp.fill_image(h[0].float().numpy(), rows, cols)).all()


@pytest.mark.gpu
def test_seg_weights_options_and_invalid_roi():
    rows, cols = 48, 80
    _, labs = _roi_planes(2, rows, cols, seed=9)
    roi_rm = np.stack([l.reshape(cols, rows).T for l in labs])                      # row-major ROI types
    cat = [0, 0, 0, 1, 1, 1, 2, 2, 2, 2]
    w_in = np.random.default_rng(2).uniform(0.5, 3.0, size=(2, rows, cols)).astype(np.float32)
    for pool in (0, 1, 2, 3):
        for dist, ws in ((2, True), (1, True), (2, False)):
            got = fill.seg_weights(torch.from_numpy(roi_rm).cuda(), cat, dist, ws, pool, 50.0,
                                   weight_in=torch.from_numpy(w_in).cuda())
            for b in range(2):
                ref = fp.seg_weights(roi_rm[b], dist=dist, weight_surrounding=ws, pool=pool, weight_max=50.0, weight_in=w_in[b])
                assert (got[b].cpu().numpy() == ref).all(), (pool, dist, ws)
    # column-major label input gives the same weights
    got_cm = fill.seg_weights(torch.from_numpy(labs).cuda().view(2, cols, rows), cat, column_major=True, rows=rows, cols=cols)
    got_rm = fill.seg_weights(torch.from_numpy(roi_rm).cuda(), cat)
    assert torch.equal(got_cm, got_rm)
    bad = labs.copy()
    bad[0, 5] = 77
    filler = fill.DeviceBatchFiller("cuda:0")
    imgs = np.zeros((2, rows * cols), dtype=np.float32)
    with pytest.raises(RuntimeError):
        filler.fill(torch.from_numpy(imgs), torch.from_numpy(bad), rows, cols, check=True)


@pytest.mark.gpu
def test_assembled_prefetcher_yields_the_batches_in_order():
    """prefetch.AssembledPrefetcher: copy + assembly one step ahead on its own stream into two alternating staging sets.
    Every yielded (image, label, weight) must equal the synchronous fill of the same host batch, also when the consumer
    keeps the main stream busy between calls (a staging set may only be refilled after its consumer's work)."""
    from ubresnet_b200.prefetch import AssembledPrefetcher
    rows, cols, batch, nb = 64, 128, 2, 5
    host = []
    for i in range(nb):
        imgs, labs = _roi_planes(batch, rows, cols, seed=100 + i)
        host.append((torch.from_numpy(imgs).pin_memory(), torch.from_numpy(labs).pin_memory()))
    filler = fill.DeviceBatchFiller("cuda:0")
    want = [tuple(t.clone() for t in filler.fill(p, r, rows, cols)) for p, r in host]
    torch.cuda.synchronize()
    pf = AssembledPrefetcher(iter(host), fill.DeviceBatchFiller("cuda:0"), rows, cols, batch, "cuda:0")
    busy = torch.zeros(1 << 22, device="cuda:0")
    got = []
    for img, lab, wgt in pf:
        got.append((img.clone(), lab.clone(), wgt.clone()))     # the consumer's work on this set (main stream)
        for _ in range(20):
            busy.add_(1.0)                                       # keep the main stream behind the prefetch stream
    torch.cuda.synchronize()
    assert len(got) == nb
    for (a, b, c), (x, y, z) in zip(got, want):
        assert torch.equal(a, x) and torch.equal(b, y) and torch.equal(c, z)
