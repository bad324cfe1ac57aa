"""Synthetic MicroBooNE-like wire-plane crops.

There is no network and no ROOT on the build/GPU boxes, so the LArCV data interface
(`larcvdataset`, out of scope) is replaced by a seeded generator that mimics what the
reference's fillers deliver to `train()` (training/train_ubresnet2018_wlarcv2.py:576-615):
  image  (B,1,H,W) float32  thresholded ADC (ADC_THRESH=10, :63), ~1-3 % non-zero
  label  (B,H,W)   int64    0 background, 1 track, 2 shower
  weight (B,H,W)   float32  class-balancing pixel weights in the spirit of
                            caffe/larcv/app/ImageMod/SegWeightTrackShower.cxx:109-272
                            (1 + N/n_class on class pixels, N/n_boundary on background
                            within 2 px of a labelled pixel, clipped at 10000).
Host-side numpy; deterministic for a given (seed, index).
"""
import numpy as np
import torch

BASE_SEED = 20180721
ADC_THRESH = 10.0
WEIGHT_MAX = 10000.0


def _draw_track(img, lab, rng, h, w):
    p0 = rng.uniform(0, [h - 1, w - 1])
    p1 = rng.uniform(0, [h - 1, w - 1])
    n = int(max(abs(p1 - p0)) * 1.5) + 2
    t = np.linspace(0.0, 1.0, n)
    rr = np.rint(p0[0] + (p1[0] - p0[0]) * t).astype(np.int64)
    cc = np.rint(p0[1] + (p1[1] - p0[1]) * t).astype(np.int64)
    adc = rng.uniform(20.0, 200.0)
    for d in (-1, 0, 1):                      # 3 px wide
        c2 = np.clip(cc + d, 0, w - 1)
        img[rr, c2] = adc * rng.uniform(0.8, 1.2, size=rr.shape)
        lab[rr, c2] = 1


def _draw_shower(img, lab, rng, h, w, npts=600, sigma=14.0):
    ctr = rng.uniform([0.1 * h, 0.1 * w], [0.9 * h, 0.9 * w])
    pts = rng.normal(ctr, sigma, size=(npts, 2))
    rr = np.clip(np.rint(pts[:, 0]).astype(np.int64), 0, h - 1)
    cc = np.clip(np.rint(pts[:, 1]).astype(np.int64), 0, w - 1)
    img[rr, cc] = rng.uniform(10.0, 120.0, size=npts)
    lab[rr, cc] = 2


def _dilate2(mask):
    """binary dilation by a (5x5) square == 'within DistSurrounding=2 px'."""
    m = mask.copy()
    for _ in range(2):
        n = m.copy()
        n[1:, :] |= m[:-1, :]; n[:-1, :] |= m[1:, :]
        n[:, 1:] |= m[:, :-1]; n[:, :-1] |= m[:, 1:]
        n[1:, 1:] |= m[:-1, :-1]; n[:-1, :-1] |= m[1:, 1:]
        n[1:, :-1] |= m[:-1, 1:]; n[:-1, 1:] |= m[1:, :-1]
        m = n
    return m


def pixel_weights(lab):
    h, w = lab.shape
    npix = float(h * w)
    wgt = np.ones((h, w), dtype=np.float32)
    labelled = lab > 0
    for c in (1, 2):
        m = lab == c
        n = int(m.sum())
        if n:
            wgt[m] += np.floor(npix / n)
    near = _dilate2(labelled) & ~labelled
    nb = int(near.sum())
    if nb:
        wgt[near] += np.floor(npix / nb)
    return np.minimum(wgt, WEIGHT_MAX)


def make_crop(seed, h=512, w=512, ntracks=6, nshowers=3):
    rng = np.random.default_rng(seed)
    img = np.zeros((h, w), dtype=np.float32)
    lab = np.zeros((h, w), dtype=np.int64)
    scale = (h * w) / float(512 * 512)
    for _ in range(max(1, int(round(ntracks * max(scale, 0.25))))):
        _draw_track(img, lab, rng, h, w)
    for _ in range(max(1, int(round(nshowers * max(scale, 0.25))))):
        _draw_shower(img, lab, rng, h, w, npts=max(40, int(600 * min(1.0, scale * 2))))
    low = img < ADC_THRESH
    img[low] = 0.0
    lab[low] = 0
    return img, lab, pixel_weights(lab)


def make_batch(batch, h=512, w=512, rank=0, step=0, dense=False):
    """Returns CPU tensors (image (B,1,H,W) f32, label (B,H,W) i64, weight (B,H,W) f32)."""
    if dense:   # worst-case numerics variant: randn image, uniform labels, rand weights
        g = torch.Generator().manual_seed(BASE_SEED + 7919 * rank + step)
        img = torch.randn(batch, 1, h, w, generator=g)
        lab = torch.randint(0, 3, (batch, h, w), generator=g)
        wgt = torch.rand(batch, h, w, generator=g) + 0.5
        return img, lab, wgt
    imgs = np.zeros((batch, 1, h, w), dtype=np.float32)
    labs = np.zeros((batch, h, w), dtype=np.int64)
    wgts = np.zeros((batch, h, w), dtype=np.float32)
    for b in range(batch):
        seed = BASE_SEED + rank * 1000003 + step * 4099 + b
        imgs[b, 0], labs[b], wgts[b] = make_crop(seed, h, w)
    return torch.from_numpy(imgs), torch.from_numpy(labs), torch.from_numpy(wgts)


def make_wholeview(seed=BASE_SEED, rows=1008, cols=3456, planes=3):
    """(planes, rows, cols) float32 event with 40 tracks / 12 showers per plane."""
    out = np.zeros((planes, rows, cols), dtype=np.float32)
    for p in range(planes):
        rng = np.random.default_rng(seed + 31 * p)
        img = out[p]
        lab = np.zeros((rows, cols), dtype=np.int64)
        for _ in range(40):
            _draw_track(img, lab, rng, rows, cols)
        for _ in range(12):
            _draw_shower(img, lab, rng, rows, cols)
        img[img < ADC_THRESH] = 0.0
    return torch.from_numpy(out)
