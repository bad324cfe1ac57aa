"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's C++ batch fillers.

The reference assembles every training batch on host threads in C++ (needs ROOT to build: not buildable here, see
DESIGN.md); these functions restate the arithmetic of the three pieces the device kernels replace, citing the lines
they follow (paths relative to /root/reference).  Parity: UNPINNED by reference tests (the fillers have none);
the `_loops` variants below are literal transcriptions of the C++ loops and pin the vectorised forms on small cases.
Never imported from `ubresnet_b200/`.
"""
import numpy as np

K_ROI_TYPE_MAX = 10            # caffe/larcv/core/DataFormat/DataFormatTypes.h:60-71 (kROIUnknown .. kROIProton)
K_INVALID = 0xFF
SHOWER_TYPES = (3, 4, 5)       # kROIEminus, kROIGamma, kROIPizero      (SegWeightTrackShower.cxx:113)
TRACK_TYPES = (6, 7, 8, 9)     # kROIMuminus, kROIKminus, kROIPiminus, kROIProton (:115)


def fill_image(pixels, rows, cols, mirror=False):
    """larcv/larcv/app/ThreadIO/BatchFillerImage2D.cxx:102-112 (index tables) + :217-256 (copy):
    out[row*cols + col] = in[col*rows + row]; mirrored: in[(cols-col-1)*rows + row].  `pixels`: flat column-major."""
    img = np.asarray(pixels).reshape(cols, rows)          # [col][row]
    if mirror:
        img = img[::-1]
    return np.ascontiguousarray(img.T).astype(np.float32)


def fill_image_loops(pixels, rows, cols, mirror=False):
    out = np.zeros(rows * cols, dtype=np.float32)
    k = 0
    for row in range(rows):
        for col in range(cols):
            src = (cols - col - 1) * rows + row if mirror else col * rows + row
            out[k] = pixels[src]
            k += 1
    return out.reshape(rows, cols)


def roitype_to_class(class_type_list, class_type_def):
    """caffe/larcv/app/APICaffe/SegFiller.cxx:49-63: unknown -> 0; ClassTypeList[i] -> i+1; every other ROI type takes the
    class of ClassTypeDef[type].  Returns a uint8 table of kROITypeMax entries (0xFF = invalid)."""
    assert len(class_type_def) == K_ROI_TYPE_MAX
    lut = [K_INVALID] * K_ROI_TYPE_MAX
    lut[0] = 0
    for i, t in enumerate(class_type_list):
        lut[t] = i + 1
    for i in range(K_ROI_TYPE_MAX):
        if lut[i] != K_INVALID:
            continue
        lut[i] = lut[class_type_def[i]]
    return np.array(lut, dtype=np.uint8)


def fill_label(pixels, rows, cols, lut, mirror=False, shift=-1):
    """SegFiller.cxx:339-363 (class_value = _roitype_to_class[(size_t)label_value]) and the `+= -1` of
    training/grid_scripts/train_ubresnet_wlarcv1_tuftsgrid.py:132.  Raises on an invalid ROI type like the C++ (larbys)."""
    lab = np.asarray(pixels).reshape(cols, rows)
    if mirror:
        lab = lab[::-1]
    idx = lab.T.astype(np.int64)
    if (idx < 0).any() or (idx >= len(lut)).any():
        raise ValueError("Found invalid ROI type in the label image")
    cls = np.asarray(lut)[idx]
    if (cls == K_INVALID).any():
        raise ValueError("Found invalid ROI type in the label image")
    return np.ascontiguousarray(cls.astype(np.int64) + shift)


def category(label, shower=SHOWER_TYPES, track=TRACK_TYPES):
    """0 none / 1 shower / 2 track per pixel"""
    lab = np.asarray(label).astype(np.int64)
    cat = np.zeros(lab.shape, dtype=np.uint8)
    cat[np.isin(lab, shower)] = 1
    cat[np.isin(lab, track)] = 2
    return cat


def seg_weights(label, shower=SHOWER_TYPES, track=TRACK_TYPES, dist=2, weight_surrounding=True, pool=0,
                weight_max=10000.0, weight_in=None):
    """caffe/larcv/app/ImageMod/SegWeightTrackShower.cxx:109-272 on a (rows, cols) label image (the window is a square,
    so the result does not depend on the pixel order convention).  pool: 0 sum, 1 max, 2 average, 3 overwrite
    (SegWeightTrackShower.h PoolType_t).  The keypoint term (:189-238, WeightVertex) is not restated (off in the only
    in-tree configuration, caffe/larcv/app/MeatSlicer/supera_segment.cfg:200-210)."""
    cat = category(label, shower, track)
    rows, cols = cat.shape
    n = np.float32(rows * cols)
    ns, nt = int((cat == 1).sum()), int((cat == 2).sum())
    w_shower = np.float32(int(n / np.float32(ns))) if ns else np.float32(1.0)          # :122-125 (int)(float/float)
    w_track = np.float32(int(n / np.float32(nt))) if nt else np.float32(1.0)
    temp = np.zeros((rows, cols), dtype=np.float32)
    temp[cat == 1] = w_shower
    temp[cat == 2] = w_track
    if weight_surrounding and dist:
        lab = cat != 0
        near = np.zeros_like(lab)
        pad = np.pad(lab, dist)
        for dr in range(2 * dist + 1):                                                # :150-176 square window, clipped
            for dc in range(2 * dist + 1):
                near |= pad[dr:dr + rows, dc:dc + cols]
        near &= ~lab
        nb = int(near.sum())
        w_surr = np.float32(int(n / np.float32(nb))) if nb else np.float32(0.0)       # :180-181
        temp = temp + near.astype(np.float32) * w_surr                                # :186-187
    w0 = np.ones((rows, cols), dtype=np.float32) if weight_in is None else np.asarray(weight_in, dtype=np.float32)
    if pool == 0:
        v = w0 + temp
    elif pool == 1:
        v = np.maximum(w0, temp)
    elif pool == 2:
        v = (w0 + temp) / np.float32(2.0)
    else:
        v = temp
    return np.minimum(v, np.float32(weight_max)).astype(np.float32)


def seg_weights_loops(label, shower=SHOWER_TYPES, track=TRACK_TYPES, dist=2, weight_max=10000.0):
    """literal transcription of the C++ loops (sum pooling, weight image painted 1.0) for small images"""
    lab = np.asarray(label).astype(np.int64)
    rows, cols = lab.shape
    is_s = lambda v: v in shower
    is_t = lambda v: v in track
    ns = sum(1 for v in lab.flat if is_s(v))
    nt = sum(1 for v in lab.flat if is_t(v))
    n = np.float32(lab.size)
    w_s = np.float32(int(n / np.float32(ns))) if ns else np.float32(1.0)
    w_t = np.float32(int(n / np.float32(nt))) if nt else np.float32(1.0)
    temp = np.zeros((rows, cols), dtype=np.float32)
    bnd = np.zeros((rows, cols), dtype=np.float32)
    for r in range(rows):
        for c in range(cols):
            v = lab[r, c]
            temp[r, c] = w_s if is_s(v) else (w_t if is_t(v) else 0.0)
    nsur = 0
    for row in range(rows):
        for col in range(cols):
            v = lab[row, col]
            if is_s(v) or is_t(v):
                continue
            flag = False
            tr = row + dist
            while tr + dist >= row:
                if tr >= rows:
                    tr -= 1
                    continue
                if tr < 0:
                    break
                tc = col + dist
                while tc + dist >= col:
                    if tc >= cols:
                        tc -= 1
                        continue
                    if tc < 0:
                        break
                    u = lab[tr, tc]
                    if is_s(u) or is_t(u):
                        flag = True
                        break
                    tc -= 1
                if flag:
                    break
                tr -= 1
            bnd[row, col] = 1.0 if flag else 0.0
            nsur += int(flag)
    w_b = np.float32(int(n / np.float32(nsur))) if nsur else np.float32(0.0)
    out = np.ones((rows, cols), dtype=np.float32) + temp + bnd * w_b
    return np.minimum(out, np.float32(weight_max))
