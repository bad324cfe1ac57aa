"""Device-side batch assembly: the reference feeds the network from C++ host threads
(`BatchFillerImage2D`, `SegFiller`, `SegWeightTrackShower`; "GPU-Util 0 % ... threadfiller isn't very efficient",
training/grid_scripts/README.md:269).  Here the host only ships the raw larcv planes -- column-major pixels, 1-byte
ROI-type labels -- and the transpose / mirror, the ROI-type -> class table with the `-1` shift, and the class-balancing
pixel weights are computed on the GPU (csrc/batch_fill.cu).  Per 32-crop 512x512 batch that is 42 MB over PCIe instead
of 134 MB (fp32 image + int64 label + fp32 weight).
"""
import torch

from . import _lib
from .ops import _ptr, _stream, require_cuda

K_ROI_TYPE_MAX = 10
_DTYPE_CODE = {torch.float32: 0, torch.float16: 1, torch.uint8: 2, torch.int64: 3}


def roitype_to_class(class_type_list, class_type_def):
    """The table SegFiller builds from its configuration (caffe/larcv/app/APICaffe/SegFiller.cxx:49-63);
    0xFF marks an invalid ROI type."""
    if len(class_type_def) != K_ROI_TYPE_MAX:
        raise ValueError("ClassTypeDef length is %d but it needs to be length kROITypeMax (%d)" %
                         (len(class_type_def), K_ROI_TYPE_MAX))
    lut = [0xFF] * K_ROI_TYPE_MAX
    lut[0] = 0
    for i, t in enumerate(class_type_list):
        if not 0 <= t < K_ROI_TYPE_MAX:
            raise ValueError("ClassTypeList contains type %d which is not a valid ROIType_t" % t)
        lut[t] = i + 1
    for i in range(K_ROI_TYPE_MAX):
        if lut[i] == 0xFF:
            lut[i] = lut[class_type_def[i]]
    return lut


class DeviceBatchFiller:
    """fill(image_planes, label_planes[, mirror]) -> (image (B,1,H,W) f32, label (B,H,W) i64, weight (B,H,W) f32), all on
    the device.  Inputs are (B, cols*rows) / (B, cols, rows) column-major larcv planes, host (pinned) or device."""

    def __init__(self, device, class_type_list=(0, 1, 2), class_type_def=(0, 0, 0, 2, 2, 2, 1, 1, 1, 1), label_shift=-1,
                 dist_surrounding=2, weight_surrounding=True, pool_type=0, weight_max=10000.0):
        self.device = torch.device(device)
        self.lut = torch.tensor(roitype_to_class(class_type_list, class_type_def), dtype=torch.uint8, device=self.device)
        self.shift = int(label_shift)
        # weights are made from the FILLED class image (1 byte per pixel, mirrored like the image): class -> 1 shower /
        # 2 track for the classes the shower / track ROI types map to (ClassTypeDef's groups)
        cat = [0] * 16
        for roi, cls in enumerate(roitype_to_class(class_type_list, class_type_def)):
            if roi in (3, 4, 5) and cls < 16:
                cat[cls] = 1
            if roi in (6, 7, 8, 9) and cls < 16:
                cat[cls] = 2
        self.cat_lut = torch.tensor(cat, dtype=torch.uint8, device=self.device)
        self.dist, self.wsurr, self.pool, self.wmax = int(dist_surrounding), int(bool(weight_surrounding)), int(pool_type), float(weight_max)
        self._bad = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._ws = None

    def fill(self, image_planes, label_planes, rows, cols, mirror=None, check=False, out=None):
        """out: optional (image, label, weight) device tensors to fill in place (e.g. the static inputs of a CUDA graph)"""
        lib = _lib.load()
        img = image_planes.to(self.device, non_blocking=True)
        lab = label_planes.to(self.device, non_blocking=True)
        require_cuda(img, "image planes")
        b = img.shape[0]
        assert img.numel() == b * rows * cols and lab.numel() == b * rows * cols
        mir = None
        if mirror is not None:
            mir = torch.as_tensor(mirror, dtype=torch.uint8).to(self.device, non_blocking=True)
        if out is not None:
            out_img, out_lab, out_w = out
            assert out_img.shape == (b, 1, rows, cols) and out_img.dtype == torch.float32 and out_img.is_contiguous()
            assert out_lab.shape == (b, rows, cols) and out_lab.dtype == torch.int64 and out_lab.is_contiguous()
            assert out_w.shape == (b, rows, cols) and out_w.dtype == torch.float32 and out_w.is_contiguous()
        else:
            out_img = torch.empty((b, 1, rows, cols), dtype=torch.float32, device=self.device)
            out_lab = torch.empty((b, rows, cols), dtype=torch.int64, device=self.device)
            out_w = torch.empty((b, rows, cols), dtype=torch.float32, device=self.device)
        if getattr(self, "_cls8", None) is None or self._cls8.shape != (b, rows, cols):
            self._cls8 = torch.empty((b, rows, cols), dtype=torch.uint8, device=self.device)
        cls8 = self._cls8
        _lib.check(lib.ubr_fill_image_batch(img.contiguous().data_ptr(), _DTYPE_CODE[img.dtype], b, rows, cols, _ptr(mir),
                                            out_img.data_ptr(), _stream()), "fill_image_batch")
        _lib.check(lib.ubr_fill_label_batch(lab.contiguous().data_ptr(), _DTYPE_CODE[lab.dtype], b, rows, cols, _ptr(mir),
                                            self.lut.data_ptr(), self.lut.numel(), self.shift, out_lab.data_ptr(),
                                            cls8.data_ptr(), self._bad.data_ptr(), _stream()), "fill_label_batch")
        need = int(lib.ubr_seg_weights_workspace_bytes(b, rows, cols))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        _lib.check(lib.ubr_seg_weights(cls8.data_ptr(), 2, 0, b, rows, cols, self.cat_lut.data_ptr(), self.cat_lut.numel(),
                                       self.dist, self.wsurr, self.pool, self.wmax, 0, out_w.data_ptr(), self._ws.data_ptr(),
                                       self._ws.numel(), _stream()), "seg_weights")
        if check and int(self._bad.item()):          # one host sync; the C++ filler throws on the first bad pixel
            raise RuntimeError("Found invalid ROI type in the label image (%d pixels)" % int(self._bad.item()))
        return out_img, out_lab, out_w


def seg_weights(label, cat_lut, dist_surrounding=2, weight_surrounding=True, pool_type=0, weight_max=10000.0, weight_in=None,
                column_major=False, rows=None, cols=None):
    """SegWeightTrackShower on a device label batch (B, rows, cols) of dtype f32 / u8 / i64 -> (B, rows, cols) f32."""
    lib = _lib.load()
    require_cuda(label, "label")
    b = label.shape[0]
    rows = rows or label.shape[1]
    cols = cols or label.shape[2]
    out = torch.empty((b, rows, cols), dtype=torch.float32, device=label.device)
    lut = torch.as_tensor(cat_lut, dtype=torch.uint8).to(label.device)
    ws = torch.empty(int(lib.ubr_seg_weights_workspace_bytes(b, rows, cols)), dtype=torch.uint8, device=label.device)
    _lib.check(lib.ubr_seg_weights(label.contiguous().data_ptr(), _DTYPE_CODE[label.dtype], int(column_major), b, rows, cols,
                                   lut.data_ptr(), lut.numel(), int(dist_surrounding), int(bool(weight_surrounding)),
                                   int(pool_type), float(weight_max), _ptr(weight_in), out.data_ptr(), ws.data_ptr(),
                                   ws.numel(), _stream()), "seg_weights")
    return out
