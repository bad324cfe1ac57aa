"""ctypes binding of libubresnet_b200.so (the C ABI in include/ubresnet_b200.h).

The library is the ONLY compute path: if it is missing or the device is not a B200 every op raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBPATH = os.path.join(_HERE, "lib", "libubresnet_b200.so")


class View(C.Structure):
    """ubr_view: NHWC bf16 channel slice."""
    _fields_ = [("ptr", C.c_void_p), ("n", C.c_int32), ("h", C.c_int32), ("w", C.c_int32), ("c", C.c_int32),
                ("ld", C.c_int32)]


_vp, _i, _u64, _i64, _f, _d = C.c_void_p, C.c_int, C.c_uint64, C.c_int64, C.c_float, C.c_double

# name -> (restype, argtypes).  Mirrors include/ubresnet_b200.h one to one.
SIGNATURES = {
    "ubr_abi_version": (_i, []),
    "ubr_last_error": (C.c_char_p, []),
    "ubr_device_check": (_i, []),
    "ubr_sync_check": (_i, [_vp]),
    "ubr_shutdown": (_i, []),
    "ubr_launch_count": (_u64, []),
    "ubr_trace_kernels": (None, [_i]),
    "ubr_take_kernel_names": (C.c_char_p, []),
    "ubr_pack_weights": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "ubr_pack_weights_batch": (_i, [_vp, _i, _vp]),
    "ubr_conv2d_fwd": (_i, [View, _vp, _vp, _i, _i, _i, View, _vp, _i, _vp]),
    "ubr_conv2d_fwd_fused": (_i, [View, _vp, _vp, _i, _i, _i, _i, View, _i, View, _vp]),
    "ubr_block_chain_fwd": (_i, [View, _vp, _vp, _vp, _vp, _i, View, View, _vp]),
    "ubr_conv2d_fwd_stats": (_i, [View, _vp, _vp, _i, _i, _i, View, _vp, _vp, _vp]),
    "ubr_conv2d_fwd_bn": (_i, [View, _vp, _vp, _i, _i, _i, View, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp]),
    "ubr_conv2d_dgrad": (_i, [View, _vp, _i, _i, _i, View, _vp]),
    "ubr_conv2d_dgrad_bnreduce_final": (_i, [View, _vp, _i, _i, _i, View, View, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp,
                                             _i, _vp]),
    "ubr_conv2d_dgrad_bnreduce": (_i, [View, _vp, _i, _i, _i, View, View, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "ubr_bn_act_bwd_apply_rows": (_i, [View, View, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _d, View, _vp, _vp, _i, _vp]),
    "ubr_conv2d_dgrad_acc": (_i, [View, _vp, _i, _i, _i, View, _vp]),
    "ubr_conv2d_wgrad": (_i, [View, View, _i, _i, _i, _vp, _i, _vp, _u64, _vp]),
    "ubr_wgrad_workspace_bytes": (_u64, [_i, _i, _i, _i, _i, _i]),
    "ubr_deconv4s2_fwd": (_i, [View, _vp, _i, View, _vp]),
    "ubr_conv2d_dilated_fwd": (_i, [View, _vp, _vp, _i, _i, View, _vp]),
    "ubr_conv2d_dilated_fwd_relu": (_i, [View, _vp, _vp, _i, _i, View, _vp]),
    "ubr_conv2d_dilated_dgrad": (_i, [View, _vp, _i, _i, View, _vp]),
    "ubr_conv2d_dilated_wgrad": (_i, [View, View, _i, _i, _vp, _i, _vp, _u64, _vp]),
    "ubr_maxpool3s1_fwd": (_i, [View, View, _vp, _vp]),
    "ubr_maxpool3s1_bwd_idx": (_i, [_vp, View, View, _i, _vp]),
    "ubr_pixel_unshuffle2": (_i, [View, View, _vp]),
    "ubr_deconv4s2_wgrad_gather": (_i, [_vp, _i, _i, _vp, _i, _vp]),
    "ubr_deconv4s2_fwd_subpixel": (_i, [View, _vp, _i, View, _vp]),
    "ubr_deconv4s2_dgrad": (_i, [View, _vp, _i, View, _vp]),
    "ubr_deconv4s2_wgrad": (_i, [View, View, _vp, _i, _vp, _u64, _vp]),
    "ubr_stem_conv_fwd": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, View, _vp]),
    "ubr_stem_conv_fwd_relu": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, View, _vp]),
    "ubr_stem_conv_wgrad": (_i, [_vp, _i, _i, _i, View, _vp, _vp, _i, _vp, _u64, _vp]),
    "ubr_stem_wgrad_workspace_bytes": (_u64, [_i, _i, _i, _i]),
    "ubr_bn_workspace_bytes": (_u64, [_i]),
    "ubr_bn_stats": (_i, [View, _vp, _vp]),
    "ubr_bn_tree_bytes": (_u64, [_i]),
    "ubr_bn_stats_final": (_i, [View, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp]),
    "ubr_bn_finalize": (_i, [_vp, _i, _d, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp]),
    "ubr_bn_finalize_rows": (_i, [_vp, _i, _i, _d, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp]),
    "ubr_bn_eval_affine": (_i, [_i, _vp, _vp, _vp, _vp, _f, _vp, _vp, _vp]),
    "ubr_bn_act": (_i, [View, _vp, _vp, _i, View, _vp]),
    "ubr_block_tail": (_i, [View, _vp, _vp, View, _vp, _vp, View, _vp]),
    "ubr_bn_act_bwd_reduce": (_i, [View, View, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "ubr_bn_act_bwd_apply": (_i, [View, View, _vp, _vp, _vp, _vp, _vp, _i, _vp, _d, View, _vp, _vp, _i, _vp]),
    "ubr_bn_act_bwd_reduce_final": (_i, [View, View, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp]),
    "ubr_bn_act_bwd_apply_final": (_i, [View, View, _vp, _vp, _vp, _vp, _i, _vp, _d, View, _vp]),
    "ubr_block_tail_bwd_reduce_final": (_i, [View, View, View, _vp, _vp, _vp, _vp, View, _vp, _vp, _vp, _vp, _i, _vp, _vp,
                                             _vp, _vp, _vp, _vp, _i, _vp]),
    "ubr_block_tail_bwd_apply_final": (_i, [View, View, View, _vp, _vp, _vp, _vp, View, _vp, _vp, _vp, _vp, _vp, _i, _vp, _d,
                                            View, View, _vp]),
    "ubr_block_tail_bwd_reduce": (_i, [View, View, View, _vp, _vp, _vp, _vp, View, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "ubr_block_tail_bwd_apply": (_i, [View, View, View, _vp, _vp, _vp, _vp, _vp, View, _vp, _vp, _vp, _vp, _vp, _i, _vp, _d,
                                      View, View, _vp, _vp, _vp, _vp, _i, _vp]),
    "ubr_maxpool3s2_fwd": (_i, [View, View, _vp]),
    "ubr_maxpool3s2_bwd": (_i, [View, View, View, _i, _vp]),
    "ubr_maxpool3s2_fwd_idx": (_i, [View, View, _vp, _vp]),
    "ubr_maxpool3s2_bwd_idx": (_i, [_vp, View, View, _i, _vp]),
    "ubr_add": (_i, [View, View, View, View, _vp]),
    "ubr_logsoftmax_fwd": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "ubr_logsoftmax_bwd": (_i, [_vp, _vp, _i, _i, _i, View, _vp]),
    "ubr_nll_workspace_bytes": (_u64, [_i, _i]),
    "ubr_pixelwise_nll_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _vp, _vp, _vp]),
    "ubr_pixelwise_nll_bwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _vp, _vp]),
    "ubr_tile_gather": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _vp]),
    "ubr_stitch_scatter": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _vp, _vp]),
    "ubr_argmax_classes": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "ubr_stitch_scatter_u8": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "ubr_tile_gather_rect": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _vp, _vp]),
    "ubr_stitch_scatter_rect": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp, _vp]),
    "ubr_accuracy_counts": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "ubr_scale_f32": (_i, [_vp, _u64, _f, _vp]),
    "ubr_channel_sum": (_i, [View, _i, _vp, _vp, _i, _vp]),
    "ubr_channel_sum_static": (_i, [View, _i, _vp, _i, _vp]),
    "ubr_increment_i64": (_i, [_vp, _i, _vp]),
    "ubr_fill_image_batch": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "ubr_fill_label_batch": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "ubr_seg_weights_workspace_bytes": (_u64, [_i, _i, _i]),
    "ubr_seg_weights": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _i, _i, _i, _f, _vp, _vp, _vp, _u64, _vp]),
    "ubr_f32_conv2d_fwd": (_i, [View, _vp, _vp, _i, _i, _i, _i, _i, View, _vp]),
    "ubr_f32_conv2d_dgrad": (_i, [View, _vp, _i, _i, _i, _i, _i, View, _i, _vp]),
    "ubr_f32_conv2d_wgrad": (_i, [View, View, _i, _i, _i, _i, _vp, _i, _vp, _u64, _vp]),
    "ubr_f32_deconv_fwd": (_i, [View, _vp, _i, _i, _i, _i, View, _vp]),
    "ubr_f32_deconv_dgrad": (_i, [View, _vp, _i, _i, _i, _i, View, _vp]),
    "ubr_f32_deconv_wgrad": (_i, [View, View, _i, _i, _i, _vp, _i, _vp, _u64, _vp]),
    "ubr_f32_bn_stats": (_i, [View, _vp, _vp]),
    "ubr_f32_bn_finalize": (_i, [_vp, _i, _d, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp]),
    "ubr_f32_channel_sum": (_i, [View, _vp, _vp, _i, _vp]),
    "ubr_f32_bn_act": (_i, [View, _vp, _vp, _i, View, _vp]),
    "ubr_f32_block_tail": (_i, [View, _vp, _vp, View, _vp, _vp, View, _vp]),
    "ubr_f32_relu_mask": (_i, [View, View, View, _vp]),
    "ubr_f32_bn_act_bwd": (_i, [View, View, _vp, _vp, _vp, _vp, _i, _vp, _d, View, _vp, _vp, _i, _vp]),
    "ubr_f32_maxpool3s2_fwd": (_i, [View, View, _vp, _vp]),
    "ubr_f32_maxpool3s2_bwd": (_i, [_vp, View, View, _i, _vp]),
    "ubr_f32_add": (_i, [View, View, View, View, _vp]),
    "ubr_f32_logsoftmax_bwd": (_i, [_vp, _vp, _i, _i, _i, View, _vp]),
    "ubr_opt_chunk_elems": (_i, []),
    "ubr_sgd_step": (_i, [_vp, _vp, _i, _f, _f, _f, _f, _i, _i, _vp]),
    "ubr_adam_step": (_i, [_vp, _vp, _i, _d, _d, _d, _d, _d, _i, _vp]),
}

_lib = None


def load():
    """dlopen the in-tree library; raises (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIBPATH):
        raise RuntimeError(
            "ubresnet_b200: %s is missing -- run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU/PyTorch fallback)" % LIBPATH)
    lib = C.CDLL(LIBPATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.ubr_abi_version() != 1:
        raise RuntimeError("ubresnet_b200: ABI version mismatch")
    _lib = lib
    return lib


class TimedLib:
    """Profiling proxy: brackets every C-ABI call with CUDA events on the launching (current) stream and notes which
    kernels the call launched (ubr_take_kernel_names) plus its View / integer arguments.
    Used by bench.py for the live per-kernel timings behind the `roofline` object."""

    _SKIP = ("ubr_last_error", "ubr_abi_version", "ubr_launch_count", "ubr_trace_kernels", "ubr_take_kernel_names")

    def __init__(self, lib):
        self._lib = lib
        self.records = []          # (function name, start event, end event, kernel names, args summary)
        lib.ubr_trace_kernels(1)

    def __getattr__(self, name):
        fn = getattr(self._lib, name)
        if not name.startswith("ubr_") or name in self._SKIP or name.endswith("_bytes"):
            return fn
        import torch
        take = self._lib.ubr_take_kernel_names

        def timed(*args):
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            take()
            e0.record()
            rc = fn(*args)
            e1.record()
            kern = take().decode("ascii", "replace")
            summ = tuple((a.n, a.h, a.w, a.c) if isinstance(a, View) else a for a in args)      # positional, raw
            self.records.append((name, e0, e1, kern, summ))
            return rc
        return timed

    def summary(self):
        """{function name: (calls, total ms)} -- call after torch.cuda.synchronize()."""
        out = {}
        for name, e0, e1, _, _ in self.records:
            c, t = out.get(name, (0, 0.0))
            out[name] = (c + 1, t + e0.elapsed_time(e1))
        return out

    def by_kernel(self):
        """{kernel names launched by one call ("a+b"): [(function, args summary, ms), ...]}."""
        out = {}
        for name, e0, e1, kern, summ in self.records:
            out.setdefault(kern, []).append((name, summ, e0.elapsed_time(e1)))
        return out


def enable_profiling():
    global _lib
    lib = load()
    if not isinstance(lib, TimedLib):
        _lib = TimedLib(lib)
    return _lib


def disable_profiling():
    global _lib
    if isinstance(_lib, TimedLib):
        _lib._lib.ubr_trace_kernels(0)
        _lib = _lib._lib


def last_error():
    return load().ubr_last_error().decode("utf-8", "replace")


def check(rc, what):
    if rc != 0:
        raise RuntimeError("ubresnet_b200.%s failed (%d): %s" % (what, rc, last_error()))
