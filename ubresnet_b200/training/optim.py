"""Multi-tensor SGD / Adam: the optimizers of the reference training scripts with one kernel launch per step.

    train_ubresnet2018_wlarcv1.py:127-129   torch.optim.SGD(model.parameters(), lr, momentum=0.9, weight_decay=1e-3)
    train_ubresnet2018_wlarcv2.py:155-157   torch.optim.Adam(model.parameters(), lr=1e-5, weight_decay=1e-4)

Both subclass `torch.optim.Optimizer` and keep torch's state layout (`momentum_buffer`; `step`, `exp_avg`,
`exp_avg_sq`), so `optimizer.state_dict()` in the reference's checkpoint dict (`train_ubresnet2018_wlarcv2.py:260-266`)
round-trips with the torch classes.  The arithmetic is torch's (L2 weight decay folded into the gradient, Adam bias
correction, no amsgrad / maximize / decoupled decay); all parameters of a group are updated by ONE launch of
`ubr_sgd_step` / `ubr_adam_step` over a device table of pointers.  CUDA fp32 parameters only.
"""
import ctypes

import torch

from .. import _lib


class _Slot(ctypes.Structure):
    _fields_ = [("param", ctypes.c_void_p), ("grad", ctypes.c_void_p), ("state1", ctypes.c_void_p),
                ("state2", ctypes.c_void_p), ("numel", ctypes.c_int64), ("vec_ok", ctypes.c_int32),
                ("pad_", ctypes.c_int32)]


class _Chunk(ctypes.Structure):
    _fields_ = [("slot", ctypes.c_int32), ("pad_", ctypes.c_int32), ("offset", ctypes.c_int64)]


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class _Tables:
    """device copies of the slot / chunk tables of one param group, rebuilt only when a pointer changes"""

    def __init__(self):
        self.key = None
        self.slots = self.chunks = None
        self.nchunks = 0

    def update(self, entries, device):
        key = tuple((p, g, s1, s2, n) for p, g, s1, s2, n in entries)
        if key == self.key:
            return
        chunk = _lib.load().ubr_opt_chunk_elems()
        slots = (_Slot * len(entries))()
        chunks = []
        for i, (p, g, s1, s2, n) in enumerate(entries):
            ptrs = [p, g] + [s for s in (s1, s2) if s]
            slots[i] = _Slot(p, g, s1 or None, s2 or None, n, int(all(q % 16 == 0 for q in ptrs)), 0)
            chunks.extend((i, 0, off) for off in range(0, n, chunk))
        carr = (_Chunk * len(chunks))(*[_Chunk(*c) for c in chunks])

        capturing = torch.cuda.is_current_stream_capturing()
        # pinned staging + asynchronous copy (legal inside CUDA-graph capture, ubresnet_b200.graphs).  Outside capture a
        # ring of 4 staging sets is REUSED -- gradient pointers may change every step (pinning memory costs
        # milliseconds) and the host must not wait for the step that is still running; under capture the source must
        # stay untouched for the life of the graph, so it gets buffers of its own
        ring = getattr(self, "_ring", None)
        if ring is None:
            ring = self._ring = {"i": 0, "sets": [dict() for _ in range(4)], "done": [None] * 4}
        keep = []
        k = ring["i"] % 4
        if not capturing:
            ring["i"] += 1
            if ring["done"][k] is not None:
                ring["done"][k].synchronize()       # the upload issued 4 updates ago has left this set

        def upload(arr, slot):
            raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            if capturing:
                host = raw.pin_memory()
                keep.append(host)
            else:
                stage = ring["sets"][k]
                host = stage.get(slot)
                if host is None or host.numel() < raw.numel():
                    host = stage[slot] = torch.empty(max(raw.numel(), 1 << 16), dtype=torch.uint8).pin_memory()
                host = host[:raw.numel()]
                host.copy_(raw)
            return host.to(device, non_blocking=True)

        self.slots, self.chunks, self.nchunks, self.key = upload(slots, 0), upload(carr, 1), len(chunks), key
        if capturing:
            self._host = getattr(self, "_host", []) + keep
        else:
            ring["done"][k] = torch.cuda.Event()
            ring["done"][k].record()


def _mark_modified(tensors):
    """The kernels write parameters and state through raw pointers: bump the tensors' version counters like an
    in-place torch op would, so that everything keyed on `_version` (the engine's packed bf16 weight cache, autograd's
    saved-tensor checks) sees the update."""
    torch.autograd.graph.increment_version(list(tensors))


class _FusedBase(torch.optim.Optimizer):
    def _collect(self, group, state_names):
        """[(param, grad, state tensors...)] of the parameters that have a gradient"""
        out = []
        for p in group["params"]:
            if p.grad is None:
                continue
            if not (p.is_cuda and p.dtype == torch.float32 and p.grad.dtype == torch.float32):
                raise RuntimeError("ubresnet_b200 optimizers update CUDA fp32 parameters only (got %s %s)"
                                   % (p.device, p.dtype))
            if p.grad.is_sparse or not p.is_contiguous():
                raise RuntimeError("ubresnet_b200 optimizers need dense contiguous parameters and gradients")
            g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
            st = self.state[p]
            out.append((p, g, [st.get(nm) for nm in state_names], st))
        return out


class SGD(_FusedBase):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                      nesterov=nesterov))
        self._tables = {}

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        for gi, group in enumerate(self.param_groups):
            items = self._collect(group, ("momentum_buffer",))
            if not items:
                continue
            mom = float(group["momentum"])
            # torch initialises momentum_buffer = grad on a parameter's first step; parameters that join later would
            # need their own first-step flag, so a mixed group is split into two launches
            for first in (True, False):
                sub = [(p, g, s, st) for p, g, s, st in items if (s[0] is None) == first] if mom != 0 else \
                      (items if first else [])
                if not sub:
                    continue
                entries = []
                for p, g, s, st in sub:
                    if mom != 0 and s[0] is None:
                        st["momentum_buffer"] = torch.empty_like(p, memory_format=torch.contiguous_format)
                    buf = st["momentum_buffer"].data_ptr() if mom != 0 else 0
                    entries.append((p.data_ptr(), g.data_ptr(), buf, 0, p.numel()))
                tab = self._tables.setdefault((gi, first), _Tables())
                tab.update(entries, sub[0][0].device)
                _lib.check(lib.ubr_sgd_step(tab.slots.data_ptr(), tab.chunks.data_ptr(), tab.nchunks, float(group["lr"]),
                                            mom, float(group["dampening"]), float(group["weight_decay"]),
                                            int(bool(group["nesterov"])), int(first and mom != 0), _stream()), "sgd_step")
                _mark_modified([p for p, _, _, _ in sub])
                if mom != 0:
                    _mark_modified([st["momentum_buffer"] for _, _, _, st in sub])
        return loss


class Adam(_FusedBase):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._tables = {}

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        for gi, group in enumerate(self.param_groups):
            items = self._collect(group, ("exp_avg", "exp_avg_sq"))
            if not items:
                continue
            by_step = {}
            for p, g, s, st in items:
                if s[0] is None:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] += 1
                by_step.setdefault(int(st["step"].item()), []).append((p, g, st))
            for step, sub in by_step.items():          # one launch (all parameters share the step count) in practice
                entries = [(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel())
                           for p, g, st in sub]
                tab = self._tables.setdefault((gi, len(by_step) > 1 and step), _Tables())
                tab.update(entries, sub[0][0].device)
                b1, b2 = group["betas"]
                _lib.check(lib.ubr_adam_step(tab.slots.data_ptr(), tab.chunks.data_ptr(), tab.nchunks, float(group["lr"]),
                                             float(b1), float(b2), float(group["eps"]), float(group["weight_decay"]),
                                             step, _stream()), "adam_step")
                _mark_modified([p for p, _, _ in sub])
                _mark_modified([t for _, _, st in sub for t in (st["exp_avg"], st["exp_avg_sq"])])
        return loss
