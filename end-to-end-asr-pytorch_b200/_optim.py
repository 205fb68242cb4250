"""The optimizer of the train step (reference src/solver.py:101-106: `torch.optim.Adadelta(params, lr, eps=1e-8)`,
stepped at solver.py:182) on the flat gradient buffer: ONE launch of `b2asr_adadelta_step` (csrc/optim.cu) instead of
torch's seven multi-tensor passes over ~60 tensors.

`Adadelta` IS a `torch.optim.Adadelta`: same constructor, same `param_groups`, and `state_dict()` /
`load_state_dict()` exchange checkpoints with the stock optimizer (per parameter `step`, `square_avg`, `acc_delta`);
only the storage differs -- the two state tensors of every parameter are views into two flat vectors laid out like
`solver.FlatGradients`.  When the gradients are not the bound flat views (somebody replaced `.grad`), the step runs
through the stock implementation on the same state.
"""
import ctypes

import torch

from . import _lib

CHUNK = 16384          # elements per table entry (one CTA walks an entry)


class Adadelta(torch.optim.Adadelta):
    def __init__(self, params, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0):
        super().__init__(params, lr=lr, rho=rho, eps=eps, weight_decay=weight_decay, foreach=False)
        self._grads = None
        self._tables = None

    # ---- binding to the flat gradient buffer --------------------------------------------------------------
    def bind(self, flat_grads):
        """flat_grads: solver.FlatGradients over (a superset of) this optimizer's parameters."""
        self._grads = flat_grads
        n = flat_grads.flat.numel()
        dev = flat_grads.flat.device
        self._sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self._acc = torch.zeros(n, dtype=torch.float32, device=dev)
        self._offset = {id(p): o for p, o in zip(flat_grads.params, flat_grads.offsets)}
        self._adopt()
        self._tables = None

    def _adopt(self):
        """Make every parameter's state a view into the flat vectors (keeping values that are already there, e.g.
        after load_state_dict)."""
        if self._grads is None:
            return
        for group in self.param_groups:
            for p in group["params"]:
                o = self._offset.get(id(p))
                if o is None:
                    continue
                st = self.state[p]
                for key, flat in (("square_avg", self._sq), ("acc_delta", self._acc)):
                    view = flat[o:o + p.numel()].view_as(p)
                    old = st.get(key)
                    if old is not None and old.data_ptr() != view.data_ptr():
                        view.copy_(old)
                    st[key] = view
                if "step" not in st:
                    st["step"] = torch.zeros((), dtype=torch.float32)

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._adopt()

    def zero_grad(self, set_to_none=True):
        if self._grads is not None:
            self._grads.zero()                      # the views stay attached
            self._grads.rebind()
        else:
            super().zero_grad(set_to_none)

    # ---- the step -------------------------------------------------------------------------------------------
    def _build_tables(self):
        tables = []
        for group in self.param_groups:
            entries, params = [], []
            for p in group["params"]:
                o = self._offset.get(id(p))
                if o is None:
                    continue
                params.append(p)
                for s in range(0, p.numel(), CHUNK):
                    entries.append((p.data_ptr() + 4 * s, o + s, min(CHUNK, p.numel() - s)))
            arr = (_lib.ParamChunk * max(len(entries), 1))()
            for i, (ptr, off, cnt) in enumerate(entries):
                arr[i].param, arr[i].offset, arr[i].count = ptr, off, cnt
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            tables.append((host.to(self._sq.device), len(entries), params, [p.data_ptr() for p in params],
                           [self.state[p]["step"] for p in params]))
        self._tables = tables

    def _flat_ok(self):
        g = self._grads
        if g is None:
            return False
        base = g.flat.data_ptr()
        for group in self.param_groups:
            for p in group["params"]:
                o = self._offset.get(id(p))
                if o is None:
                    if p.grad is not None:
                        return False
                    continue
                if (p.grad is None or p.grad.data_ptr() != base + 4 * o or not p.is_contiguous()
                        or p.dtype != torch.float32 or not p.is_cuda):
                    return False
        return True

    def can_clip_on_device(self):
        """True when step_clipped() may replace clip_grad_norm_ + NaN guard + step() (one parameter group: the clip
        coefficient is applied inside the group's launch)."""
        return self._flat_ok() and len(self.param_groups) == 1

    @torch.no_grad()
    def step(self, closure=None, _clip=None):
        if not self._flat_ok():
            return super().step(closure)
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self._tables is None or any(ptrs != [p.data_ptr() for p in params] for _, _, params, ptrs, _ in self._tables):
            self._adopt()
            self._build_tables()
        for group, (table, n, params, _, steps) in zip(self.param_groups, self._tables):
            if group.get("maximize", False):
                raise _lib.B2asrError("b2asr Adadelta: maximize=True is not supported")
            if _clip is None:
                _lib.call("b2asr_adadelta_step", 1, _lib.lib().b2asr_adadelta_step, table.data_ptr(), n,
                          self._grads.flat.data_ptr(), self._sq.data_ptr(), self._acc.data_ptr(),
                          float(group["lr"]), float(group["rho"]), float(group["eps"]), float(group["weight_decay"]),
                          _lib.stream_ptr())
            else:
                norm, pre_scale, max_norm, skipped = _clip
                _lib.call("b2asr_adadelta_step_clipped", 1, _lib.lib().b2asr_adadelta_step_clipped, table.data_ptr(), n,
                          self._grads.flat.data_ptr(), self._sq.data_ptr(), self._acc.data_ptr(),
                          float(group["lr"]), float(group["rho"]), float(group["eps"]), float(group["weight_decay"]),
                          norm.data_ptr(), float(pre_scale), float(max_norm), _lib.ptr(skipped), _lib.stream_ptr())
            if steps:
                torch._foreach_add_(steps, 1)       # (counts a skipped step too; Adadelta's update does not use it)
        return loss

    def step_clipped(self, grad_norm, pre_scale, max_norm, skipped=None):
        """clip_grad_norm_(max_norm) + NaN guard + step() of reference solver.py:178-182 in the one launch.
        grad_norm: 0-dim/1-element fp32 DEVICE tensor, the 2-norm of the flat gradient as it stands; the gradient
        that is clipped and applied is flat * pre_scale; skipped: int32 device counter of NaN-skipped steps."""
        return self.step(_clip=(grad_norm, pre_scale, max_norm, skipped))
