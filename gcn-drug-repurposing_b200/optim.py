"""Adam for the drop-in path: same constructor arguments, update rule and state layout as
torch.optim.Adam (train.py:139-141: lr=args.lr — 1e-4 by default, 3e-4 in train.sh — betas (0.9, 0.999), eps 1e-8,
weight_decay 0, amsgrad False),
but all parameter tensors are updated by ONE libgdr kernel (gdr_adam_step_multi_f32) instead of torch's
foreach pipeline (7 launches per step at 4 tensors).  state_dict() is interchangeable with torch.optim.Adam's:
per-parameter 'step' (0-d fp32 CPU tensor), 'exp_avg', 'exp_avg_sq'."""
import ctypes

import torch

from . import _ffi


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise RuntimeError("gdr Adam: weight_decay and amsgrad are not on the reference path (train.py:139-141)")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False,
                                      foreach=None, capturable=False, differentiable=False, fused=None))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _ffi.lib()
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            for p in ps:
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("gdr Adam: parameters must be contiguous fp32 CUDA tensors")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            steps = {int(self.state[p]["step"].item()) for p in ps}
            b1, b2 = group["betas"]
            for s0 in sorted(steps):                       # normally one value: all tensors stepped together
                sel = [p for p in ps if int(self.state[p]["step"].item()) == s0]
                for k in range(0, len(sel), 16):
                    chunk = sel[k:k + 16]
                    n = len(chunk)
                    arr = ctypes.c_void_p * n
                    grads = [p.grad.contiguous() for p in chunk]
                    rc = L.gdr_adam_step_multi_f32(
                        n, arr(*[p.data_ptr() for p in chunk]), arr(*[g.data_ptr() for g in grads]),
                        arr(*[self.state[p]["exp_avg"].data_ptr() for p in chunk]),
                        arr(*[self.state[p]["exp_avg_sq"].data_ptr() for p in chunk]),
                        (ctypes.c_int64 * n)(*[p.numel() for p in chunk]), s0 + 1, float(group["lr"]), float(b1),
                        float(b2), float(group["eps"]), _ffi.stream())
                    _ffi.check(rc, "adam_step_multi")
                for p in sel:
                    self.state[p]["step"] += 1
        return loss
