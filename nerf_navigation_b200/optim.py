"""FusedAdam — torch.optim.Adam-compatible optimiser for the training configuration (BASELINE config 5), one fused
kernel per parameter tensor (csrc/adam.cu): gradient scaling (the all-reduce mean's 1/world), Adam moments + update with
torch's single-tensor arithmetic, and optionally the EMA shadow (decay 0.95 with torch_ema's warm-up).

torch.amp.GradScaler (the reference's fp16 recipe: scaler.scale(loss).backward(); scaler.step(optimizer); scaler.update(),
nerf/utils.py:784-791) is supported the way torch's own fused optimisers support it: `_step_supports_amp_scaling` makes
GradScaler.step hand over its device-side `grad_scale` and `found_inf` tensors instead of unscaling on the host; the
kernel divides by the scale and skips the step (moments, step count and EMA untouched) when an inf / nan was found.

Reference: main_nerf.py:126 `torch.optim.Adam(model.get_params(lr), betas=(0.9, 0.99), eps=1e-15)`, stepped by
nerf/utils.py:784-791 and followed by `ema.update()` (:814-815).  State keys (`step`, `exp_avg`, `exp_avg_sq`) are
torch.optim.Adam's, so an optimiser state_dict from a reference checkpoint loads unchanged (nerf/utils.py:1040-1046)."""
import torch

from . import _lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, ema_decay=None):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self.ema_decay = ema_decay
        self.grad_mult = 1.0   # set by the training step: 1/world (the gradient SUM was all-reduced)
        self._step_supports_amp_scaling = True   # GradScaler.step sets .grad_scale / .found_inf (device tensors) around step()
        self.ema_updates = 0

    def ema_shadow(self, p):
        return self.state[p].get("ema")

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _lib.lib()
        # torch.amp.GradScaler.step sets these two attributes around the call (and deletes them afterwards)
        amp_scale = getattr(self, "grad_scale", None)
        found_inf = getattr(self, "found_inf", None)
        skipped = None
        if found_inf is not None:
            found_inf = found_inf.float().reshape(1) if found_inf.dtype != torch.float32 else found_inf.reshape(1)
            skipped = bool(found_inf.item() != 0)   # host copy of the flag: the step counters must not advance on a skip
        if amp_scale is not None:
            amp_scale = amp_scale.float().reshape(1)
        if skipped:
            return loss
        ema_t = 0.0
        if self.ema_decay is not None:
            # torch_ema.ExponentialMovingAverage(use_num_updates=True), the reference's default (nerf/utils.py:380):
            # decay_t = min(decay, (1 + n) / (10 + n)) with n counted from 1
            self.ema_updates += 1
            ema_t = min(float(self.ema_decay), (1.0 + self.ema_updates) / (10.0 + self.ema_updates))
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous fp32 CUDA tensors (no CPU path)")
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    if self.ema_decay is not None:
                        st["ema"] = p.detach().clone()
                st["step"] = int(st["step"]) + 1
                g = p.grad.contiguous()
                with torch.cuda.device(p.device):
                    _lib.check(L.ngp_adam_step(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(),
                                               _lib.ptr(st.get("ema")), p.numel(), float(group["lr"]), float(b1), float(b2),
                                               float(group["eps"]), st["step"], float(self.grad_mult), float(ema_t),
                                               _lib.ptr(amp_scale), _lib.ptr(found_inf), _lib.stream()), "adam_step")
        return loss
