"""Multi-tensor Adam on the B200 kernel `csb_adam_step`.

Drop-in for `torch.optim.Adam(param_groups, lr=0.0, eps=1e-15)` as MultiGaussianMesh.training_setup builds it
(scene_reconstruction/gaussian_mesh.py:121-135) and train_step drives it (train_utils.py:312-318): same constructor,
`param_groups` (so `update_learning_rate` and the densification code that edits groups keep working), same per-parameter
state keys (`step`, `exp_avg`, `exp_avg_sq`) and therefore the same `state_dict()`.  One kernel launch per 48 parameter
tensors instead of ~8 elementwise launches per tensor.  amsgrad / weight decay / maximize are not used by the
reference and raise.  No CPU fallback: parameters must be CUDA fp32 tensors."""
import ctypes

import torch

import csb_cabi as _abi


def _bind(lib):
    if getattr(lib, "_adam_bound", False):
        return
    c_int, c_void_p, c_double = ctypes.c_int, ctypes.c_void_p, ctypes.c_double
    lib.csb_adam_step.restype = c_int
    lib.csb_adam_step.argtypes = [c_int] + [c_void_p] * 6 + [c_double, c_double, c_double, c_int, c_int, c_void_p]
    lib.csb_adam_step_capturable.restype = c_int
    lib.csb_adam_step_capturable.argtypes = [c_int] + [c_void_p] * 6 + [c_double, c_double, c_double, c_void_p, c_void_p,
                                                                     c_int, c_void_p]
    lib._adam_bound = True


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, *,
                 maximize=False, zero_grads=False):
        if weight_decay != 0 or amsgrad or maximize:
            raise NotImplementedError("csb_optim.Adam: weight_decay / amsgrad / maximize are not on the hot path")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False)
        super().__init__(params, defaults)
        self.zero_grads = bool(zero_grads)
        self._capturable = None  # {(device, beta1, beta2, eps): (step int64[1], bias float64[2])} while in capturable mode

    # ---- CUDA-graph capture ------------------------------------------------------------------------------------
    def make_capturable(self):
        """From here on the step count lives on the device (csb_adam_step_capturable): `step()` has no host-side state
        that changes from call to call, so it can be captured in a CUDA graph and replayed.  All parameters of a
        (betas, eps) bucket must share their step count (they do in the refinement loop: the densification code moves the
        state rows together).  `leave_capturable()` writes the device counts back into `state[p]['step']`."""
        if self._capturable is not None:
            return
        cap = {}
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                key = (p.device, float(b1), float(b2), float(group["eps"]))
                step = int(st["step"])
                if key not in cap:
                    cap[key] = (torch.full((1,), step, dtype=torch.int64, device=p.device),
                                torch.zeros(2, dtype=torch.float64, device=p.device), step)
                elif cap[key][2] != step:
                    raise ValueError("csb_optim.Adam.make_capturable: parameters of one bucket have different step counts")
        self._capturable = {k: (a, b) for k, (a, b, _) in cap.items()}

    def leave_capturable(self):
        if self._capturable is None:
            return
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                key = (p.device, float(b1), float(b2), float(group["eps"]))
                if key in self._capturable and p in self.state:
                    self.state[p]["step"] = int(self._capturable[key][0].item())
        self._capturable = None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _abi.lib()
        _bind(lib)
        # tensors that share (betas, eps, step) go out in one call
        buckets = {}
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise ValueError("csb_optim.Adam needs CUDA fp32 parameters and gradients; there is no CPU path")
                if p.grad.is_sparse or not p.is_contiguous() or not p.grad.is_contiguous():
                    raise ValueError("csb_optim.Adam needs dense contiguous parameters and gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                if self._capturable is None:
                    st["step"] = int(st["step"]) + 1
                    key = (p.device, float(b1), float(b2), float(group["eps"]), st["step"])
                else:
                    key = (p.device, float(b1), float(b2), float(group["eps"]), None)
                buckets.setdefault(key, []).append((p, st, float(group["lr"])))
        for (dev, b1, b2, eps, step), items in buckets.items():
            n = len(items)
            P = (ctypes.c_void_p * n)(*[p.data_ptr() for p, _, _ in items])
            G = (ctypes.c_void_p * n)(*[p.grad.data_ptr() for p, _, _ in items])
            M = (ctypes.c_void_p * n)(*[s["exp_avg"].data_ptr() for _, s, _ in items])
            V = (ctypes.c_void_p * n)(*[s["exp_avg_sq"].data_ptr() for _, s, _ in items])
            N = (ctypes.c_int64 * n)(*[p.numel() for p, _, _ in items])
            LR = (ctypes.c_double * n)(*[lr for _, _, lr in items])
            with torch.cuda.device(dev):
                if step is None:
                    step_dev, bias_dev = self._capturable[(dev, b1, b2, eps)]
                    rc = lib.csb_adam_step_capturable(n, P, G, M, V, N, LR, b1, b2, eps, step_dev.data_ptr(),
                                                      bias_dev.data_ptr(), int(self.zero_grads), _abi.current_stream_ptr())
                else:
                    rc = lib.csb_adam_step(n, P, G, M, V, N, LR, b1, b2, eps, step, int(self.zero_grads),
                                           _abi.current_stream_ptr())
            _abi.check(rc, "csb_adam_step")
            # the kernel wrote through raw pointers: tell autograd (saved-tensor checks) and every cache keyed on
            # tensor versions (EncodeProcessDecode's packed bf16 weights) that these parameters changed
            torch.autograd.graph.increment_version([p for p, _, _ in items])
            if self.zero_grads:
                torch.autograd.graph.increment_version([p.grad for p, _, _ in items])
        return loss
