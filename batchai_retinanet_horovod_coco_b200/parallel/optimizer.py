"""Keras-semantics Adam and the Horovod-style ``DistributedOptimizer``.

Reference: ``hvd.DistributedOptimizer(keras.optimizers.adam(lr=1e-5, clipnorm=0.001))``
(``train.py:103-104``).  Order of operations reproduced here: Keras'
``get_gradients`` clips each rank's gradients by their LOCAL global norm, Horovod then averages
them across ranks, then Adam applies the update

    lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t);   p -= lr_t * m / (sqrt(v) + eps),   eps = 1e-7.

This module is the plain-PyTorch (CPU-capable) implementation and the oracle for the fused CUDA
optimizer kernels used by the native engine (``csrc/optim.cu``).
"""

from __future__ import annotations

import math

import torch


class KerasAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, beta_1: float = 0.9, beta_2: float = 0.999, epsilon: float = 1e-7,
                 decay: float = 0.0, clipnorm: float | None = None, clipvalue: float | None = None):
        defaults = dict(lr=lr, beta_1=beta_1, beta_2=beta_2, epsilon=epsilon, decay=decay)
        super().__init__(params, defaults)
        self.clipnorm = clipnorm
        self.clipvalue = clipvalue
        self.iterations = 0

    # -- Keras get_gradients(): clipping happens BEFORE any allreduce -------------------------
    @torch.no_grad()
    def clip_local_gradients(self) -> float:
        grads = [p.grad for g in self.param_groups for p in g["params"] if p.grad is not None]
        if not grads:
            return 0.0
        norm = math.sqrt(sum(float((g.double() ** 2).sum()) for g in grads))
        if self.clipnorm and self.clipnorm > 0 and norm >= self.clipnorm:
            s = self.clipnorm / norm
            for g in grads:
                g.mul_(s)
        if self.clipvalue and self.clipvalue > 0:
            for g in grads:
                g.clamp_(-self.clipvalue, self.clipvalue)
        return norm

    @torch.no_grad()
    def apply_gradients(self):
        self.iterations += 1
        t = self.iterations
        for group in self.param_groups:
            lr = group["lr"]
            if group["decay"] > 0:
                lr = lr * (1.0 / (1.0 + group["decay"] * (t - 1)))
            b1, b2, eps = group["beta_1"], group["beta_2"], group["epsilon"]
            lr_t = lr * math.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["m"] = torch.zeros_like(p)
                    st["v"] = torch.zeros_like(p)
                g = p.grad
                st["m"].mul_(b1).add_(g, alpha=1 - b1)
                st["v"].mul_(b2).addcmul_(g, g, value=1 - b2)
                p.addcdiv_(st["m"], st["v"].sqrt().add_(eps), value=-lr_t)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.clip_local_gradients()
        self.apply_gradients()
        return loss

    def get_config(self) -> dict:
        g = self.param_groups[0]
        cfg = {"lr": g["lr"], "beta_1": g["beta_1"], "beta_2": g["beta_2"], "epsilon": g["epsilon"],
               "decay": g["decay"], "amsgrad": False}
        if self.clipnorm:
            cfg["clipnorm"] = self.clipnorm
        if self.clipvalue:
            cfg["clipvalue"] = self.clipvalue
        return cfg

    def state_dict(self):
        sd = super().state_dict()
        sd["iterations"] = self.iterations
        sd["clipnorm"], sd["clipvalue"] = self.clipnorm, self.clipvalue
        return sd

    def load_state_dict(self, sd):
        sd = dict(sd)
        self.iterations = sd.pop("iterations", 0)
        self.clipnorm = sd.pop("clipnorm", self.clipnorm)
        self.clipvalue = sd.pop("clipvalue", self.clipvalue)
        super().load_state_dict(sd)


class DistributedOptimizer:
    """Wraps an optimizer so that gradients are averaged over all ranks before the update.

    * optimizers with ``clip_local_gradients`` (KerasAdam): clip locally -> fused allreduce of every
      gradient (enqueued together so the runtime's tensor fusion packs them) -> apply.  This is the
      reference's exact order, and it is why nothing can overlap with backward there.
    * any other ``torch.optim.Optimizer``: post-accumulate-grad hooks enqueue each gradient's
      allreduce as soon as autograd produces it (overlapping backward), ``step()`` waits for them.
    """

    def __init__(self, optimizer, named_parameters=None, average: bool = True, compression=None):
        from . import allreduce_async_, is_initialized, size
        self._optimizer = optimizer
        self._average = average
        if compression is None:
            self._allreduce_async_ = allreduce_async_
        else:       # hvd.Compression.fp16: CUDA gradients cross NVLink in 16 bits
            self._allreduce_async_ = lambda t, avg, name=None: allreduce_async_(t, avg, name=name, compression=compression)
        self._size = size if is_initialized() else (lambda: 1)
        params = [p for g in optimizer.param_groups for p in g["params"]]
        if named_parameters is not None:
            names = {id(p): n for n, p in named_parameters}
        else:
            names = {}
        self._names = {id(p): names.get(id(p), f"param.{i}") for i, p in enumerate(params)}
        self._params = params
        self._handles = {}
        self._hooks = []
        self._keras_order = hasattr(optimizer, "clip_local_gradients")
        if not self._keras_order and self._size() > 1:
            for p in params:
                if p.requires_grad:
                    self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook()))

    def _make_hook(self):
        def hook(p):
            if id(p) in self._handles:
                raise AssertionError("gradient computed twice before step(); call synchronize() between backwards")
            self._handles[id(p)] = self._allreduce_async_(p.grad, self._average, name=f"grad.{self._names[id(p)]}")
        return hook

    def synchronize(self):
        for h in self._handles.values():
            h.wait()
        self._handles.clear()

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        if self._keras_order:
            self._optimizer.clip_local_gradients()
            if self._size() > 1:
                hs = [self._allreduce_async_(p.grad, self._average, name=f"grad.{self._names[id(p)]}")
                      for p in self._params if p.grad is not None]
                for h in hs:
                    h.wait()
            self._optimizer.apply_gradients()
        else:
            if self._size() > 1:
                # parameters whose hook did not fire (unused in this step) still need a matching call
                for p in self._params:
                    if p.requires_grad and p.grad is not None and id(p) not in self._handles:
                        self._handles[id(p)] = self._allreduce_async_(p.grad, self._average,
                                                                      name=f"grad.{self._names[id(p)]}")
                self.synchronize()
            self._optimizer.step()
        return loss

    def zero_grad(self, set_to_none: bool = False):
        # keep gradient storage stable (in-place allreduce targets) unless asked otherwise
        self._optimizer.zero_grad(set_to_none=set_to_none)

    def __getattr__(self, name):
        return getattr(self._optimizer, name)
