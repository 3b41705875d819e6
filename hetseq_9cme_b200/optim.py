"""Optimizers.

Public surface mirrors hetseq/optim.py (``_Optimizer`` wrapper API at :6-80, ``_Adam`` / ``_Adadelta``
wrappers at :83-131, raw ``Adam`` / ``Adadelta`` ``torch.optim.Optimizer`` classes at :133-304) and the
update rules are the reference's, not torch's:

* Adam: ``m, v`` moments; ``denom = sqrt(v) + eps`` (eps outside the sqrt, v not bias-corrected separately);
  ``step_size = lr * sqrt(1 - b2^t) / (1 - b1^t)``; decoupled weight decay ``p *= (1 - wd*lr)`` applied before
  the moment update is subtracted.
* Adadelta: stock rule with L2 weight decay folded into the gradient.

Two execution paths share the same state layout (so checkpoints are interchangeable):

* **arena path** (CUDA + ``ParamArena``): moments live in flat fp32 buffers parallel to the parameter arena;
  ``fused_step`` runs scale -> global-norm clip -> update -> zero_grad as one sm_100a kernel driven by device
  scalars (no host sync).  ``state[p]['exp_avg']`` etc. are views into the flat buffers.
* **tensor path** (CPU, or parameters outside an arena): per-tensor ``torch._foreach`` arithmetic.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from .options import _literal


def _foreach_ok(tensors) -> bool:
    return len(tensors) > 0 and hasattr(torch, "_foreach_mul_")


# -------------------------------------------------------------------------------------------------
# raw optimizers
# -------------------------------------------------------------------------------------------------
class _ArenaMixin:
    """Shared plumbing for optimizers whose state can live in flat buffers next to a ParamArena."""

    _state_names = ()

    def attach_arena(self, arena):
        self.arena = arena
        self._flat_state = {name: torch.zeros(arena.numel, dtype=torch.float32, device=arena.device)
                            for name in self._state_names}
        self._global_step = 0
        for p in arena.params:
            o, n = arena.offsets[id(p)], p.numel()
            st = self.state[p]
            st["step"] = 0
            for name in self._state_names:
                st[name] = self._flat_state[name][o:o + n].view(p.shape)

    @property
    def has_arena(self) -> bool:
        return getattr(self, "arena", None) is not None

    def _load_state_exact(self, state_dict):
        """``load_state_dict`` for the arena path.  torch's implementation casts every state tensor to the dtype of its
        parameter — in bf16 mode that is the bf16 working copy, which would round the fp32 moments to 8 bits on every
        resume.  Here the hyper-parameters go through torch (empty state) and the moments are copied into the flat
        fp32 buffers as they are."""
        groups = state_dict["param_groups"]
        saved_ids = [pid for g in groups for pid in g["params"]]
        ours = [p for g in self.param_groups for p in g["params"]]
        if len(saved_ids) != len(ours):
            raise ValueError("loaded state dict has a different number of parameters ({} vs {})".format(len(saved_ids), len(ours)))
        torch.optim.Optimizer.load_state_dict(self, {"state": {}, "param_groups": groups})
        arena = self.arena
        steps = []
        for pid, p in zip(saved_ids, ours):
            saved = state_dict["state"].get(pid, {})
            o, n = arena.offsets[id(p)], p.numel()
            st = self.state[p]
            for name in self._state_names:
                flat = self._flat_state[name][o:o + n].view(p.shape)
                if name in saved:
                    flat.copy_(saved[name].to(device=flat.device, dtype=torch.float32).reshape(p.shape))
                else:
                    flat.zero_()
                st[name] = flat
            st["step"] = int(saved.get("step", 0))
            steps.append(st["step"])
        self._global_step = max(steps) if steps else 0

    def _load_flat_state(self):
        """After ``load_state_dict`` the per-parameter tensors are fresh copies: move them back into the flat
        buffers and re-establish the views."""
        arena = self.arena
        steps = []
        for p in arena.params:
            st = self.state.get(p, {})
            o, n = arena.offsets[id(p)], p.numel()
            for name in self._state_names:
                flat = self._flat_state[name][o:o + n].view(p.shape)
                if name in st and st[name].data_ptr() != flat.data_ptr():
                    flat.copy_(st[name].to(device=flat.device, dtype=torch.float32))
                st[name] = flat
            steps.append(int(st.get("step", 0)))
            self.state[p] = st
        self._global_step = max(steps) if steps else 0


class Adam(_ArenaMixin, torch.optim.Optimizer):
    """Adam with the reference's update rule (hetseq/optim.py:162-231)."""

    _state_names = ("exp_avg", "exp_avg_sq")

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad)
        super().__init__(params, defaults)
        self.arena = None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self.has_arena and self.arena.device.type == "cuda":
            self.fused_step()
            return loss
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            lr, eps, wd, amsgrad = group["lr"], group["eps"], group["weight_decay"], group["amsgrad"]
            todo = [p for p in group["params"] if p.grad is not None]
            if not todo:
                continue
            for p in todo:
                st = self.state[p]
                if len(st) == 0 or "exp_avg" not in st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                    if amsgrad:
                        st["max_exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                st["step"] += 1
            # all tensors of a group share the step count
            t = self.state[todo[0]]["step"]
            step_size = lr * math.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
            grads = [p.grad.float() if p.grad.dtype != torch.float32 else p.grad for p in todo]
            ms = [self.state[p]["exp_avg"] for p in todo]
            vs = [self.state[p]["exp_avg_sq"] for p in todo]
            master = [p.float() if p.dtype != torch.float32 else p for p in todo]
            same_step = all(self.state[p]["step"] == t for p in todo)
            if _foreach_ok(todo) and same_step:
                torch._foreach_mul_(ms, beta1)
                torch._foreach_add_(ms, grads, alpha=1.0 - beta1)
                torch._foreach_mul_(vs, beta2)
                torch._foreach_addcmul_(vs, grads, grads, value=1.0 - beta2)
                if amsgrad:
                    mx = [self.state[p]["max_exp_avg_sq"] for p in todo]
                    torch._foreach_maximum_(mx, vs)
                    denom = torch._foreach_sqrt(mx)
                else:
                    denom = torch._foreach_sqrt(vs)
                torch._foreach_add_(denom, eps)
                if wd != 0:
                    torch._foreach_mul_(master, 1.0 - wd * lr)
                torch._foreach_addcdiv_(master, ms, denom, value=-step_size)
            else:
                for p, g, m, v, w in zip(todo, grads, ms, vs, master):
                    ti = self.state[p]["step"]
                    ss = lr * math.sqrt(1.0 - beta2 ** ti) / (1.0 - beta1 ** ti)
                    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
                    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
                    if amsgrad:
                        mx = self.state[p]["max_exp_avg_sq"]
                        torch.maximum(mx, v, out=mx)
                        d = mx.sqrt().add_(eps)
                    else:
                        d = v.sqrt().add_(eps)
                    if wd != 0:
                        w.mul_(1.0 - wd * lr)
                    w.addcdiv_(m, d, value=-ss)
            for p, w in zip(todo, master):
                if w is not p:
                    p.copy_(w)
        return loss

    def hyper_values(self, step: int):
        """[lr, bias-corrected step size, decoupled decay factor, 0] for update number ``step`` (1-based)."""
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        return [g["lr"], g["lr"] * math.sqrt(1.0 - b2 ** step) / (1.0 - b1 ** step), 1.0 - g["weight_decay"] * g["lr"], 0.0]

    def fused_step(self, *, grad_scale_mult: float = 1.0, max_norm: float = 0.0, norm_sq=None, sample_size=None,
                   stats_out=None, zero_grad: bool = False, hyper_dev=None, advance: bool = True):
        """Arena update in one kernel; scale / clip coefficients are derived on the device from
        ``norm_sq`` (sum g^2 of the unscaled gradient) and ``sample_size`` (both device tensors, optional).
        ``hyper_dev`` (device float[4], see ``hyper_values``) makes the launch replayable from a CUDA graph."""
        from .ops import kernels
        g = self.param_groups[0]
        if advance:
            self._global_step += 1
        a = self.arena
        kernels.adam_flat(a.master, a.grad, self._flat_state["exp_avg"], self._flat_state["exp_avg_sq"], a.numel,
                          lr=g["lr"], beta1=g["betas"][0], beta2=g["betas"][1], eps=g["eps"], weight_decay=g["weight_decay"],
                          step=max(self._global_step, 1), grad_scale_mult=grad_scale_mult, max_norm=max_norm, norm_sq=norm_sq,
                          sample_size=sample_size, stats_out=stats_out, zero_grad=zero_grad, p_bf16=a.work, hyper_dev=hyper_dev)
        if advance:
            self.mark_step()

    def sharded_step(self, reducer, *, grad_scale_mult: float = 1.0, max_norm: float = 0.0, norm_sq=None, sample_size=None,
                     stats_out=None, hyper_dev=None, advance: bool = True, gate=None, pipelined: bool = False):
        """The same update restricted to the arena slices this rank owns, with the new parameters written into every
        rank by the kernel itself (``parallel/p2p.py``); moments of slices owned elsewhere are not maintained here —
        ``reducer.gather_sharded`` makes them whole before a checkpoint."""
        from .ops import kernels
        g = self.param_groups[0]
        if advance:
            self._global_step += 1
        args = kernels.make_adam_args(lr=g["lr"], beta1=g["betas"][0], beta2=g["betas"][1], eps=g["eps"],
                                      weight_decay=g["weight_decay"], step=max(self._global_step, 1),
                                      grad_scale_mult=grad_scale_mult, max_norm=max_norm, norm_sq=norm_sq,
                                      sample_size=sample_size, stats_out=stats_out, zero_grad=True, hyper_dev=hyper_dev)
        launch = reducer.sharded_adam_pipelined if pipelined else reducer.sharded_adam
        launch(args, self._flat_state["exp_avg"], self._flat_state["exp_avg_sq"], gate=gate)
        if advance:
            self.mark_step()

    def flat_state_tensors(self):
        return [self._flat_state[n] for n in self._state_names] if self.has_arena else []

    def mark_step(self):
        for p in self.arena.params:
            self.state[p]["step"] = self._global_step

    def load_state_dict(self, state_dict):
        if self.has_arena:
            return self._load_state_exact(state_dict)
        super().load_state_dict(state_dict)


class Adadelta(_ArenaMixin, torch.optim.Optimizer):
    """Adadelta (hetseq/optim.py:263-304)."""

    _state_names = ("square_avg", "acc_delta")

    def __init__(self, params, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0):
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if not 0.0 <= rho <= 1.0:
            raise ValueError("Invalid rho value: {}".format(rho))
        if eps < 0.0:
            raise ValueError("Invalid epsilon value: {}".format(eps))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        super().__init__(params, dict(lr=lr, rho=rho, eps=eps, weight_decay=weight_decay))
        self.arena = None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self.has_arena and self.arena.device.type == "cuda":
            self.fused_step()
            return loss
        for group in self.param_groups:
            rho, eps, lr, wd = group["rho"], group["eps"], group["lr"], group["weight_decay"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad
                st = self.state[p]
                if len(st) == 0 or "square_avg" not in st:
                    st["step"] = 0
                    st["square_avg"] = torch.zeros_like(p)
                    st["acc_delta"] = torch.zeros_like(p)
                st["step"] += 1
                if wd != 0:
                    g = g.add(p, alpha=wd)
                sq, acc = st["square_avg"], st["acc_delta"]
                sq.mul_(rho).addcmul_(g, g, value=1.0 - rho)
                std = sq.add(eps).sqrt_()
                delta = acc.add(eps).sqrt_().div_(std).mul_(g)
                p.add_(delta, alpha=-lr)
                acc.mul_(rho).addcmul_(delta, delta, value=1.0 - rho)
        return loss

    def hyper_values(self, step: int):
        return [self.param_groups[0]["lr"], 0.0, 0.0, 0.0]

    def fused_step(self, *, grad_scale_mult: float = 1.0, max_norm: float = 0.0, norm_sq=None, sample_size=None,
                   stats_out=None, zero_grad: bool = False, hyper_dev=None, advance: bool = True):
        from .ops import kernels
        g = self.param_groups[0]
        if advance:
            self._global_step += 1
        a = self.arena
        kernels.adadelta_flat(a.master, a.grad, self._flat_state["square_avg"], self._flat_state["acc_delta"], a.numel,
                              lr=g["lr"], rho=g["rho"], eps=g["eps"], weight_decay=g["weight_decay"],
                              grad_scale_mult=grad_scale_mult, max_norm=max_norm, norm_sq=norm_sq, sample_size=sample_size,
                              stats_out=stats_out, zero_grad=zero_grad, p_bf16=a.work, hyper_dev=hyper_dev)
        if advance:
            self.mark_step()

    def mark_step(self):
        for p in self.arena.params:
            self.state[p]["step"] = self._global_step

    def load_state_dict(self, state_dict):
        if self.has_arena:
            return self._load_state_exact(state_dict)
        super().load_state_dict(state_dict)


# -------------------------------------------------------------------------------------------------
# wrappers used by the Controller
# -------------------------------------------------------------------------------------------------
class _Optimizer:
    """Uniform optimizer facade (lr access, gradient scaling / clipping, state dict)."""

    def __init__(self, args):
        self.args = args

    @property
    def optimizer(self):
        opt = getattr(self, "_optimizer", None)
        if opt is None:
            raise NotImplementedError
        if not isinstance(opt, torch.optim.Optimizer):
            raise ValueError("_optimizer must be an instance of torch.optim.Optimizer")
        return opt

    @property
    def params(self):
        for group in self.optimizer.param_groups:
            yield from group["params"]

    @property
    def arena(self):
        return getattr(self.optimizer, "arena", None)

    def attach_arena(self, arena):
        self.optimizer.attach_arena(arena)

    def get_lr(self):
        return self.optimizer.param_groups[0]["lr"]

    def set_lr(self, lr):
        for group in self.optimizer.param_groups:
            group["lr"] = lr

    def state_dict(self):
        return self.optimizer.state_dict()

    def load_state_dict(self, state_dict, optimizer_overrides=None):
        """Loads state and hyper-parameters from the checkpoint, then applies ``--optimizer-overrides`` to every group
        (ref: hetseq/optim.py:39-53; scheduled values such as lr are set again by the scheduler's next update)."""
        self.optimizer.load_state_dict(state_dict)
        if optimizer_overrides:
            for group in self.optimizer.param_groups:
                group.update(optimizer_overrides)

    def backward(self, loss):
        loss.backward()

    def _grads(self):
        return [p.grad for p in self.params if p.grad is not None]

    def multiply_grads(self, c):
        """grad *= c for every parameter: one kernel on the arena, one foreach call otherwise."""
        arena = self.arena
        if arena is not None:
            if arena.device.type == "cuda":
                from .ops import kernels
                kernels.scale_(arena.grad, float(c))
            else:
                arena.grad.mul_(c)
            return
        grads = self._grads()
        if grads:
            if _foreach_ok(grads):
                torch._foreach_mul_(grads, c)
            else:
                for g in grads:
                    g.mul_(c)

    def grad_norm(self) -> torch.Tensor:
        """Global L2 norm of all gradients as a 0-d tensor (no host sync)."""
        arena = self.arena
        if arena is not None:
            if arena.device.type == "cuda":
                from .ops import kernels
                ws = kernels.Workspace.get(arena.device)
                kernels.sumsq(arena.grad, ws.norm_sq)
                return ws.norm_sq.sqrt().squeeze(0)
            return arena.grad.norm()
        grads = self._grads()
        if not grads:
            return torch.zeros(())
        return torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g.float()) for g in grads]))

    def clip_grad_norm(self, max_norm):
        """Clip to ``max_norm`` (coefficient ``max_norm / (norm + 1e-6)`` applied only when norm > max_norm, as
        torch.nn.utils.clip_grad_norm_ does); returns the pre-clip norm (ref: hetseq/optim.py:65-70)."""
        norm = self.grad_norm()
        if max_norm > 0:
            coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
            arena = self.arena
            if arena is not None:
                arena.grad.mul_(coef.to(arena.grad.device))
            else:
                grads = self._grads()
                if grads:
                    torch._foreach_mul_(grads, coef.to(grads[0].device)) if _foreach_ok(grads) else [g.mul_(coef) for g in grads]
        return norm

    def step(self, closure=None):
        self.optimizer.step(closure)

    def zero_grad(self):
        """Arena: one memset, gradient views stay alive.  Otherwise ``p.grad = None`` like the reference."""
        arena = self.arena
        if arena is not None:
            arena.zero_grad()
            arena.rebind()
            return
        for p in self.params:
            p.grad = None


class _Adam(_Optimizer):
    def __init__(self, args, params):
        super().__init__(args)
        self._optimizer = Adam(params, **self.optimizer_config)

    @property
    def optimizer_config(self):
        betas = self.args.adam_betas
        if isinstance(betas, str):
            betas = _literal(betas)
        return {"lr": self.args.lr[0], "betas": tuple(betas), "eps": self.args.adam_eps,
                "weight_decay": self.args.weight_decay}


class _Adadelta(_Optimizer):
    def __init__(self, args, params):
        super().__init__(args)
        self._optimizer = Adadelta(params, **self.optimizer_config)

    @property
    def optimizer_config(self):
        return {"lr": self.args.lr[0], "rho": self.args.adadelta_rho, "eps": self.args.adadelta_eps,
                "weight_decay": getattr(self.args, "dadelta_weight_decay", 0.0)}


# name (value of --optimizer) -> wrapper class; extensions add theirs here and their flags to options._OPT_ROWS
OPTIMIZER_REGISTRY = {"adam": _Adam, "adadelta": _Adadelta}


def build_optimizer(args, params):
    name = getattr(args, "optimizer", "adam")
    if name not in OPTIMIZER_REGISTRY:
        raise ValueError("unsupported optimizer: {}".format(name))
    return OPTIMIZER_REGISTRY[name](args, params)
