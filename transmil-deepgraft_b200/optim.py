"""Optimizer side of the training step: the reference's optimizer factory for the two optimizers its
TransMIL configs name, and the gradient exchange of bag-data-parallel training.

Reference: `create_optimizer(args, model)` (code/MyOptimizer/optim_factory.py:40-123) -- `opt: radam` builds
`torch.optim.RAdam` over the `add_weight_decay` parameter groups (:25-37: biases and 1-D tensors undecayed), `lookahead_radam`
wraps it in `Lookahead(alpha=0.5, k=6)` (code/MyOptimizer/lookahead.py:12-52); DeepGraft/TransMIL_feat_norm_rest.yaml:49-55
sets lr 1e-4, weight_decay 0.01.  Here both updates are one kernel launch over all 29 parameter tensors
(tm_radam_step / tm_lookahead_sync, csrc/backward.cu).  CUDA only, no fallback.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import capi


def add_weight_decay(model, weight_decay=1e-5, skip_list=()):
    """code/MyOptimizer/optim_factory.py:25-37."""
    decay, no_decay = [], []
    for name, param in model.named_parameters():
        if not param.requires_grad:
            continue
        if len(param.shape) == 1 or name.endswith(".bias") or name in skip_list:
            no_decay.append(param)
        else:
            decay.append(param)
    return [{"params": no_decay, "weight_decay": 0.0}, {"params": decay, "weight_decay": weight_decay}]


class FusedRAdam(torch.optim.Optimizer):
    """torch.optim.RAdam semantics (L2 weight decay, not decoupled), one launch per step for every tensor."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, lib=None):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._lib = lib
        self._step = 0

    def _tensors(self, p_list, slot_m, slot_v, wd):
        arr = (capi.OptTensor * len(p_list))()
        for i, p in enumerate(p_list):
            st = self.state[p]
            for key in (slot_m, slot_v):
                if key is not None and key not in st:
                    st[key] = torch.zeros_like(p.data, memory_format=torch.contiguous_format)
            if p.dtype != torch.float32 or not p.data.is_contiguous() or (p.grad is not None and not p.grad.is_contiguous()):
                raise capi.TransmilError("FusedRAdam needs contiguous float32 parameters and gradients")
            arr[i].p, arr[i].n, arr[i].weight_decay = p.data.data_ptr(), p.numel(), wd
            arr[i].g = p.grad.data_ptr() if p.grad is not None else None
            arr[i].m = st[slot_m].data_ptr() if slot_m is not None else None
            arr[i].v = st[slot_v].data_ptr() if slot_v is not None else None
        return arr

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        L = self._lib if self._lib is not None else capi.lib()
        self._step += 1
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            if self._lib is None and not ps[0].is_cuda:
                raise capi.TransmilError("FusedRAdam runs on CUDA parameters only (there is no CPU path)")
            stream = torch.cuda.current_stream(ps[0].device).cuda_stream if ps[0].is_cuda else None
            for i0 in range(0, len(ps), capi.TM_OPT_MAX_TENSORS):
                chunk = ps[i0:i0 + capi.TM_OPT_MAX_TENSORS]
                arr = self._tensors(chunk, "exp_avg", "exp_avg_sq", float(group["weight_decay"] or 0.0))
                status = L.tm_radam_step(arr, len(chunk), self._step, group["lr"], group["betas"][0], group["betas"][1], group["eps"], stream)
                if status != 0:
                    raise capi.TransmilError("tm_radam_step: " + L.tm_last_error().decode())
        return loss


class FusedLookahead:
    """Lookahead(alpha, k) around FusedRAdam (code/MyOptimizer/lookahead.py): every k-th step the slow weights move alpha of
    the way to the fast weights and the fast weights are reset to them.  The slow buffer is created at the first sync, as a copy
    of the fast weights of that moment (lookahead.py:33-35)."""

    def __init__(self, base, alpha=0.5, k=6):
        self.base, self.alpha, self.k, self.steps = base, alpha, k, 0
        self.param_groups = base.param_groups

    def zero_grad(self, set_to_none=True):
        self.base.zero_grad(set_to_none=set_to_none)

    def step(self, closure=None):
        loss = self.base.step(closure)
        self.steps += 1
        if self.steps % self.k == 0:
            self.sync()
        return loss

    @torch.no_grad()
    def sync(self):
        b = self.base
        L = b._lib if b._lib is not None else capi.lib()
        ps = [p for g in b.param_groups for p in g["params"] if p.grad is not None]
        for p in ps:
            if "slow_buffer" not in b.state[p]:
                b.state[p]["slow_buffer"] = p.data.clone(memory_format=torch.contiguous_format)
        stream = torch.cuda.current_stream(ps[0].device).cuda_stream if ps and ps[0].is_cuda else None
        for i0 in range(0, len(ps), capi.TM_OPT_MAX_TENSORS):
            chunk = ps[i0:i0 + capi.TM_OPT_MAX_TENSORS]
            arr = b._tensors(chunk, "slow_buffer", None, 0.0)
            status = L.tm_lookahead_sync(arr, len(chunk), self.alpha, stream)
            if status != 0:
                raise capi.TransmilError("tm_lookahead_sync: " + L.tm_last_error().decode())


def create_optimizer(args, model, filter_bias_and_bn=True, lib=None):
    """code/MyOptimizer/optim_factory.py:40-123 for the optimizers the TransMIL configs select ('radam', 'lookahead_radam').
    `args` carries .opt, .lr, .weight_decay and optionally .opt_eps / .opt_betas like the reference's cfg.Optimizer."""
    opt_lower = args.opt.lower()
    weight_decay = getattr(args, "weight_decay", 0.0) or 0.0
    if weight_decay and filter_bias_and_bn:
        skip = model.no_weight_decay() if hasattr(model, "no_weight_decay") else {}
        parameters = add_weight_decay(model, weight_decay, skip)
        weight_decay = 0.0
    else:
        parameters = list(model.parameters())
    opt_args = dict(lr=args.lr, weight_decay=weight_decay)
    if getattr(args, "opt_eps", None) is not None:
        opt_args["eps"] = args.opt_eps
    if getattr(args, "opt_betas", None) is not None:
        opt_args["betas"] = tuple(args.opt_betas)
    split = opt_lower.split("_")
    if split[-1] != "radam":
        raise ValueError(f"transmil_b200.optim builds 'radam' and 'lookahead_radam' (the TransMIL configs' optimizers), not {args.opt!r}")
    optimizer = FusedRAdam(parameters, lib=lib, **opt_args)
    if len(split) > 1 and split[0] == "lookahead":
        optimizer = FusedLookahead(optimizer)
    return optimizer


def allreduce_gradients(params, group=None, average=True):
    """Bag-data-parallel training: every rank differentiates its own bags; the 29 gradient tensors
    (4.77 M floats, 19 MB) are flattened into ONE bucket, summed with a single allreduce (NCCL over NVLink on the GPU box, gloo
    in the CPU tests) and scattered back -- what Lightning DDP does for the reference (train.py:184), without per-bucket
    launches.  With accumulate_grad_batches (train.py:199) call it on the last micro-batch only."""
    ps = [p for p in params if p.grad is not None]
    if not ps or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return 0
    flat = torch.cat([p.grad.reshape(-1) for p in ps])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    if average:
        flat /= dist.get_world_size(group)
    off = 0
    for p in ps:
        n = p.numel()
        p.grad.copy_(flat[off:off + n].view_as(p.grad))
        off += n
    return flat.numel()
