"""Optimizer side of the training step: the reference's optimizer factory for the two optimizers its
TransMIL configs name, and the gradient exchange of bag-data-parallel training.

Reference: `create_optimizer(args, model)` (code/MyOptimizer/optim_factory.py:40-123) -- `opt: radam` builds
`torch.optim.RAdam` over the `add_weight_decay` parameter groups (:25-37: biases and 1-D tensors undecayed), `lookahead_radam`
wraps it in `Lookahead(alpha=0.5, k=6)` (code/MyOptimizer/lookahead.py:12-52); DeepGraft/TransMIL_feat_norm_rest.yaml:49-55
sets lr 1e-4, weight_decay 0.01.  Here both updates are one kernel launch over all 29 parameter tensors
(tm_radam_step / tm_lookahead_sync, csrc/backward.cu).  CUDA only, no fallback.
"""
import ctypes as C
from collections import OrderedDict, defaultdict

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
    """torch.optim.RAdam semantics (L2 weight decay, not decoupled), one launch per step for every tensor.  The state is kept
    the way torch.optim.RAdam keeps it -- per parameter `step` (a float32 scalar tensor), `exp_avg`, `exp_avg_sq` -- so
    `state_dict()` / `load_state_dict()` round-trip, resume continues the bias correction and rectification where it stopped,
    and a checkpoint written by torch.optim.RAdam (what the reference's factory builds, optim_factory.py:78-80) loads."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, lib=None):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._lib = lib

    @staticmethod
    def _check(p):
        if p.dtype != torch.float32 or not p.data.is_contiguous() or (p.grad is not None and not p.grad.is_contiguous()):
            raise capi.TransmilError("FusedRAdam needs contiguous float32 parameters and gradients")

    def _tensors(self, p_list, state, slot_m, slot_v, wd):
        arr = (capi.OptTensor * len(p_list))()
        for i, p in enumerate(p_list):
            self._check(p)
            st = state[p]
            for key in (slot_m, slot_v):
                if key is not None and key not in st:
                    st[key] = torch.zeros_like(p.data, memory_format=torch.contiguous_format)
            arr[i].p, arr[i].n, arr[i].weight_decay = p.data.data_ptr(), p.numel(), wd
            arr[i].g = p.grad.data_ptr() if p.grad is not None else None
            arr[i].m = st[slot_m].data_ptr() if slot_m is not None else None
            arr[i].v = st[slot_v].data_ptr() if slot_v is not None else None
        return arr

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        L = self._lib if self._lib is not None else capi.lib()
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            if self._lib is None and not ps[0].is_cuda:
                raise capi.TransmilError("FusedRAdam runs on CUDA parameters only (there is no CPU path)")
            stream = torch.cuda.current_stream(ps[0].device).cuda_stream if ps[0].is_cuda else None
            by_step = {}
            for p in ps:                                  # parameters share a launch when they are at the same step (the normal case)
                st = self.state[p]
                if "step" not in st:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                st["step"] += 1
                by_step.setdefault(int(st["step"].item()), []).append(p)
            for step, plist in by_step.items():
                for i0 in range(0, len(plist), capi.TM_OPT_MAX_TENSORS):
                    chunk = plist[i0:i0 + capi.TM_OPT_MAX_TENSORS]
                    arr = self._tensors(chunk, self.state, "exp_avg", "exp_avg_sq", float(group["weight_decay"] or 0.0))
                    status = L.tm_radam_step(arr, len(chunk), step, group["lr"], group["betas"][0], group["betas"][1], group["eps"], stream)
                    if status != 0:
                        raise capi.TransmilError("tm_radam_step: " + L.tm_last_error().decode())
        return loss


class FusedLookahead(torch.optim.Optimizer):
    """Lookahead(alpha, k) around FusedRAdam, the reference's wrapper (code/MyOptimizer/lookahead.py:12-92) with the slow-weight
    update as one launch (tm_lookahead_sync): every k-th step the slow weights move alpha of the way to the fast weights and
    the fast weights are reset to them; the slow buffer is created at the first sync as a copy of the fast weights of that
    moment (lookahead.py:33-35).  Like the reference it IS a torch Optimizer (LR schedulers such as ReduceLROnPlateau,
    model_interface.py:873, and checkpointing accept it), shares `param_groups` with the base optimizer, keeps the slow buffers
    in its own `state` and writes them as `slow_state` next to the base optimizer's `state` (lookahead.py:53-66)."""

    def __init__(self, base, alpha=0.5, k=6):
        if not 0.0 <= alpha <= 1.0:
            raise ValueError(f"Invalid slow update rate: {alpha}")
        if not 1 <= k:
            raise ValueError(f"Invalid lookahead steps: {k}")
        defaults = dict(lookahead_alpha=alpha, lookahead_k=k, lookahead_step=0)
        self.base_optimizer = self.base = base
        self.param_groups = base.param_groups
        self.defaults = base.defaults
        self.defaults.update(defaults)
        self.state = defaultdict(dict)
        for name, default in defaults.items():
            for group in self.param_groups:
                group.setdefault(name, default)
        # what torch.optim.Optimizer.__init__ would have set up (hooks used by LR schedulers / Lightning)
        self._optimizer_step_pre_hooks, self._optimizer_step_post_hooks = OrderedDict(), OrderedDict()
        self._optimizer_state_dict_pre_hooks, self._optimizer_state_dict_post_hooks = OrderedDict(), OrderedDict()
        self._optimizer_load_state_dict_pre_hooks, self._optimizer_load_state_dict_post_hooks = OrderedDict(), OrderedDict()

    def zero_grad(self, set_to_none=True):
        self.base.zero_grad(set_to_none=set_to_none)

    def add_param_group(self, param_group):
        self.base.add_param_group(param_group)
        for name in ("lookahead_alpha", "lookahead_k", "lookahead_step"):
            self.param_groups[-1].setdefault(name, self.defaults[name])

    def step(self, closure=None):
        loss = self.base.step(closure)
        for group in self.param_groups:
            group["lookahead_step"] += 1
            if group["lookahead_step"] % group["lookahead_k"] == 0:
                self.update_slow(group)
        return loss

    @torch.no_grad()
    def update_slow(self, group):
        b = self.base
        L = b._lib if b._lib is not None else capi.lib()
        ps = [p for p in group["params"] if p.grad is not None]
        if not ps:
            return
        for p in ps:
            if "slow_buffer" not in self.state[p]:
                self.state[p]["slow_buffer"] = p.data.clone(memory_format=torch.contiguous_format)
        stream = torch.cuda.current_stream(ps[0].device).cuda_stream if ps[0].is_cuda else None
        for i0 in range(0, len(ps), capi.TM_OPT_MAX_TENSORS):
            chunk = ps[i0:i0 + capi.TM_OPT_MAX_TENSORS]
            arr = b._tensors(chunk, self.state, "slow_buffer", None, 0.0)
            status = L.tm_lookahead_sync(arr, len(chunk), float(group["lookahead_alpha"]), stream)
            if status != 0:
                raise capi.TransmilError("tm_lookahead_sync: " + L.tm_last_error().decode())

    def sync_lookahead(self):
        for group in self.param_groups:
            self.update_slow(group)

    sync = sync_lookahead

    def state_dict(self):
        fast = self.base.state_dict()
        index = {id(p): i for i, p in enumerate(p for g in self.param_groups for p in g["params"])}
        slow_state = {index[id(p)]: v for p, v in self.state.items()}
        return {"state": fast["state"], "slow_state": slow_state, "param_groups": fast["param_groups"]}

    def load_state_dict(self, state_dict):
        self.base.load_state_dict({"state": state_dict["state"], "param_groups": state_dict["param_groups"]})
        self.param_groups = self.base.param_groups            # both refer to the same container again
        params = [p for g in self.param_groups for p in g["params"]]
        self.state = defaultdict(dict)
        for idx, v in state_dict.get("slow_state", {}).items():
            p = params[int(idx)]
            self.state[p] = {k: (t.to(device=p.device, dtype=p.dtype) if torch.is_tensor(t) else t) for k, t in v.items()}
        for name in ("lookahead_alpha", "lookahead_k", "lookahead_step"):
            for group in self.param_groups:
                group.setdefault(name, self.defaults[name])


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
