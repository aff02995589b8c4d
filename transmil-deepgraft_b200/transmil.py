"""Drop-in `TransMIL` for the DeepGraft plugin loader (`models.TransMIL`).

Mirrors /root/reference/code/models/TransMIL.py: same constructor
`TransMIL(n_classes, in_features, out_features=512)` (TransMIL.py:79), same
`forward(x, return_attn=False)` (TransMIL.py:167, plus the upstream `data=`
keyword), same sub-module names and state_dict keys/shapes, sub-modules created
in the same order so seeded initialisation is bit-identical (TransMIL.py:91-155).

The sub-modules are parameter containers: the arithmetic of the whole forward
runs in libtransmil_b200.so (hand-written sm_100a CUDA) through the C ABI in
include/transmil_b200.h.  There is no PyTorch or CPU fallback.
"""
import ctypes as C
import math
import os

import numpy as np
import torch
import torch.nn as nn

from . import capi

__all__ = ["TransMIL", "MDMIL", "CTMIL", "TransformerMIL", "TransLayer", "PPEG", "NystromAttention", "LazyAttention"]


class NystromAttention(nn.Module):
    """Parameter container with the layout of nystrom-attention 0.0.11
    (constructor call at TransMIL.py:26-34): to_qkv, to_out.0, res_conv."""

    def __init__(self, dim, dim_head=64, heads=8, num_landmarks=256, pinv_iterations=6,
                 residual=True, residual_conv_kernel=33, eps=1e-8, dropout=0.0):
        super().__init__()
        if (dim, dim_head, heads, num_landmarks, pinv_iterations, residual, residual_conv_kernel) != (512, 64, 8, 256, 6, True, 33):
            raise ValueError("libtransmil_b200 implements the TransMIL configuration only: dim=512, heads=8, "
                             "dim_head=64, num_landmarks=256, pinv_iterations=6, residual conv 33")
        inner = heads * dim_head
        self.eps = eps
        self.num_landmarks = num_landmarks
        self.pinv_iterations = pinv_iterations
        self.heads = heads
        self.scale = dim_head ** -0.5
        self.to_qkv = nn.Linear(dim, inner * 3, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, dim), nn.Dropout(dropout))
        self.residual = residual
        self.res_conv = nn.Conv2d(heads, heads, (residual_conv_kernel, 1), padding=(residual_conv_kernel // 2, 0),
                                  groups=heads, bias=False)


class TransLayer(nn.Module):
    """TransMIL.py:19-57.  `norm` stays a real nn.LayerNorm sub-module (GradCAM hook
    point, code/train_loop.py:87); `attn` holds the NystromAttention parameters."""

    def __init__(self, norm_layer=nn.LayerNorm, dim=512):
        super().__init__()
        self.norm = norm_layer(dim)
        attention_heads = 8
        self.attn = NystromAttention(dim=dim, dim_head=dim // attention_heads, heads=attention_heads,
                                     num_landmarks=dim // 2, pinv_iterations=6, residual=True, dropout=0.7)


class PPEG(nn.Module):
    """TransMIL.py:60-75: parameter container for the 7/5/3 depthwise convs."""

    def __init__(self, dim=512):
        super().__init__()
        self.proj = nn.Conv2d(dim, dim, 7, 1, 7 // 2, groups=dim)
        self.proj1 = nn.Conv2d(dim, dim, 5, 1, 5 // 2, groups=dim)
        self.proj2 = nn.Conv2d(dim, dim, 3, 1, 3 // 2, groups=dim)


def _fc1_1024_else_form():
    """the reference's in_features==1024 branch (TransMIL.py:117-121)
    raises at the first forward; 1024-d inputs use Sequential(Linear(1024,512), GELU())
    (= code/models/MDMIL.py:66 = TransMIL's else branch).  Drawn from a LOCAL generator seeded
    initial_seed()+1 (nn.Linear.reset_parameters restated with generator=), so no global RNG
    state -- CPU or CUDA -- is touched and the init stream of every other parameter is unchanged."""
    g = torch.Generator().manual_seed(torch.initial_seed() + 1)
    lin = torch.nn.utils.skip_init(nn.Linear, 1024, 512)
    with torch.no_grad():
        nn.init.kaiming_uniform_(lin.weight, a=math.sqrt(5), generator=g)
        bound = 1.0 / math.sqrt(1024)
        nn.init.uniform_(lin.bias, -bound, bound, generator=g)
    return nn.Sequential(lin, nn.GELU())


class TransMIL(nn.Module):
    HEAD = "_fc"   # attribute / state_dict name of the classifier Linear (TransMIL.py:155); MDMIL calls it _fc2
    PPEG_NAME = "pos_layer"   # ... and of the PPEG block (TransMIL.py:91); CTMIL / TransformerMIL call it pos_layer_0

    def __init__(self, n_classes, in_features, out_features=512):
        super().__init__()
        if out_features != 512:
            raise ValueError("libtransmil_b200 implements out_features=512 (the only value the reference configs use)")
        norm_layer = nn.LayerNorm
        self.pos_layer = PPEG(dim=out_features)
        if in_features == 2048:      # TransMIL.py:100-111
            self._fc1 = nn.Sequential(
                nn.Linear(in_features, int(in_features / 2)), nn.GELU(), norm_layer(int(in_features / 2)),
                nn.Linear(int(in_features / 2), out_features), nn.GELU())
            self._fc1_kind = capi.TM_FC1_2048
        elif in_features == 1024:    # TransMIL.py:117-121 consumes RNG for two Linears, then is unusable
            nn.Linear(in_features, in_features)
            nn.Linear(in_features, out_features)
            self._fc1 = _fc1_1024_else_form()
            self._fc1_kind = capi.TM_FC1_SINGLE
        elif in_features == 768:     # TransMIL.py:122-126
            self._fc1 = nn.Sequential(
                nn.Linear(in_features, int(in_features)), nn.GELU(), nn.Dropout(p=0.6), norm_layer(in_features),
                nn.Linear(in_features, out_features), nn.GELU(), nn.Dropout(p=0.6), norm_layer(out_features))
            self._fc1_kind = capi.TM_FC1_768
        else:                        # TransMIL.py:128-133
            self._fc1 = nn.Sequential(nn.Linear(in_features, out_features), nn.GELU())
            self._fc1_kind = capi.TM_FC1_SINGLE
        self.cls_token = nn.Parameter(torch.randn(1, 1, out_features))
        self.n_classes = n_classes
        self.in_features = in_features
        self.layer1 = TransLayer(norm_layer=norm_layer, dim=out_features)
        self.layer2 = TransLayer(norm_layer=norm_layer, dim=out_features)
        self.norm = norm_layer(out_features)
        self._fc = nn.Linear(out_features, self.n_classes)
        self.flags = 0          # capi.TM_FLAG_* bits OR-ed into every call (the public knob is the `engine` property)
        if os.environ.get("TRANSMIL_B200_ENGINE"):
            self.engine = os.environ["TRANSMIL_B200_ENGINE"]
        self.keep_graph = False  # True: differentiate the eval-mode forward too (GradCAM-style consumers); default eval = fast path
        # True: the second TransLayer's out-projection is run for the cls row's tiles only (TransMIL.py:197-204 reads nothing else
        # of it); logits keep their bits.  Off by default because those rows are work the reference does.
        self.skip_unread_rows = False
        self._ws = None         # cached device workspace (uint8 tensor)
        self._stage_key = None

    # ---- arithmetic engine -------------------------------------------------------
    ENGINES = {
        # name: (flag bits, what it computes in, measured error against the fp32 CPU oracle)
        "tensor-core": (0, "tcgen05: tf32 operands for fc1 (the fp32 bag is read as is), fp16 operands for everything behind the first "
                           "LayerNorm (xn, q, k, v, landmarks, softmax probabilities, the 6 pinv iterations, merged, to_qkv / to_out "
                           "weights), fp32 accumulation, statistics, softmax and residual stream.  Default-init weights: max|dlogit| "
                           "<= 2e-4, max|dY_prob| <= 5e-5 (50 k ... 200 k patches); weights after training steps: tests/"
                           "test_zzzz_trained_weights_gpu.py; to_qkv x 8 (\"sharp softmax\" stress set): up to 2.4e-2, the same size as "
                           "TF32 -- the reference's own GPU setting, code/train.py:94 -- gives there (DESIGN.md, Precision)"),
        "fp32": (capi.TM_FLAG_SIMT_ONLY, "fp32 CUDA-core kernels everywhere (no tensor cores): the on-device cross-check, 1e-5-class "
                                         "agreement with the oracle, about 30x slower at 50 k patches"),
    }

    @property
    def engine(self):
        """Which arithmetic the forward uses: "tensor-core" (default) or "fp32" (see TransMIL.ENGINES for what each computes in
        and its measured error).  Settable; the environment variable TRANSMIL_B200_ENGINE sets the initial value."""
        return "fp32" if int(self.flags) & capi.TM_FLAG_SIMT_ONLY else "tensor-core"

    @engine.setter
    def engine(self, name):
        if name not in self.ENGINES:
            raise ValueError(f"engine must be one of {sorted(self.ENGINES)}, not {name!r}")
        self.flags = (int(self.flags) & ~capi.TM_FLAG_SIMT_ONLY) | self.ENGINES[name][0]

    # ---- parameters -> C struct ------------------------------------------------
    def _ptr(self, t, shape=None):
        if not t.is_cuda:
            raise capi.TransmilError("TransMIL parameters must live on a CUDA device (model.cuda()); there is no CPU path")
        if t.dtype != torch.float32 or not t.is_contiguous():
            raise capi.TransmilError("TransMIL parameters must be contiguous float32")
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError(f"parameter has shape {tuple(t.shape)}, expected {tuple(shape)}")
        return t.data_ptr()

    def _weights(self):
        w = capi.Weights()
        w.fc1_kind, w.in_features, w.n_classes = self._fc1_kind, self.in_features, self.n_classes
        f = getattr(self, "_fc1", None)
        D = self.in_features
        if self._fc1_kind == capi.TM_FC1_2048:
            w.fc1_w0, w.fc1_b0 = self._ptr(f[0].weight, (1024, D)), self._ptr(f[0].bias)
            w.fc1_ln_w, w.fc1_ln_b = self._ptr(f[2].weight, (1024,)), self._ptr(f[2].bias)
            w.fc1_w1, w.fc1_b1 = self._ptr(f[3].weight, (512, 1024)), self._ptr(f[3].bias)
        elif self._fc1_kind == capi.TM_FC1_768:
            w.fc1_w0, w.fc1_b0 = self._ptr(f[0].weight, (768, 768)), self._ptr(f[0].bias)
            w.fc1_ln_w, w.fc1_ln_b = self._ptr(f[3].weight, (768,)), self._ptr(f[3].bias)
            w.fc1_w1, w.fc1_b1 = self._ptr(f[4].weight, (512, 768)), self._ptr(f[4].bias)
            w.fc1_ln2_w, w.fc1_ln2_b = self._ptr(f[7].weight, (512,)), self._ptr(f[7].bias)
        elif self._fc1_kind == capi.TM_FC1_TOKENS:
            pass     # the stem runs outside the library (conv stem of CTMIL): the bag already holds 512-d tokens
        else:
            w.fc1_w0, w.fc1_b0 = self._ptr(f[0].weight, (512, D)), self._ptr(f[0].bias)
        w.cls_token = self._ptr(self.cls_token, (1, 1, 512))
        for name in ("layer1", "layer2"):
            layer, lw = getattr(self, name), getattr(w, name)
            lw.norm_w, lw.norm_b = self._ptr(layer.norm.weight, (512,)), self._ptr(layer.norm.bias, (512,))
            lw.qkv_w = self._ptr(layer.attn.to_qkv.weight, (1536, 512))
            lw.out_w, lw.out_b = self._ptr(layer.attn.to_out[0].weight, (512, 512)), self._ptr(layer.attn.to_out[0].bias)
            lw.conv_w = self._ptr(layer.attn.res_conv.weight, (8, 1, 33, 1))
        p = getattr(self, self.PPEG_NAME)
        w.ppeg_w7, w.ppeg_b7 = self._ptr(p.proj.weight, (512, 1, 7, 7)), self._ptr(p.proj.bias)
        w.ppeg_w5, w.ppeg_b5 = self._ptr(p.proj1.weight, (512, 1, 5, 5)), self._ptr(p.proj1.bias)
        w.ppeg_w3, w.ppeg_b3 = self._ptr(p.proj2.weight, (512, 1, 3, 3)), self._ptr(p.proj2.bias)
        w.norm_w, w.norm_b = self._ptr(self.norm.weight, (512,)), self._ptr(self.norm.bias, (512,))
        head = getattr(self, self.HEAD)
        w.fc_w, w.fc_b = self._ptr(head.weight, (self.n_classes, 512)), self._ptr(head.bias)
        return w

    def _workspace(self, nbytes, device):
        ws = self._ws
        if ws is None or ws.device != device or ws.numel() < nbytes:
            self._ws = ws = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
        return ws

    # ---- forward ---------------------------------------------------------------
    def forward(self, x=None, return_attn=False, **kwargs):
        if x is None:
            x = kwargs.get("data")          # upstream TransMIL: forward(**kwargs) with kwargs['data']
        if x is None:
            raise TypeError("forward() needs the bag tensor (positional `x` or keyword `data`)")
        hooked = self._hooked_norms()
        graph = torch.is_grad_enabled() and (self.training or self.keep_graph or bool(hooked))
        if graph and return_attn and not (self.training or self.keep_graph):
            # eval() + hooks + return_attn with autograd on: code/visualize_mil.py:453,580 and code/gradcam_sus.py:570 build GradCAM
            # objects (forward hooks on a norm) and then ask for the attention of the same module.  Inference path; the hooks are
            # fired with the kept activations (no graph on this call)
            graph = False
        if graph:
            # training_step (code/models/model_interface.py:333-383): forward with every stage kept + the library's backward
            # (csrc/backward.cu) behind a torch.autograd.Function; train mode applies the to_out Dropout(0.7)
            if return_attn:
                raise NotImplementedError("libtransmil_b200: return_attn=True is an inference feature (eval() / no_grad())")
            return self._forward_train(x)
        # TransMIL.py:170-174
        if x.dim() > 3:
            x = x.squeeze(0)
        elif x.dim() == 2:
            x = x.unsqueeze(0)
        if x.dim() != 3 or x.shape[-1] != self.in_features:
            raise ValueError(f"expected a bag tensor [B, n, {self.in_features}], got {tuple(x.shape)}")
        device = self.cls_token.device
        if device.type != "cuda":
            raise capi.TransmilError("TransMIL must be on a CUDA device (model.cuda()); there is no CPU path")
        if self.training and not return_attn:
            # train() under no_grad: the reference's Dropout layers (to_out 0.7, the 768-d stem's 0.6) are still active, so this is
            # the training forward without a graph, not the eval fast path
            from . import training
            xd = x.detach().to(device=device, dtype=torch.float32, non_blocking=True).contiguous()
            seeds = tuple(int(s) for s in torch.randint(0, 2 ** 62, (4,)))
            with torch.cuda.device(device):
                logits, saved, _w = training.forward_keep(self, xd, True, seeds)
                if hooked:
                    training.fire_hooks_no_graph(self, hooked, {"layer1": saved["h0"], "layer2": saved["h2"], "norm": saved["h3"]})
            return logits
        L = capi.lib()
        B, n, D = (int(s) for s in x.shape)
        flags = int(self.flags) | (capi.TM_FLAG_KEEP_STAGES if (return_attn or hooked) else 0)
        if getattr(self, "skip_unread_rows", False) and not return_attn:
            flags |= capi.TM_FLAG_CLS_ONLY_L2
        # fp16 feature bags (half-precision feature files) are consumed as they are: x.float() of
        # TransMIL.py:174 is an exact upcast, so fc1 reads the halves directly (half the H2D and HBM bytes)
        half_in = x.dtype == torch.float16 and not (flags & capi.TM_FLAG_SIMT_ONLY)
        if half_in:
            flags |= capi.TM_FLAG_X_F16
        with torch.cuda.device(device):
            w = self._weights()
            nbytes = C.c_size_t()
            capi.check(L.tm_workspace_bytes(B, n, C.byref(w), flags, C.byref(nbytes)))
            ws = self._workspace(nbytes.value, device)
            logits = torch.empty((B, self.n_classes), dtype=torch.float32, device=device)
            stream = torch.cuda.current_stream(device).cuda_stream
            xd = x.detach()
            want = torch.float16 if half_in else torch.float32
            if xd.dtype != want or not xd.is_contiguous() or not xd.is_cuda:
                # x.float() (TransMIL.py:174); host bags are copied on the current stream
                xd = xd.to(device=device, dtype=want, non_blocking=True).contiguous()
            capi.check(L.tm_transmil_fwd(C.byref(w), xd.data_ptr(), B, n, logits.data_ptr(), ws.data_ptr(),
                                         ws.numel(), flags, stream))
            self._stage_key = (B, n, flags)
            if hooked:
                # forward hooks on layer{1,2}.norm / norm without a graph (no_grad, or the return_attn call above): fired with the
                # LayerNorm of the kept layer inputs, as torch fires them in the reference (TransMIL.py:47,202)
                from . import training
                training.fire_hooks_no_graph(self, hooked, {nm: _stage_tensor(self, ws, w, st, B, n, flags)
                                                            for nm, st in (("layer1", "h0"), ("layer2", "h2"), ("norm", "h3"))})
        if return_attn:
            plan = capi.make_plan(n)
            attn = LazyAttention(self, ws, w, B, n, flags, plan)
            self._ws = None   # the lazy view owns this workspace now
            return logits, (attn, int(plan.padding))
        return logits

    HOOK_POINTS = ("layer1", "layer2", "norm")

    def _norm_module(self, name):
        return self.norm if name == "norm" else getattr(self, name).norm

    def _hooked_norms(self):
        """The reference's GradCAM consumers hook `layer1.norm` / `layer2.norm` (code/train_loop.py:87,
        code/test_visualize.py:121-127) or the final `norm` (code/visualize_mil.py:227, code/gradcam_sus.py).  Those sub-modules are
        parameter containers here -- their forward never runs.  Forward hooks on them are honoured all the same: with autograd on by
        the training path (training.forward_with_norm_hooks: activation and its gradient, which is what GradCAM reads), without a
        graph by training.fire_hooks_no_graph (activation only, as under no_grad in the reference); any other hook kind is refused
        rather than silently skipped."""
        hooked = []
        for name in self.HOOK_POINTS:
            norm = self._norm_module(name)
            label = name if name == "norm" else name + ".norm"
            if norm._forward_pre_hooks or norm._backward_hooks or getattr(norm, "_backward_pre_hooks", None):
                raise NotImplementedError(f"libtransmil_b200: only forward hooks are supported on {label}")
            if norm._forward_hooks:
                hooked.append(name)
        return hooked

    def _forward_train(self, x):
        from . import training
        if x.dim() > 3:          # TransMIL.py:170-174
            x = x.squeeze(0)
        elif x.dim() == 2:
            x = x.unsqueeze(0)
        if x.dim() != 3 or x.shape[-1] != self.in_features:
            raise ValueError(f"expected a bag tensor [B, n, {self.in_features}], got {tuple(x.shape)}")
        device = self.cls_token.device
        if device.type != "cuda":
            raise capi.TransmilError("TransMIL must be on a CUDA device (model.cuda()); there is no CPU path")
        hooked = self._hooked_norms()
        if x.requires_grad:
            # the bag (or the tokens of a torch stem, CTMIL) is part of a graph: d(loss)/dx comes back from the library's backward
            # (one more contraction through _fc1), as the reference's autograd would deliver it
            if hooked:
                raise NotImplementedError("libtransmil_b200: norm hooks together with a bag that requires grad are not built")
            return training.TransMILFunction.apply(self, x.to(device=device, dtype=torch.float32).contiguous(), *self._path_parameters())
        xd = x.detach().to(device=device, dtype=torch.float32, non_blocking=True).contiguous()   # x.float(), TransMIL.py:174
        if hooked:
            return training.forward_with_norm_hooks(self, xd, hooked)
        return training.TransMILFunction.apply(self, xd, *self._path_parameters())

    def _path_named_parameters(self):
        """The parameters the library's path reads and differentiates: all of them for TransMIL / MDMIL; the sibling models
        with a torch stem (CTMIL, TransformerMIL) keep further sub-modules the path never touches."""
        return list(self.named_parameters())

    def _path_parameters(self):
        return [p for _, p in self._path_named_parameters()]

    def stage(self, name):
        """Per-stage buffer of the last forward (needs flags |= TM_FLAG_KEEP_STAGES); parity tests only."""
        if self._ws is None or self._stage_key is None:
            raise capi.TransmilError("no forward has been run")
        B, n, flags = self._stage_key
        return _stage_tensor(self, self._ws, self._weights(), name, B, n, flags)


_STAGE_SHAPES = {
    "h": lambda B, p: (B, p.S, 512), "p": lambda B, p: (B, p.Sp, 512), "q": lambda B, p: (B, 8, p.Sp, 64),
    "l": lambda B, p: (B, 8, 256, 64), "m": lambda B, p: (B, 8, 256, 256)}
_STAGE_KIND = {"h0": "h", "h1": "h", "h2": "h", "h3": "h", "xn": "p", "merged": "p", "q": "q", "k": "q", "v": "q",
               "q_lm": "l", "k_lm": "l", "W3": "l", "M": "l", "attn2": "m", "Z": "m",
               "q_lm1": "l", "k_lm1": "l", "W31": "l", "M1": "l", "attn21": "m", "Z1": "m"}


def _stage_tensor(model, ws, w, name, B, n, flags):
    ptr, numel = C.c_void_p(), C.c_int64()
    capi.check(capi.lib().tm_stage_ptr(name.encode(), B, n, C.byref(w), flags, ws.data_ptr(), C.byref(ptr), C.byref(numel)))
    off = ptr.value - ws.data_ptr()
    plan = capi.make_plan(n)
    shape = _STAGE_SHAPES[_STAGE_KIND[name]](B, plan)
    if capi.lib().tm_stage_is_f16(name.encode(), flags):   # q, k, v of the tensor-core engine: returned as an fp32 copy
        return ws[off:off + 2 * numel.value].view(torch.float16).view(shape).float()
    return ws[off:off + 4 * numel.value].view(torch.float32).view(shape)


class LazyAttention:
    """Stands in for the dense `attn1 @ attn2_inv @ attn3` [B,8,Sp,Sp] tensor the
    reference returns with return_attn=True (TransMIL.py:47,209-210).  Consumers read
    single rows (`attn[0, :, padding+1, padding+1:padding+1+H]`, code/visualize_mil.py:581);
    rows are computed on demand by tm_attn_rows_fwd from layer 2's q, k, landmarks and Z."""

    MAX_DENSE_BYTES = 8 << 30

    def __init__(self, model, ws, w, B, n, flags, plan):
        self._ws, self._w, self._B, self._n, self._flags = ws, w, B, n, flags
        self._model = model
        self.Sp = int(plan.Sp)
        self.shape = (B, 8, self.Sp, self.Sp)
        self.device = ws.device
        self.dtype = torch.float32

    def size(self, dim=None):
        return self.shape if dim is None else self.shape[dim]

    def dim(self):
        return 4

    def rows(self, rows):
        """[B, 8, len(rows), Sp] for the given padded row indices."""
        rows = [int(r) % self.Sp for r in rows]
        L = capi.lib()
        st = {k: _stage_tensor(self._model, self._ws, self._w, k, self._B, self._n, self._flags)
              for k in ("q", "k", "q_lm", "k_lm", "Z")}
        out = torch.empty((self._B, 8, len(rows), self.Sp), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            for i0 in range(0, len(rows), 16):
                chunk = rows[i0:i0 + 16]
                nr = len(chunk)
                scratch = torch.empty(int(L.tm_attn_rows_scratch_floats(self._B, self.Sp, nr)), dtype=torch.float32,
                                      device=self.device)
                tmp = torch.empty((self._B, 8, nr, self.Sp), dtype=torch.float32, device=self.device)
                arr = (C.c_int64 * nr)(*chunk)
                capi.check(L.tm_attn_rows_fwd(st["q"].data_ptr(), st["k"].data_ptr(), st["q_lm"].data_ptr(),
                                              st["k_lm"].data_ptr(), st["Z"].data_ptr(), self._B, self.Sp, arr, nr,
                                              tmp.data_ptr(), scratch.data_ptr(), stream))
                out[:, :, i0:i0 + nr] = tmp
        return out

    def __getitem__(self, idx):
        if not isinstance(idx, tuple):
            idx = (idx,)
        idx = idx + (slice(None),) * (4 - len(idx))
        b, h, r, c = idx
        if isinstance(r, slice):
            rr = list(range(*r.indices(self.Sp)))
            squeeze = False
        elif isinstance(r, (list, tuple)):
            rr, squeeze = list(r), False
        else:
            rr, squeeze = [int(r)], True
        full = self.rows(rr)                      # [B,8,len,Sp]
        full = full[b, h] if not (isinstance(b, slice) and isinstance(h, slice)) else full
        # after indexing b,h the row axis moved left by the number of integer indices among (b,h)
        lead = sum(1 for t in (b, h) if isinstance(t, slice))
        sel = (slice(None),) * lead + ((0,) if squeeze else (slice(None),)) + (c,)
        return full[sel]

    def dense(self):
        nbytes = self._B * 8 * self.Sp * self.Sp * 4
        if nbytes > self.MAX_DENSE_BYTES:
            raise MemoryError(f"dense attention would need {nbytes / 2**30:.1f} GiB; index rows instead")
        return self.rows(range(self.Sp))


class MDMIL(TransMIL):
    """Sibling model of the same path: /root/reference/code/models/MDMIL.py:60-113.

    `MDMIL(n_classes)` is TransMIL's path with a fixed 1024-d single-Linear stem (MDMIL.py:63-66), the classifier named
    `_fc2` (MDMIL.py:74) and a forward that always returns `(logits, attn2)` (MDMIL.py:77,113) -- layer 2's
    `attn1 @ attn2_inv @ attn3` [B,8,Sp,Sp], which the reference materialises densely and this module returns as lazy
    rows (LazyAttention).  Sub-modules are created in the reference's order (pos_layer, _fc1, cls_token, layer1, layer2,
    norm, _fc2) so the seeded init and the state_dict keys are identical.  Same kernels, no new arithmetic.
    In train mode (autograd on) the forward returns `(logits, None)`: the dense attention has no gradient consumer in the
    reference and the training forward does not keep it."""
    HEAD = "_fc2"

    def __init__(self, n_classes):
        nn.Module.__init__(self)
        in_features, out_features = 1024, 512
        self.pos_layer = PPEG(dim=out_features)
        self._fc1 = nn.Sequential(nn.Linear(in_features, out_features), nn.GELU())
        self._fc1_kind = capi.TM_FC1_SINGLE
        self.cls_token = nn.Parameter(torch.randn(1, 1, out_features))
        self.n_classes = n_classes
        self.in_features = in_features
        self.layer1 = TransLayer(dim=out_features)
        self.layer2 = TransLayer(dim=out_features)
        self.norm = nn.LayerNorm(out_features)
        self._fc2 = nn.Linear(out_features, self.n_classes)
        self.flags = 0
        self.keep_graph = False
        self._ws = None
        self._stage_key = None

    def forward(self, x):
        if torch.is_grad_enabled() and (self.training or self.keep_graph or self._hooked_norms()):
            return TransMIL.forward(self, x), None
        logits, (attn, _padding) = TransMIL.forward(self, x, return_attn=True)
        return logits, attn


class CTMIL(TransMIL):
    """Sibling model of the same path: /root/reference/code/models/CTMIL.py:76-165.

    `CTMIL(n_classes, in_features)` takes image-like tiles `[1, N, in_features, H, H]`, runs a two-stage conv stem on every
    tile (CTMIL.py:90-101,137-142: Conv3x3 -> BatchNorm -> GELU -> MaxPool(3,2,1), twice), REINTERPRETS the `[N,512,h,h]`
    result as `[N, h*h, 512]` tokens with `.view` (CTMIL.py:142 -- a view, not a transpose), and sends every tile as its own
    bag through duplicate pad + cls, TransLayer, PPEG, TransLayer, LayerNorm, `_fc2` (CTMIL.py:143-165): logits `[N, n_classes]`.
    The conv stem is four stock torch modules (cuDNN / ATen: library code, outside SURVEY 8's hot path; "zero new kernels");
    everything from the tokens on is ONE `tm_transmil_fwd` call with `TM_FC1_TOKENS`.  Sub-modules are created in the
    reference's order (pos_layer_0, conv1, conv2, _fc1 -- which the reference builds but never calls --, cls_token, layer1,
    layer2, norm, _fc2) so the seeded init and the state_dict keys are identical.  With autograd on, d(loss)/d(tokens) comes back
    from the library's backward and torch differentiates the conv stem."""
    HEAD = "_fc2"
    PPEG_NAME = "pos_layer_0"
    _PATH_PREFIXES = ("pos_layer_0.", "cls_token", "layer1.", "layer2.", "norm.", "_fc2.")

    def __init__(self, n_classes, in_features, out_features=512):
        nn.Module.__init__(self)
        if out_features != 512:
            raise ValueError("libtransmil_b200 implements out_features=512 (the only value the reference configs use)")
        self.pos_layer_0 = PPEG(dim=out_features)
        norm_layer = nn.LayerNorm
        self._make_conv_stem(in_features, out_features)
        if in_features == 2048:      # CTMIL.py:103-108 (built, never called by forward)
            self._fc1 = nn.Sequential(nn.Linear(in_features, int(in_features / 2), bias=True),
                                      nn.Linear(int(in_features / 2), out_features), nn.GELU())
        elif in_features == 1024:    # CTMIL.py:109-113
            self._fc1 = nn.Sequential(nn.Linear(in_features, out_features), nn.GELU(), nn.Dropout(p=0.6), norm_layer(out_features))
        elif in_features == 768:     # CTMIL.py:114-115
            self._fc1 = nn.Sequential(nn.Linear(in_features, 512, bias=True), nn.ReLU())
        self._finish(n_classes, out_features)

    def _make_conv_stem(self, in_features, out_features):
        half = int(in_features / 2)  # CTMIL.py:90-101
        self.conv1 = nn.Sequential(nn.Conv2d(in_features, half, kernel_size=3, stride=1, padding=1, bias=False),
                                   nn.BatchNorm2d(half), nn.GELU(), nn.MaxPool2d(kernel_size=3, stride=2, padding=1))
        self.conv2 = nn.Sequential(nn.Conv2d(half, out_features, kernel_size=3, stride=1, padding=1, bias=False),
                                   nn.BatchNorm2d(out_features), nn.GELU(), nn.MaxPool2d(kernel_size=3, stride=2, padding=1))

    def _finish(self, n_classes, out_features):
        self.cls_token = nn.Parameter(torch.randn(1, 1, out_features))
        self.n_classes = n_classes
        self.in_features = out_features          # what the library's path is fed: 512-d tokens
        self._fc1_kind = capi.TM_FC1_TOKENS
        self.layer1 = TransLayer(dim=out_features)
        self.layer2 = TransLayer(dim=out_features)
        self.norm = nn.LayerNorm(out_features)
        self._fc2 = nn.Linear(out_features, self.n_classes)
        self.flags = 0
        self.keep_graph = False
        self.skip_unread_rows = False
        self._ws = None
        self._stage_key = None

    def _path_named_parameters(self):
        return [(n, p) for n, p in self.named_parameters() if n.startswith(self._PATH_PREFIXES)]

    def tokens(self, x):
        """CTMIL.py:137-142: tiles [1,N,C,H,H] (or [N,C,H,H]) -> tokens [N, h*h, 512] through the torch conv stem."""
        if x.dim() == 5:
            x = x.squeeze(0)
        if x.dim() != 4:
            raise ValueError(f"expected image tiles [1, N, C, H, H], got {tuple(x.shape)}")
        device = self.cls_token.device
        if device.type != "cuda":
            raise capi.TransmilError("CTMIL must be on a CUDA device (model.cuda()); there is no CPU path")
        h = self.conv2(self.conv1(x.to(device=device, dtype=torch.float32)))
        return h.view(h.shape[0], h.shape[2] * h.shape[2], h.shape[1])

    def forward(self, x):
        return TransMIL.forward(self, self.tokens(x))


class _PreNorm(nn.Module):
    """Parameter container of code/models/_transformer.py:6-13."""

    def __init__(self, dim, fn):
        super().__init__()
        self.norm = nn.LayerNorm(dim)
        self.fn = fn


class _Attention(nn.Module):
    """Parameter container of code/models/_transformer.py:16-42 (plain softmax attention: not on the Nystrom path)."""

    def __init__(self, dim, heads, dim_head, dropout):
        super().__init__()
        inner = dim_head * heads
        self.to_qkv = nn.Linear(dim, inner * 3, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, dim), nn.Dropout(dropout))


class _FeedForward(nn.Module):
    """Parameter container of code/models/_transformer.py:45-57."""

    def __init__(self, dim, hidden_dim, dropout):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(dim, hidden_dim), nn.GELU(), nn.Dropout(dropout), nn.Linear(hidden_dim, dim), nn.Dropout(dropout))


class _Transformer(nn.Module):
    """Parameter container of TransformerMIL.py:18-26."""

    def __init__(self, dim, depth, heads, dim_head, mlp_dim, dropout=0.0):
        super().__init__()
        self.layers = nn.ModuleList([nn.ModuleList([_PreNorm(dim, _Attention(dim, heads, dim_head, dropout)),
                                                    _PreNorm(dim, _FeedForward(dim, mlp_dim, dropout))]) for _ in range(depth)])


class TransformerMIL(CTMIL):
    """The Nystrom forward of /root/reference/code/models/TransformerMIL.py.

    The reference file holds two forwards: the live one (TransformerMIL.py:137-151, a plain softmax-attention transformer over
    `fc1` features -- not the Nystrom path, not built: `forward_plain` raises) and the commented-out one
    (TransformerMIL.py:155-183), which is CTMIL's: conv stem on tiles, tokens by `.view`, duplicate pad + cls, TransLayer,
    PPEG, TransLayer, LayerNorm, `_fc2`.  `forward` here is that Nystrom forward on this class's own parameters.  The
    constructor mirrors TransformerMIL.py:76-135 sub-module by sub-module (pos_layer_0, conv1, conv2, fc1, cls_token, layer1,
    layer2, norm, _fc2, transformer1, transformer2) so the seeded init and the state_dict keys are identical; `fc1`,
    `transformer1/2` are parameter containers the Nystrom forward never reads."""

    def __init__(self, n_classes, in_features, out_features=512):
        nn.Module.__init__(self)
        if out_features != 512:
            raise ValueError("libtransmil_b200 implements out_features=512 (the only value the reference configs use)")
        self.pos_layer_0 = PPEG(dim=out_features)
        norm_layer = nn.LayerNorm
        self._make_conv_stem(in_features, out_features)
        if in_features == 2048:      # TransformerMIL.py:106-110
            self.fc1 = nn.Sequential(nn.Linear(in_features, int(in_features / 2)), nn.GELU(), nn.Dropout(p=0.6), norm_layer(int(in_features / 2)),
                                     nn.Linear(int(in_features / 2), out_features), nn.GELU())
        elif in_features == 1024:    # TransformerMIL.py:111-115
            self.fc1 = nn.Sequential(nn.Linear(in_features, out_features), nn.GELU(), nn.Dropout(p=0.6), norm_layer(out_features))
        elif in_features in (768, 384):   # TransformerMIL.py:116-119
            self.fc1 = nn.Sequential(nn.Linear(in_features, 512, bias=True), nn.ReLU())
        self._finish(n_classes, out_features)
        dropout, emb_dropout = 0.5, 0.5   # TransformerMIL.py:129-135
        self.transformer1 = _Transformer(dim=out_features, depth=2, dim_head=64, heads=8, mlp_dim=512, dropout=dropout)
        self.transformer2 = _Transformer(dim=out_features, depth=2, dim_head=64, heads=8, mlp_dim=512, dropout=dropout)
        self.dropout = nn.Dropout(emb_dropout)
        self.to_latent = nn.Identity()
        self.pool = "cls"

    def forward_plain(self, x):
        raise NotImplementedError("libtransmil_b200: the plain softmax-attention forward of TransformerMIL.py:137-151 is not on the "
                                  "Nystrom path and is not built; forward() is the Nystrom forward of TransformerMIL.py:155-183")


def y_prob_y_hat(logits):
    """ModelInterface.step post-processing (code/models/model_interface.py:323-324)."""
    return torch.softmax(logits, dim=1), torch.argmax(logits, dim=1)
