"""MEASUREMENT TOOL, not product: SURVEY 8d/D6 -- "stock PyTorch on B200", the thing to beat on the same box.

The reference ships no kernels: on a GPU it runs TransMIL.forward through torch eager (cuBLAS / cuDNN / ATen).  This
file states that forward as a plain `nn.Module` in ordinary torch calls (same state_dict keys as
code/models/TransMIL.py, attention per nystrom-attention 0.0.11, dense [Sp,Sp] attention not materialised = the
survey's "variant B"), loads the weights of the library's module, and times it with CUDA events in the three
precisions the survey names: fp32 ('highest'), TF32 ('high', the reference's own setting, code/train.py:94) and bf16
autocast.  One JSON line per mode, with the library's own time for the same bag beside it.

    python tools/bench_torch_eager.py --n 50000 --features 1024 [--steps 10 --warmup 3] [--modes fp32,tf32,bf16]

tests/test_torch_eager_tool.py checks the module against the CPU oracle (it must be the same function).
"""
import argparse
import json
import math
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class StockAttention(nn.Module):
    def __init__(self, dim=512, heads=8, landmarks=256, iters=6, kernel=33):
        super().__init__()
        self.heads, self.m, self.iters = heads, landmarks, iters
        self.to_qkv = nn.Linear(dim, dim * 3, bias=False)
        self.to_out = nn.Sequential(nn.Linear(dim, dim), nn.Dropout(0.7))
        self.res_conv = nn.Conv2d(heads, heads, (kernel, 1), padding=(kernel // 2, 0), groups=heads, bias=False)

    def forward(self, x):
        B, S, dim = x.shape
        h, m = self.heads, self.m
        P = (m - S % m) % m
        x = F.pad(x, (0, 0, P, 0))
        Sp = S + P
        q, k, v = (t.reshape(B, Sp, h, dim // h).transpose(1, 2) for t in self.to_qkv(x).chunk(3, dim=-1))
        q = q * (dim // h) ** -0.5
        l = math.ceil(S / m)
        q_lm = q.reshape(B, h, m, l, -1).sum(3) / l
        k_lm = k.reshape(B, h, m, l, -1).sum(3) / l
        attn1 = (q @ k_lm.transpose(-1, -2)).softmax(-1)
        attn2 = (q_lm @ k_lm.transpose(-1, -2)).softmax(-1)
        attn3 = (q_lm @ k.transpose(-1, -2)).softmax(-1)
        a = attn2.abs()
        z = attn2.transpose(-1, -2) / (a.sum(-1).max() * a.sum(-2).max())
        eye = torch.eye(m, device=x.device, dtype=attn2.dtype)[None]
        for _ in range(self.iters):
            xz = attn2 @ z
            z = 0.25 * z @ (13 * eye - (xz @ (15 * eye - (xz @ (7 * eye - xz)))))
        out = (attn1 @ z) @ (attn3 @ v) + self.res_conv(v)
        out = self.to_out(out.transpose(1, 2).reshape(B, Sp, dim))
        return out[:, -S:]


class StockLayer(nn.Module):
    def __init__(self, dim=512):
        super().__init__()
        self.norm = nn.LayerNorm(dim)
        self.attn = StockAttention(dim)

    def forward(self, x):
        return x + self.attn(self.norm(x))


class StockPPEG(nn.Module):
    def __init__(self, dim=512):
        super().__init__()
        self.proj = nn.Conv2d(dim, dim, 7, 1, 3, groups=dim)
        self.proj1 = nn.Conv2d(dim, dim, 5, 1, 2, groups=dim)
        self.proj2 = nn.Conv2d(dim, dim, 3, 1, 1, groups=dim)

    def forward(self, x, G):
        B, _, C = x.shape
        cls, feat = x[:, :1], x[:, 1:]
        img = feat.transpose(1, 2).contiguous().view(B, C, G, G)
        img = self.proj(img) + img + self.proj1(img) + self.proj2(img)
        return torch.cat((cls, img.flatten(2).transpose(1, 2)), dim=1)


class StockTransMIL(nn.Module):
    """Takes its sub-module layout (and so its state_dict keys) from a module of the library."""

    def __init__(self, like):
        super().__init__()
        import copy
        self.pos_layer = StockPPEG()
        self._fc1 = copy.deepcopy(like._fc1)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, 512))
        self.layer1, self.layer2 = StockLayer(), StockLayer()
        self.norm = nn.LayerNorm(512)
        self._fc = nn.Linear(512, like.n_classes)
        sd = {k.replace(like.HEAD + ".", "_fc."): v for k, v in like.state_dict().items()}
        self.load_state_dict(sd)

    def forward(self, x):
        h = self._fc1(x.float())
        n = h.shape[1]
        G = int(math.ceil(math.sqrt(n)))
        h = torch.cat([h, h[:, :G * G - n]], dim=1)
        h = torch.cat((self.cls_token.expand(h.shape[0], -1, -1), h), dim=1)
        h = self.layer2(self.pos_layer(self.layer1(h), G))
        return self._fc(self.norm(h)[:, 0])


def measure(ours, x, steps=5, warmup=2, modes=("fp32", "tf32", "bf16"), oracle_logits=None):
    """D6 for bench.py: the stock-torch module with the weights of `ours` on the device bag x [B,n,D]; median ms per step of each
    precision mode (L2 flushed between iterations), the logit gap to the library and, when given, to the CPU oracle."""
    dev = x.device
    stock = StockTransMIL(ours).eval().to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    prev_prec, prev_cudnn = torch.get_float32_matmul_precision(), torch.backends.cudnn.allow_tf32

    def timed(fn):
        for _ in range(warmup):
            fn()
        ms = []
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn()
            e1.record()
            torch.cuda.synchronize(dev)
            ms.append(e0.elapsed_time(e1))
        return sorted(ms)[len(ms) // 2], out

    res = {"what": "SURVEY 8d/D6: the same forward as a plain torch eager module (cuBLAS / cuDNN / ATen, dense [Sp,Sp] attention not "
                   "materialised) on this GPU; tf32 = torch.set_float32_matmul_precision('high'), the reference's setting (code/train.py:94)",
           "torch": torch.__version__, "modes": {}}
    try:
        with torch.no_grad():
            lib_ms, lib_logits = timed(lambda: ours(x))
            res["library_ms_per_step"] = lib_ms
            for mode in modes:
                torch.set_float32_matmul_precision("highest" if mode == "fp32" else "high")
                torch.backends.cudnn.allow_tf32 = mode != "fp32"
                fn = (lambda: torch.autocast("cuda", torch.bfloat16)(stock)(x)) if mode == "bf16" else (lambda: stock(x))
                ms, logits = timed(fn)
                m = {"ms_per_step": ms, "library_speedup": ms / lib_ms,
                     "max_abs_dlogit_vs_library": float((logits.float() - lib_logits).abs().max())}
                if oracle_logits is not None:
                    m["max_abs_dlogit_vs_oracle"] = float((logits.float().cpu() - oracle_logits).abs().max())
                res["modes"][mode] = m
    finally:
        torch.set_float32_matmul_precision(prev_prec)
        torch.backends.cudnn.allow_tf32 = prev_cudnn
    return res


def main():
    sys.path.insert(0, ROOT)
    import transmil_b200 as T
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=50000)
    ap.add_argument("--features", type=int, default=1024)
    ap.add_argument("--classes", type=int, default=2)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--modes", default="fp32,tf32,bf16")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(2021)
    ours = T.TransMIL(a.classes, a.features).eval().to(dev)
    stock = StockTransMIL(ours).eval().to(dev)
    x = torch.rand(1, a.n, a.features, generator=torch.Generator().manual_seed(1)).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn):
        for _ in range(a.warmup):
            fn()
        ms = []
        for _ in range(a.steps):
            flush.zero_()                                   # L2 flush between timed iterations
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            out = fn()
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        return sorted(ms)[len(ms) // 2], out

    with torch.no_grad():
        lib_ms, lib_logits = timed(lambda: ours(x))
        for mode in a.modes.split(","):
            torch.set_float32_matmul_precision("highest" if mode == "fp32" else "high")
            torch.backends.cudnn.allow_tf32 = mode != "fp32"
            if mode == "bf16":
                fn = lambda: torch.autocast("cuda", torch.bfloat16)(stock)(x)
            else:
                fn = lambda: stock(x)
            ms, logits = timed(fn)
            print(json.dumps({"tool": "torch_eager_baseline", "mode": mode, "n": a.n, "features": a.features,
                              "ms_per_bag_torch_eager": ms, "ms_per_bag_library": lib_ms, "ratio": ms / lib_ms,
                              "max_abs_logit_diff_vs_library": float((logits.float() - lib_logits).abs().max()),
                              "torch": torch.__version__, "device": torch.cuda.get_device_name(0)}))


if __name__ == "__main__":
    main()
