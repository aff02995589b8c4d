"""Error budget: logits error of the path when the operands of each contraction are rounded
to tf32 / fp16 / bf16 (fp32 accumulate emulated by fp32 matmul of rounded operands).
Run on CPU:  python tools/precision_budget.py [n] [D]
TEST/DESIGN TOOL ONLY (imports the oracle)."""
import math
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import transmil_oracle as O
import transmil_b200 as T


def r_tf32(x):
    xi = x.contiguous().view(torch.int32)
    r = ((xi >> 13) & 1) + 0xFFF
    return ((xi + r) & ~0x1FFF).view(torch.float32)


def r_fp16(x):
    return x.half().float()


def r_bf16(x):
    return x.bfloat16().float()


def ident(x):
    return x


def forward(sd, x, R, which, pinv_R=ident):
    """which: set of stage names whose operands are rounded with R."""
    def rr(name, t):
        return R(t) if name in which else t
    h = x
    D = x.shape[-1]
    if "_fc1.3.weight" in sd and D == 2048:
        h = F.gelu(F.linear(rr("fc1", h), rr("fc1", sd["_fc1.0.weight"]), sd["_fc1.0.bias"]))
        h = F.layer_norm(h, (h.shape[-1],), sd["_fc1.2.weight"], sd["_fc1.2.bias"], 1e-5)
        h = F.gelu(F.linear(rr("fc1", h), rr("fc1", sd["_fc1.3.weight"]), sd["_fc1.3.bias"]))
    else:
        h = F.gelu(F.linear(rr("fc1", h), rr("fc1", sd["_fc1.0.weight"]), sd["_fc1.0.bias"]))
    h, p = O.square_pad_cls(sd, h)
    for li, name in enumerate(("layer1", "layer2")):
        xn = F.layer_norm(h, (512,), sd[name + ".norm.weight"], sd[name + ".norm.bias"], 1e-5)
        B, S, _ = xn.shape
        P = p["P"]; Sp = p["Sp"]; l = p["l"]
        xp = F.pad(xn, (0, 0, P, 0))
        pre = name + ".attn."
        qkv = F.linear(rr("qkv", xp), rr("qkv", sd[pre + "to_qkv.weight"]))
        q, k, v = qkv.chunk(3, dim=-1)
        q, k, v = (t.reshape(B, Sp, 8, 64).permute(0, 2, 1, 3) for t in (q, k, v))
        q = q * 0.125
        # storage rounding of q,k,v (what the kernels will read)
        q, k, v = rr("store_qkv", q), rr("store_qkv", k), rr("store_qkv", v)
        q_lm = q.reshape(B, 8, 256, l, 64).sum(3) / l
        k_lm = k.reshape(B, 8, 256, l, 64).sum(3) / l
        attn1 = (rr("sim1", q) @ rr("sim1", k_lm).transpose(-1, -2)).softmax(-1)
        attn2 = (q_lm @ k_lm.transpose(-1, -2)).softmax(-1)
        s3 = rr("sim3", q_lm) @ rr("sim3", k).transpose(-1, -2)
        s3 = s3 - s3.amax(-1, keepdim=True)
        e3 = s3.exp()
        # pinv
        A = attn2
        absA = A.abs()
        scale = absA.sum(-1).max() * absA.sum(-2).max()
        Z = A.transpose(-1, -2) / scale
        I = torch.eye(256).unsqueeze(0)
        for _ in range(6):
            XZ = pinv_R(A) @ pinv_R(Z)
            T1 = 7 * XZ - pinv_R(XZ) @ pinv_R(XZ)
            T2 = 15 * XZ - pinv_R(XZ) @ pinv_R(T1)
            Z = 3.25 * Z - 0.25 * (pinv_R(Z) @ pinv_R(T2))
        W3 = (rr("pv3", e3) @ rr("pv3", v)) / e3.sum(-1, keepdim=True)
        M = Z @ W3
        out = rr("pm", attn1) @ rr("pm", M)
        conv = F.conv2d(v, rr("convw", sd[pre + "res_conv.weight"]), None, padding=(16, 0), groups=8)
        out = out + conv
        merged = out.permute(0, 2, 1, 3).reshape(B, Sp, 512)
        y = F.linear(rr("to_out", merged), rr("to_out", sd[pre + "to_out.0.weight"]), sd[pre + "to_out.0.bias"])
        h = h + y[:, -S:]
        if li == 0:
            h = O.ppeg(sd, h, p["G"])
    hn = F.layer_norm(h, (512,), sd["norm.weight"], sd["norm.bias"], 1e-5)[:, 0]
    return F.linear(hn, sd["_fc.weight"], sd["_fc.bias"])


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    torch.manual_seed(2021)
    sd = {k: v.detach() for k, v in T.TransMIL(n_classes=2, in_features=D).state_dict().items()}
    ALL = {"fc1", "qkv", "sim1", "sim3", "pv3", "pm", "to_out", "store_qkv", "convw"}
    for gen in ("rand", "randn"):
        g = torch.Generator().manual_seed(7)
        x = (torch.rand if gen == "rand" else torch.randn)(1, n, D, generator=g)
        with torch.no_grad():
            ref = O.forward(sd, x)
            ref64 = O.forward(sd, x, dtype=torch.float64).float()
            print(f"n={n} D={D} {gen}: fp32-vs-fp64 {float((ref-ref64).abs().max()):.2e}  logits {ref.tolist()}")
            for nm, R in (("tf32", r_tf32), ("fp16", r_fp16), ("bf16", r_bf16)):
                full = forward(sd, x, R, ALL)
                e = float((full - ref).abs().max())
                ep = float((full.softmax(1) - ref.softmax(1)).abs().max())
                line = f"  {nm}: ALL dlogit={e:.2e} dprob={ep:.2e} |"
                for w in sorted(ALL):
                    o = forward(sd, x, R, {w})
                    line += f" {w}={float((o-ref).abs().max()):.1e}"
                pz = forward(sd, x, ident, set(), pinv_R=R)
                line += f" | pinv-only={float((pz-ref).abs().max()):.1e}"
                print(line)


if __name__ == "__main__":
    main()
