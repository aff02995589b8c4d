"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the TransMIL bag-level forward.

A functional restatement (torch CPU ops, fp32 or fp64) of
/root/reference/code/models/TransMIL.py:167-211 plus the un-vendored
nystrom-attention 0.0.11 (see oracle/nystrom_attention/__init__.py), driven by
a reference-layout state_dict.  Unlike oracle/ref_loader.py it does not need
/root/reference, so it travels to the GPU box, and it exposes every per-stage
intermediate the parity tests compare against.

Pinning: tests/test_oracle.py checks this restatement (a) bit-for-bit against
the unmodified reference module executed in the build container
(oracle/ref_loader.py) whenever /root/reference exists and (b) against the
committed fixtures tests/golden/*.json that were generated FROM the reference
module by tests/golden/make_golden.py.  The NystromAttention arithmetic itself
is "parity unpinned" (third-party, absent, no reference test) -- see DESIGN.md.

The dense [B,8,Sp,Sp] attention the reference always builds (TransMIL.py:47)
is only produced when `return_attn=True`; logits do not depend on it.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference leg may import this module.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

HEADS = 8
DIM = 512
DHEAD = 64
LANDMARKS = 256
PINV_ITERS = 6
CONV_K = 33


def plan(n):
    """Index contract of the path for a bag of n patches.

    G, N, dup, S: TransMIL.py:177-179 (float64 sqrt, ceil); cls at TransMIL.py:184-186.
    l, Sp, P: nystrom-attention 0.0.11 front pad / landmark grouping.
    padding: TransMIL.py:190-193.
    """
    G = int(np.ceil(np.sqrt(n)))
    N = G * G
    dup = N - n
    S = N + 1
    rem = S % LANDMARKS
    P = (LANDMARKS - rem) if rem > 0 else 0
    Sp = S + P
    l = math.ceil(S / LANDMARKS)
    padding = (256 - S % 256) if S % 256 > 0 else 0
    return dict(n=n, G=G, N=N, dup=dup, S=S, l=l, Sp=Sp, P=P, padding=padding)


def fc1(sd, x, masks=None):
    """TransMIL._fc1 (TransMIL.py:100-133).  Eval mode: Dropout = identity.  `masks` = (m0, m4), the training-mode keep masks
    of the two Dropout(0.6) of the 768-d stem (TransMIL.py:123-126), already scaled by 1/(1-p), multiplied onto the GELU
    outputs where the reference's Sequential drops them (pinned to the reference module in train mode by tests/test_oracle.py)."""
    keys = sorted(k for k in sd if k.startswith("_fc1."))
    idx = sorted({int(k.split(".")[1]) for k in keys})
    D = x.shape[-1]
    if idx == [0, 2, 3] and D == 2048:
        # Linear(2048,1024) GELU LayerNorm(1024) Linear(1024,512) GELU   (TransMIL.py:109-110)
        h = F.gelu(F.linear(x, sd["_fc1.0.weight"], sd["_fc1.0.bias"]))
        h = F.layer_norm(h, (h.shape[-1],), sd["_fc1.2.weight"], sd["_fc1.2.bias"], 1e-5)
        h = F.gelu(F.linear(h, sd["_fc1.3.weight"], sd["_fc1.3.bias"]))
        return h
    if idx == [0, 3, 4, 7] and D == 768:
        # Linear GELU Dropout LN(768) Linear GELU Dropout LN(512)  (TransMIL.py:123-126)
        h = F.gelu(F.linear(x, sd["_fc1.0.weight"], sd["_fc1.0.bias"]))
        if masks is not None:
            h = h * masks[0]
        h = F.layer_norm(h, (h.shape[-1],), sd["_fc1.3.weight"], sd["_fc1.3.bias"], 1e-5)
        h = F.gelu(F.linear(h, sd["_fc1.4.weight"], sd["_fc1.4.bias"]))
        if masks is not None:
            h = h * masks[1]
        h = F.layer_norm(h, (h.shape[-1],), sd["_fc1.7.weight"], sd["_fc1.7.bias"], 1e-5)
        return h
    if idx == [0]:
        # Linear(D,512) GELU  (TransMIL.py:130-133; MDMIL.py:66 for D=1024)
        return F.gelu(F.linear(x, sd["_fc1.0.weight"], sd["_fc1.0.bias"]))
    raise RuntimeError(f"unsupported _fc1 layout {idx} for in_features={D}")


def square_pad_cls(sd, h):
    """TransMIL.py:177-186."""
    n = h.shape[1]
    p = plan(n)
    h = torch.cat([h, h[:, :p["dup"], :]], dim=1)
    cls = sd["cls_token"].to(h.dtype).expand(h.shape[0], -1, -1)
    return torch.cat([cls, h], dim=1), p


def pinv(A, iters=PINV_ITERS):
    """moore_penrose_iter_pinv of nystrom-attention 0.0.11 (global scalar)."""
    absA = A.abs()
    scale = absA.sum(-1).max() * absA.sum(-2).max()
    Z = A.transpose(-1, -2) / scale
    I = torch.eye(A.shape[-1], dtype=A.dtype).unsqueeze(0)
    for _ in range(iters):
        XZ = A @ Z
        Z = 0.25 * Z @ (13 * I - (XZ @ (15 * I - (XZ @ (7 * I - XZ)))))
    return Z


def nystrom(sd, prefix, xn, stages=None, return_attn=False):
    """NystromAttention.forward (eval), x already layer-normed.  [B,S,512] -> [B,S,512]."""
    B, S, _ = xn.shape
    m = LANDMARKS
    rem = S % m
    P = m - rem if rem > 0 else 0
    x = F.pad(xn, (0, 0, P, 0), value=0.0)
    Sp = S + P
    qkv = F.linear(x, sd[prefix + "to_qkv.weight"])
    q, k, v = qkv.chunk(3, dim=-1)
    q, k, v = (t.reshape(B, Sp, HEADS, DHEAD).permute(0, 2, 1, 3) for t in (q, k, v))
    q = q * (DHEAD ** -0.5)
    l = math.ceil(S / m)
    q_lm = q.reshape(B, HEADS, m, l, DHEAD).sum(3) / l
    k_lm = k.reshape(B, HEADS, m, l, DHEAD).sum(3) / l
    attn1 = (q @ k_lm.transpose(-1, -2)).softmax(-1)
    attn2 = (q_lm @ k_lm.transpose(-1, -2)).softmax(-1)
    attn3 = (q_lm @ k.transpose(-1, -2)).softmax(-1)
    Z = pinv(attn2)
    W3 = attn3 @ v
    out = (attn1 @ Z) @ W3
    conv = F.conv2d(v, sd[prefix + "res_conv.weight"], None, padding=(CONV_K // 2, 0), groups=HEADS)
    out = out + conv
    merged = out.permute(0, 2, 1, 3).reshape(B, Sp, DIM)
    y = F.linear(merged, sd[prefix + "to_out.0.weight"], sd[prefix + "to_out.0.bias"])
    y = y[:, -S:]
    if stages is not None:
        stages.update(q=q, k=k, v=v, q_lm=q_lm, k_lm=k_lm, attn2=attn2, Z=Z, W3=W3,
                      merged=merged, P=P, l=l)
    if return_attn:
        return y, attn1 @ Z @ attn3
    return y


def trans_layer(sd, name, h, stages=None, return_attn=False):
    """TransLayer.forward (TransMIL.py:45-57): x + attn(norm(x))."""
    xn = F.layer_norm(h, (DIM,), sd[name + ".norm.weight"], sd[name + ".norm.bias"], 1e-5)
    if stages is not None:
        stages["xn"] = xn
    r = nystrom(sd, name + ".attn.", xn, stages, return_attn)
    if return_attn:
        return h + r[0], r[1]
    return h + r


def ppeg(sd, h, G):
    """PPEG.forward (TransMIL.py:60-75)."""
    B, _, C = h.shape
    cls, feat = h[:, 0], h[:, 1:]
    img = feat.transpose(1, 2).contiguous().view(B, C, G, G)
    y = (F.conv2d(img, sd["pos_layer.proj.weight"], sd["pos_layer.proj.bias"], 1, 3, groups=C)
         + img
         + F.conv2d(img, sd["pos_layer.proj1.weight"], sd["pos_layer.proj1.bias"], 1, 2, groups=C)
         + F.conv2d(img, sd["pos_layer.proj2.weight"], sd["pos_layer.proj2.bias"], 1, 1, groups=C))
    y = y.flatten(2).transpose(1, 2).contiguous()
    return torch.cat((cls.unsqueeze(1), y), dim=1)


@torch.no_grad()
def ctmil_tokens(sd, x, training=False):
    """Conv stem of CTMIL.forward (code/models/CTMIL.py:137-142; same lines in the commented forward of TransformerMIL.py:157-162):
    tiles [1,N,C,H,H] -> conv1, conv2 (each Conv3x3 no bias -> BatchNorm2d -> GELU -> MaxPool(3,2,1), CTMIL.py:90-101) ->
    `.view(N, h*h, 512)`: a REINTERPRETATION of the [N,512,h,h] memory, not a transpose.  sd: CTMIL state_dict.  Eval mode
    (BatchNorm on its running statistics) unless `training` (batch statistics; the running buffers are not updated here)."""
    h = x.squeeze(0) if x.dim() == 5 else x
    for name in ("conv1", "conv2"):
        h = F.conv2d(h, sd[f"{name}.0.weight"], None, stride=1, padding=1)
        h = F.batch_norm(h, sd[f"{name}.1.running_mean"], sd[f"{name}.1.running_var"], sd[f"{name}.1.weight"], sd[f"{name}.1.bias"],
                         training, 0.1, 1e-5)
        h = F.max_pool2d(F.gelu(h), kernel_size=3, stride=2, padding=1)
    h = h.contiguous()
    return h.view(h.shape[0], h.shape[2] * h.shape[2], h.shape[1])


def ctmil_state(sd):
    """CTMIL / TransformerMIL state_dict -> the TransMIL key names forward() reads (PPEG block pos_layer_0, classifier _fc2;
    CTMIL.py:82,122), without the stem and the sub-modules the Nystrom forward never touches."""
    out = {}
    for k, v in sd.items():
        if k.startswith("pos_layer_0."):
            out["pos_layer." + k[len("pos_layer_0."):]] = v
        elif k.startswith("_fc2."):
            out["_fc." + k[len("_fc2."):]] = v
        elif k.startswith(("cls_token", "layer1.", "layer2.", "norm.")):
            out[k] = v
    return out


def ctmil_forward(sd, x, dtype=torch.float32):
    """CTMIL.forward (CTMIL.py:135-165), eval mode: logits [N, n_classes], one row per tile."""
    sd = {k: (v.detach().to(dtype) if v.is_floating_point() else v) for k, v in sd.items()}
    return forward(ctmil_state(sd), ctmil_tokens(sd, x.to(dtype)), dtype=dtype, tokens=True)


def forward(sd, x, return_attn=False, return_stages=False, dtype=torch.float32, tokens=False):
    """TransMIL.forward (TransMIL.py:167-211), eval mode.

    sd: reference-layout state_dict (any float dtype; cast to `dtype`).
    x:  [B,n,D] / [n,D] / [1,B,n,D] as accepted at TransMIL.py:170-173.
    tokens: x already holds the [B,n,512] tokens of a stem outside this function (CTMIL.py:142): `_fc1` is skipped.
    """
    sd = {k: v.detach().to(dtype) for k, v in sd.items()}
    if x.dim() > 3:
        x = x.squeeze(0)
    elif x.dim() == 2:
        x = x.unsqueeze(0)
    h = x.to(dtype)
    st = {} if return_stages else None
    if not tokens:
        h = fc1(sd, h)
    if st is not None:
        st["fc1"] = h
    h, p = square_pad_cls(sd, h)
    if st is not None:
        st["h0"] = h
    s1 = {} if st is not None else None
    h = trans_layer(sd, "layer1", h, s1)
    if st is not None:
        st["layer1"] = s1
        st["h1"] = h
    h = ppeg(sd, h, p["G"])
    if st is not None:
        st["h2"] = h
    s2 = {} if st is not None else None
    r = trans_layer(sd, "layer2", h, s2, return_attn)
    attn = None
    if return_attn:
        h, attn = r
    else:
        h = r
    if st is not None:
        st["layer2"] = s2
        st["h3"] = h
    hn = F.layer_norm(h, (DIM,), sd["norm.weight"], sd["norm.bias"], 1e-5)[:, 0]
    logits = F.linear(hn, sd["_fc.weight"], sd["_fc.bias"])
    out = [logits]
    if return_attn:
        out.append((attn, p["padding"]))
    if return_stages:
        out.append(st)
    return out[0] if len(out) == 1 else tuple(out)


def y_prob(logits):
    """ModelInterface.step (code/models/model_interface.py:323-324)."""
    return F.softmax(logits, dim=1)
