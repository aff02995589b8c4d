"""TEST INFRASTRUCTURE ONLY -- explicit CPU restatement of the BACKWARD of the TransMIL bag-level path.

The reference has no backward code of its own: gradients come from torch.autograd through
/root/reference/code/models/TransMIL.py:167-211 and nystrom-attention 0.0.11.  This file writes that backward out stage by
stage (the contractions a CUDA backward has to run, in the order it has to run them), in eval mode (Dropout = identity;
the training-mode masks of TransMIL.py:100-133 / to_out enter as an elementwise factor on the same tensors).
Pinning: tests/test_oracle_backward.py checks every function here against torch.autograd of the forward oracle
(oracle/transmil_oracle.py, itself pinned to the reference module) in float64.  No CUDA backward exists yet; nothing outside
tests/ may import this module.
"""
import math

import torch
import torch.nn.functional as F

from . import transmil_oracle as O

HEADS, DIM, DHEAD, LANDMARKS, PINV_ITERS, CONV_K = O.HEADS, O.DIM, O.DHEAD, O.LANDMARKS, O.PINV_ITERS, O.CONV_K


def softmax_backward(P, dP):
    """d/dS of P = softmax(S, -1): dS = P * (dP - sum(dP * P, -1))."""
    return P * (dP - (dP * P).sum(-1, keepdim=True))


def layernorm_backward(x, gamma, dy, eps=1e-5):
    """y = (x - mean) * rstd * gamma + beta over the last axis.  Returns dx, dgamma, dbeta (summed over all rows)."""
    mean = x.mean(-1, keepdim=True)
    xc = x - mean
    rstd = torch.rsqrt((xc * xc).mean(-1, keepdim=True) + eps)
    xhat = xc * rstd
    g = dy * gamma
    dx = rstd * (g - g.mean(-1, keepdim=True) - xhat * (g * xhat).mean(-1, keepdim=True))
    red = tuple(range(x.dim() - 1))
    return dx, (dy * xhat).sum(red), dy.sum(red)


def gelu_backward(x, dy):
    """erf GELU (nn.GELU default): d/dx [x Phi(x)] = Phi(x) + x phi(x)."""
    cdf = 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))
    pdf = torch.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)
    return dy * (cdf + x * pdf)


def pinv_backward(A, dZ, iters=PINV_ITERS):
    """Backward of oracle.pinv (moore_penrose_iter_pinv, global init scalar).  A: [B,h,m,m] (softmax output), dZ: grad of
    the returned Z.  The forward iterates are recomputed (a CUDA backward would keep the 6 iterates or recompute likewise)."""
    absA = A.abs()
    rows, cols = absA.sum(-1), absA.sum(-2)
    r, c = rows.max(), cols.max()
    scale = r * c
    I = torch.eye(A.shape[-1], dtype=A.dtype).expand_as(A)
    Zs = [A.transpose(-1, -2) / scale]
    for _ in range(iters):
        Z = Zs[-1]
        XZ = A @ Z
        Zs.append(0.25 * Z @ (13 * I - XZ @ (15 * I - XZ @ (7 * I - XZ))))
    dA = torch.zeros_like(A)
    T = lambda M: M.transpose(-1, -2)
    for it in reversed(range(iters)):
        Z = Zs[it]
        XZ = A @ Z
        T1 = 7 * I - XZ
        T2 = 15 * I - XZ @ T1
        T3 = 13 * I - XZ @ T2
        dT3 = 0.25 * T(Z) @ dZ
        dZ = 0.25 * dZ @ T(T3)
        dXZ = -dT3 @ T(T2)
        dT2 = -T(XZ) @ dT3
        dXZ = dXZ - dT2 @ T(T1)
        dT1 = -T(XZ) @ dT2
        dXZ = dXZ - dT1
        dA = dA + dXZ @ T(Z)
        dZ = dZ + T(A) @ dXZ
    # Z0 = A^T / (r c); r, c are maxima over every (bag, head) pair: their gradient lands on the maximal row(s) / column(s)
    dA = dA + T(dZ) / scale
    dscale = -(dZ * Zs[0]).sum() / scale
    sgn = torch.sign(A)
    # torch's full-reduction max spreads its gradient evenly over ties (every softmax row sums to 1 up to rounding, so ties
    # among the row sums are the normal case; their share is a per-row constant, which the softmax backward annihilates)
    rmask = (rows == r).to(A.dtype)
    cmask = (cols == c).to(A.dtype)
    dA = dA + (dscale * c / rmask.sum()) * rmask.unsqueeze(-1) * sgn
    dA = dA + (dscale * r / cmask.sum()) * cmask.unsqueeze(-2) * sgn
    return dA


def nystrom_backward(sd, prefix, xn, dy):
    """Backward of oracle.nystrom (NystromAttention.forward, eval).  xn: [B,S,512] layer-normed input, dy: [B,S,512].
    Returns dxn and the gradients of to_qkv.weight, to_out.0.weight, to_out.0.bias, res_conv.weight."""
    B, S, _ = xn.shape
    m = LANDMARKS
    st = {}
    O.nystrom(sd, prefix, xn, st)                      # forward intermediates (q already scaled)
    q, k, v, q_lm, k_lm, A2, Z, W3, merged, P, l = (st[n] for n in ("q", "k", "v", "q_lm", "k_lm", "attn2", "Z", "W3", "merged", "P", "l"))
    Sp = S + P
    T = lambda M: M.transpose(-1, -2)
    A1 = (q @ T(k_lm)).softmax(-1)
    A3 = (q_lm @ T(k)).softmax(-1)
    Wo, Wc, Wqkv = sd[prefix + "to_out.0.weight"], sd[prefix + "res_conv.weight"], sd[prefix + "to_qkv.weight"]
    # to_out and the slice out[:, -S:]: pad rows receive no gradient
    dyp = F.pad(dy, (0, 0, P, 0))
    g_Wo = torch.einsum("bso,bsi->oi", dyp, merged)
    g_bo = dyp.sum((0, 1))
    dO = (dyp @ Wo).reshape(B, Sp, HEADS, DHEAD).permute(0, 2, 1, 3)
    # depthwise 33-tap conv along the token axis (cross-correlation, zero padding 16): out[t] += sum_j w[j] v[t + j - 16]
    w = Wc[:, 0, :, 0]                                                   # [h, 33]
    vp = F.pad(v, (0, 0, CONV_K // 2, CONV_K // 2))
    g_Wc = torch.stack([torch.einsum("bhtd,bhtd->h", dO, vp[:, :, j:j + Sp]) for j in range(CONV_K)], 1).reshape(Wc.shape)
    dOp = F.pad(dO, (0, 0, CONV_K // 2, CONV_K // 2))
    dv = sum(w[None, :, j, None, None] * dOp[:, :, CONV_K - 1 - j:CONV_K - 1 - j + Sp] for j in range(CONV_K))
    # out = A1 (Z W3)
    M = Z @ W3
    dA1 = dO @ T(M)
    dM = T(A1) @ dO
    dZ = dM @ T(W3)
    dW3 = T(Z) @ dM
    dA3 = dW3 @ T(v)
    dv = dv + T(A3) @ dW3
    dA2 = pinv_backward(A2, dZ)
    dS1, dS2, dS3 = softmax_backward(A1, dA1), softmax_backward(A2, dA2), softmax_backward(A3, dA3)
    dq = dS1 @ k_lm
    dk_lm = T(dS1) @ q + T(dS2) @ q_lm
    dq_lm = dS2 @ k_lm + dS3 @ k
    dk = T(dS3) @ q_lm
    # landmark means: every row of a group receives d(lm) / l
    dq = dq + (dq_lm / l).unsqueeze(3).expand(B, HEADS, m, l, DHEAD).reshape(B, HEADS, Sp, DHEAD)
    dk = dk + (dk_lm / l).unsqueeze(3).expand(B, HEADS, m, l, DHEAD).reshape(B, HEADS, Sp, DHEAD)
    dq = dq * (DHEAD ** -0.5)
    dqkv = torch.cat([t.permute(0, 2, 1, 3).reshape(B, Sp, DIM) for t in (dq, dk, dv)], -1)
    xp = F.pad(xn, (0, 0, P, 0))
    g_Wqkv = torch.einsum("bso,bsi->oi", dqkv, xp)
    dxn = (dqkv @ Wqkv)[:, P:]
    return dxn, {prefix + "to_qkv.weight": g_Wqkv, prefix + "to_out.0.weight": g_Wo, prefix + "to_out.0.bias": g_bo,
                 prefix + "res_conv.weight": g_Wc}


def trans_layer_backward(sd, name, h, dy):
    """Backward of oracle.trans_layer: h + attn(norm(h))."""
    gamma = sd[name + ".norm.weight"]
    xn = F.layer_norm(h, (DIM,), gamma, sd[name + ".norm.bias"], 1e-5)
    dxn, grads = nystrom_backward(sd, name + ".attn.", xn, dy)
    dh, g_gamma, g_beta = layernorm_backward(h, gamma, dxn)
    grads[name + ".norm.weight"] = g_gamma
    grads[name + ".norm.bias"] = g_beta
    return dy + dh, grads


def ppeg_backward(sd, h, G, dy):
    """Backward of oracle.ppeg: identity + three depthwise convs on the G x G token image, cls row passed through."""
    B, _, C = h.shape
    img = h[:, 1:].transpose(1, 2).reshape(B, C, G, G)
    dimg = dy[:, 1:].transpose(1, 2).reshape(B, C, G, G)
    dx = dimg.clone()
    grads = {}
    for nm, pad in (("pos_layer.proj", 3), ("pos_layer.proj1", 2), ("pos_layer.proj2", 1)):
        w = sd[nm + ".weight"]
        ksz = w.shape[-1]
        dx = dx + F.conv_transpose2d(dimg, w, None, 1, pad, groups=C)
        ip = F.pad(img, (pad, pad, pad, pad))
        gw = torch.stack([torch.stack([(dimg * ip[:, :, i:i + G, j:j + G]).sum((0, 2, 3)) for j in range(ksz)], -1) for i in range(ksz)], -2)
        grads[nm + ".weight"] = gw.reshape(w.shape)
        grads[nm + ".bias"] = dimg.sum((0, 2, 3))
    dh = torch.cat([dy[:, :1], dx.flatten(2).transpose(1, 2)], 1)
    return dh, grads


def fc1_backward(sd, x, dh, masks=None):
    """Backward of oracle.fc1 (eval; with `masks` = (m0, m4) the training mode of the 768-d stem, see oracle.fc1).  Returns dx
    and the stem's parameter gradients."""
    keys = sorted(k for k in sd if k.startswith("_fc1."))
    idx = sorted({int(k.split(".")[1]) for k in keys})
    lin = lambda a, i: F.linear(a, sd[f"_fc1.{i}.weight"], sd[f"_fc1.{i}.bias"])
    g = {}

    def lin_bwd(a, i, dz):
        g[f"_fc1.{i}.weight"] = torch.einsum("bno,bni->oi", dz, a)
        g[f"_fc1.{i}.bias"] = dz.sum((0, 1))
        return dz @ sd[f"_fc1.{i}.weight"]

    def ln_bwd(a, i, dz):
        da, gg, gb = layernorm_backward(a, sd[f"_fc1.{i}.weight"], dz)
        g[f"_fc1.{i}.weight"], g[f"_fc1.{i}.bias"] = gg, gb
        return da

    if idx == [0]:                                   # Linear GELU
        z0 = lin(x, 0)
        return lin_bwd(x, 0, gelu_backward(z0, dh)), g
    if idx == [0, 2, 3]:                             # Linear GELU LN Linear GELU
        z0 = lin(x, 0); a0 = F.gelu(z0)
        a1 = F.layer_norm(a0, (a0.shape[-1],), sd["_fc1.2.weight"], sd["_fc1.2.bias"], 1e-5)
        z3 = lin(a1, 3)
        d = lin_bwd(a1, 3, gelu_backward(z3, dh))
        d = ln_bwd(a0, 2, d)
        return lin_bwd(x, 0, gelu_backward(z0, d)), g
    if idx == [0, 3, 4, 7]:                          # Linear GELU (Dropout) LN Linear GELU (Dropout) LN
        m0, m4 = masks if masks is not None else (1.0, 1.0)
        z0 = lin(x, 0); a0 = F.gelu(z0) * m0
        a1 = F.layer_norm(a0, (a0.shape[-1],), sd["_fc1.3.weight"], sd["_fc1.3.bias"], 1e-5)
        z4 = lin(a1, 4); a4 = F.gelu(z4) * m4
        d = ln_bwd(a4, 7, dh) * m4
        d = lin_bwd(a1, 4, gelu_backward(z4, d))
        d = ln_bwd(a0, 3, d) * m0
        return lin_bwd(x, 0, gelu_backward(z0, d)), g
    raise RuntimeError(f"unsupported _fc1 layout {idx}")


def backward(sd, x, dlogits, dtype=torch.float64):
    """Gradients of sum(logits * dlogits) with respect to every parameter (reference state_dict keys) and to x, for the eval
    forward of oracle.forward.  Stage order = reverse of TransMIL.forward (TransMIL.py:167-211)."""
    sd = {k: v.detach().to(dtype) for k, v in sd.items()}
    if x.dim() == 2:
        x = x.unsqueeze(0)
    x = x.to(dtype)
    # forward, keeping the layer inputs
    hf = O.fc1(sd, x)
    h0, p = O.square_pad_cls(sd, hf)
    h1 = O.trans_layer(sd, "layer1", h0)
    h2 = O.ppeg(sd, h1, p["G"])
    h3 = O.trans_layer(sd, "layer2", h2)
    grads = {}
    # head: logits = _fc(LN(h3)[:, 0])
    cls = h3[:, 0]
    cn = F.layer_norm(cls, (DIM,), sd["norm.weight"], sd["norm.bias"], 1e-5)
    grads["_fc.weight"] = dlogits.t() @ cn
    grads["_fc.bias"] = dlogits.sum(0)
    dcls, grads["norm.weight"], grads["norm.bias"] = layernorm_backward(cls, sd["norm.weight"], dlogits @ sd["_fc.weight"])
    dh3 = torch.zeros_like(h3)
    dh3[:, 0] = dcls
    dh2, g = trans_layer_backward(sd, "layer2", h2, dh3); grads.update(g)
    dh1, g = ppeg_backward(sd, h1, p["G"], dh2); grads.update(g)
    dh0, g = trans_layer_backward(sd, "layer1", h0, dh1); grads.update(g)
    # cls token + duplicate pad (TransMIL.py:177-186): the first `dup` rows also collect the gradient of their copies
    grads["cls_token"] = dh0[:, :1].sum(0, keepdim=True)
    n, dup = p["n"], p["dup"]
    dhf = dh0[:, 1:1 + n].clone()
    dhf[:, :dup] += dh0[:, 1 + n:]
    dx, g = fc1_backward(sd, x, dhf); grads.update(g)
    return grads, dx
