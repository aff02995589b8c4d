"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the third-party package
`nystrom-attention` 0.0.11 (lucidrains), which the reference imports at
code/models/TransMIL.py:5 but does not vendor.

This is NOT reference code and NOT product code.  It restates the published
algorithm of that package (spec: SURVEY.md section 8c) so that the reference's
own `code/models/TransMIL.py` can be imported on a box where the package is
absent, and so that `oracle/transmil_oracle.py` has an attention to call.

PARITY UNPINNED at this boundary: the reference holds no test, golden vector
or fixture for NystromAttention, and the package itself cannot be installed
offline.  What the reference tree does pin -- constructor kwargs
(TransMIL.py:26-34), the (out, attn) tuple (TransMIL.py:47), the dense attn
shape [B,8,Sp,Sp] and the FRONT zero-pad indexing (TransMIL.py:200,221;
code/visualize_mil.py:581) -- is honoured here and checked in
tests/test_oracle.py.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference leg may import this module.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["NystromAttention", "moore_penrose_iter_pinv"]


def moore_penrose_iter_pinv(x, iters=6):
    """Iterative pseudo-inverse, nystrom-attention 0.0.11.

    z0 = x^T / (max(colsum) * max(rowsum)), both maxima over ALL leading dims
    (batch and heads), then `iters` steps of
        xz = x z ;  z = 1/4 z (13 I - xz (15 I - xz (7 I - xz))).
    """
    abs_x = torch.abs(x)
    col = abs_x.sum(dim=-1)
    row = abs_x.sum(dim=-2)
    z = x.transpose(-1, -2) / (torch.max(col) * torch.max(row))
    eye = torch.eye(x.shape[-1], device=x.device, dtype=x.dtype).unsqueeze(0)
    for _ in range(iters):
        xz = x @ z
        z = 0.25 * z @ (13 * eye - (xz @ (15 * eye - (xz @ (7 * eye - xz)))))
    return z


class NystromAttention(nn.Module):
    """Constructor signature as used at code/models/TransMIL.py:26-34."""

    def __init__(self, dim, dim_head=64, heads=8, num_landmarks=256,
                 pinv_iterations=6, residual=True, residual_conv_kernel=33,
                 eps=1e-8, dropout=0.0):
        super().__init__()
        self.eps = eps
        inner_dim = heads * dim_head
        self.num_landmarks = num_landmarks
        self.pinv_iterations = pinv_iterations
        self.heads = heads
        self.scale = dim_head ** -0.5
        self.to_qkv = nn.Linear(dim, inner_dim * 3, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner_dim, dim), nn.Dropout(dropout))
        self.residual = residual
        if residual:
            kernel_size = residual_conv_kernel
            padding = residual_conv_kernel // 2
            self.res_conv = nn.Conv2d(heads, heads, (kernel_size, 1),
                                      padding=(padding, 0), groups=heads, bias=False)

    def forward(self, x, mask=None, return_attn=False):
        b, n, _ = x.shape
        h, m, iters = self.heads, self.num_landmarks, self.pinv_iterations
        # FRONT zero-pad so the token count divides the landmark count
        remainder = n % m
        if remainder > 0:
            padding = m - remainder
            x = F.pad(x, (0, 0, padding, 0), value=0)
            if mask is not None:
                mask = F.pad(mask, (padding, 0), value=False)
        assert mask is None, "TransMIL never passes a mask (TransMIL.py:47)"
        q, k, v = self.to_qkv(x).chunk(3, dim=-1)
        q, k, v = (t.reshape(b, -1, h, t.shape[-1] // h).permute(0, 2, 1, 3) for t in (q, k, v))
        q = q * self.scale
        # landmarks: mean over l consecutive PADDED positions, l = ceil(n / m)
        l = math.ceil(n / m)
        q_landmarks = q.reshape(b, h, m, l, -1).sum(dim=3) / l
        k_landmarks = k.reshape(b, h, m, l, -1).sum(dim=3) / l
        sim1 = torch.einsum("bhid,bhjd->bhij", q, k_landmarks)
        sim2 = torch.einsum("bhid,bhjd->bhij", q_landmarks, k_landmarks)
        sim3 = torch.einsum("bhid,bhjd->bhij", q_landmarks, k)
        attn1, attn2, attn3 = (t.softmax(dim=-1) for t in (sim1, sim2, sim3))
        attn2_inv = moore_penrose_iter_pinv(attn2, iters)
        out = (attn1 @ attn2_inv) @ (attn3 @ v)
        if self.residual:
            out = out + self.res_conv(v)
        out = out.permute(0, 2, 1, 3).reshape(b, out.shape[2], -1)
        out = self.to_out(out)
        out = out[:, -n:]
        if return_attn:
            attn = attn1 @ attn2_inv @ attn3
            return out, attn
        return out
