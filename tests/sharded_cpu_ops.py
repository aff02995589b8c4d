"""TEST INFRASTRUCTURE ONLY -- torch CPU restatement of the per-kernel operations the token-sharded host logic
(transmil-deepgraft_b200/sharded.py) calls, built from the oracle.  It lets the world_size>1 gloo tests drive the real
shard plan, halo exchanges and collectives without a GPU.  Never imported by the package."""
import torch
import torch.nn.functional as F

from oracle import transmil_oracle as O

LM, HEADS, DHEAD, DIM = 256, 8, 64, 512


class OracleOps:
    def __init__(self, sd, n_classes):
        self.sd = {k: v.detach().float() for k, v in sd.items()}
        self.n_classes = n_classes

    def f32(self, *shape):
        return torch.zeros(*shape, dtype=torch.float32)

    def comm_buffer(self, *shape, dtype=torch.float32):
        return torch.zeros(*shape, dtype=dtype)

    def side_stream(self):
        import contextlib
        return contextlib.nullcontext()

    def join_side_stream(self):
        pass

    def to_device(self, x):
        return x.float().contiguous()

    def fc1_rows(self, x_rows, out):
        out.copy_(O.fc1(self.sd, x_rows[None])[0])

    def cls_row(self, out_row):
        out_row.copy_(self.sd["cls_token"].view(DIM))

    def ln_pad(self, h, name, S_local, P_local):
        xn = F.layer_norm(h[:S_local], (DIM,), self.sd[name + ".norm.weight"], self.sd[name + ".norm.bias"], 1e-5)
        return torch.cat([torch.zeros(P_local, DIM), xn], 0)

    def qkv(self, xn, name):
        qkv = F.linear(xn, self.sd[name + ".attn.to_qkv.weight"])
        q, k, v = (t.reshape(-1, HEADS, DHEAD).permute(1, 0, 2).contiguous() for t in qkv.chunk(3, -1))
        return q * DHEAD ** -0.5, k, v

    def v_edges(self, v):
        return v[:, :16].contiguous(), v[:, -16:].contiguous()

    def landmark_sums(self, q, k, row0, l, sums):
        g = (row0 + torch.arange(q.shape[1])) // l
        sums.zero_()
        sums[0].index_add_(1, g, q)
        sums[1].index_add_(1, g, k)

    def landmark_finish(self, sums, l):
        return dict(q_lm=sums[0] / l, k_lm=sums[1] / l)

    def pinv(self, lm):
        attn2 = (lm["q_lm"] @ lm["k_lm"].transpose(-1, -2)).softmax(-1)
        return O.pinv(attn2[None])[0]

    def partial_floats(self):
        return HEADS * LM * (DHEAD + 2)

    def attn3v_partial(self, lm, k, v, partial):
        s = lm["q_lm"] @ k.transpose(-1, -2)
        m = s.max(-1).values
        e = (s - m[..., None]).exp()
        partial.copy_(torch.cat([m.reshape(-1), e.sum(-1).reshape(-1), (e @ v).reshape(-1)]))

    def attn3v_merge(self, gathered, world):
        nb = HEADS * LM
        m = gathered[:, :nb]
        s = gathered[:, nb:2 * nb]
        num = gathered[:, 2 * nb:].reshape(world, nb, DHEAD)
        M = m.max(0).values
        w = (m - M).exp()
        return ((w[..., None] * num).sum(0) / (w * s).sum(0)[:, None]).reshape(HEADS, LM, DHEAD)

    def zw(self, Z, W3):
        return Z @ W3

    def _conv(self, v, name):
        w = self.sd[name + ".attn.res_conv.weight"]
        return F.conv2d(v[None], w, None, padding=(0, 0), groups=HEADS)[0]

    def nystrom_out(self, q, lm, M, v, name):
        attn1 = (q @ lm["k_lm"].transpose(-1, -2)).softmax(-1)
        z = torch.zeros(HEADS, 16, DHEAD)
        out = attn1 @ M + self._conv(torch.cat([z, v, z], 1), name)       # zero padding at the SHARD edges
        return out.permute(1, 0, 2).reshape(-1, DIM).contiguous()

    def conv_halo_fix(self, merged, v_prev, v_next, name):
        z = torch.zeros(HEADS, 32, DHEAD)
        if v_prev is not None:
            merged[:16] += self._conv(torch.cat([v_prev, z], 1), name).permute(1, 0, 2).reshape(16, DIM)
        if v_next is not None:
            merged[-16:] += self._conv(torch.cat([z, v_next], 1), name).permute(1, 0, 2).reshape(16, DIM)

    def to_out(self, merged, name, h_in, S_local, P_local, h_out):
        y = F.linear(merged[P_local:], self.sd[name + ".attn.to_out.0.weight"], self.sd[name + ".attn.to_out.0.bias"])
        h_out.copy_(h_in[:S_local] + y)

    def ppeg_rows(self, in_rows, y_in0, y_in1, G, tok0, tok1, out):
        ny = y_in1 - y_in0
        img = in_rows[:ny * G].reshape(ny, G, DIM).permute(2, 0, 1)[None]
        sd = self.sd
        y = (F.conv2d(img, sd["pos_layer.proj.weight"], sd["pos_layer.proj.bias"], 1, 3, groups=DIM) + img
             + F.conv2d(img, sd["pos_layer.proj1.weight"], sd["pos_layer.proj1.bias"], 1, 2, groups=DIM)
             + F.conv2d(img, sd["pos_layer.proj2.weight"], sd["pos_layer.proj2.bias"], 1, 1, groups=DIM))
        y = y[0].permute(1, 2, 0).reshape(ny * G, DIM)
        out.copy_(y[tok0 - y_in0 * G: tok1 - y_in0 * G])

    def head(self, h, S_local):
        hn = F.layer_norm(h[:1], (DIM,), self.sd["norm.weight"], self.sd["norm.bias"], 1e-5)
        return F.linear(hn, self.sd["_fc.weight"], self.sd["_fc.bias"])

    def empty_logits(self):
        return torch.zeros(1, self.n_classes)
