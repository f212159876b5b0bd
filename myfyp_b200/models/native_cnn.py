"""Hand-written training step for the CNN: convolutions as patch gather + tcgen05 GEMM, no autograd, no cuDNN.

Attached to a CUDA ``TorchP2PFLModel`` wrapping :class:`~myfyp_b200.models.cnn.CNN` (conv3x3-ReLU-maxpool2 blocks followed
by linear layers).  Parameters stay the fp32 slab in PyTorch's own layouts - that is what is aggregated and exchanged -
and are only *read* here: one launch per step repacks every conv / fc1 weight into bf16 ``[cout, (kh, kw, cin)]``, the patch
order in which a tap's channels are contiguous (16-byte vector gathers), and the weight-gradient GEMMs scatter back.

Activations are NHWC bf16, i.e. the ``[pixels, channels]`` matrices the GEMM reads and writes.  Per conv block:

    forward   Y = relu(col W^T + b)  (GEMM epilogue)       col_next = im2col(maxpool2(Y))  (ONE kernel, argmax kept, P never stored)
    backward  dZ = unpool(col2im(dcol_next)) * (Y > 0), db = colsum(dZ)                     (ONE kernel)
              dW  = dZ^T col                  (MN-major operands, split-K over pixels, fp32 atomics into the gradient slab)
              dcol = dZ W  (W MN-major)

The first linear layer after the feature map is a convolution whose kernel covers the whole map; the same weight repack
maps ``nn.Flatten``'s NCHW feature order onto the NHWC data.  Reference path: the same ``nn.Module`` through torch autograd / cuDNN.
"""

from __future__ import annotations

from typing import Dict, List

import torch
from torch import nn

from myfyp_b200 import ops
from myfyp_b200.models._branch import WgradBranch


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


class _ConvSpec:
    def __init__(self, name: str, cin: int, cout: int, k: int, stride: int, pad: int, h: int, w: int, pool: bool) -> None:
        self.name, self.cin, self.cout, self.k, self.stride, self.pad, self.h, self.w, self.pool = name, cin, cout, k, stride, pad, h, w, pool
        self.oh = (h + 2 * pad - k) // stride + 1
        self.ow = (w + 2 * pad - k) // stride + 1
        self.kdim = cin * k * k


class NativeCNN:
    def __init__(self, model, image_size: int = 32) -> None:
        self.model = model
        module = model.get_model()
        dev = model.device
        lay = model.layout
        self.convs: List[_ConvSpec] = []
        h = w = image_size
        feats = list(module.features)
        i = 0
        while i < len(feats):
            m = feats[i]
            if isinstance(m, nn.Conv2d):
                pool = i + 2 < len(feats) and isinstance(feats[i + 2], nn.MaxPool2d)
                if m.kernel_size[0] != m.kernel_size[1] or not isinstance(feats[i + 1], nn.ReLU) or not pool:
                    raise ValueError("NativeCNN supports conv(k x k) - ReLU - MaxPool2d(2) blocks")
                spec = _ConvSpec(f"features.{i}", m.in_channels, m.out_channels, m.kernel_size[0], m.stride[0], m.padding[0], h, w, True)
                self.convs.append(spec)
                h, w = spec.oh // 2, spec.ow // 2
                i += 3
            else:
                i += 1
        linears = [(j, m) for j, m in enumerate(module.classifier) if isinstance(m, nn.Linear)]
        if len(linears) != 2:
            raise ValueError("NativeCNN expects Flatten - Linear - ReLU - Linear")
        cfeat = self.convs[-1].cout
        j1, l1 = linears[0]
        # fc1 == a convolution whose kernel is the whole (h x w) map: keeps nn.Flatten's (c, h, w) feature order
        self.fc1 = _ConvSpec(f"classifier.{j1}", cfeat, l1.out_features, h, 1, 0, h, w, False)
        assert self.fc1.kdim == l1.in_features, "classifier input does not match the feature map"
        j2, l2 = linears[1]
        self.fc2_name, self.fc2_in, self.fc2_out = f"classifier.{j2}", l2.in_features, l2.out_features
        self.in_channels = self.convs[0].cin

        def views(name: str, rows: int, cols: int):
            ew, eb = lay.entries[name + ".weight"], lay.entries[name + ".bias"]
            return (
                model.slab[ew.offset : ew.offset + ew.numel].view(rows, cols), model.slab[eb.offset : eb.offset + eb.numel],
                model.grad_slab[ew.offset : ew.offset + ew.numel].view(rows, cols), model.grad_slab[eb.offset : eb.offset + eb.numel],
            )

        self.params = {s.name: views(s.name, s.cout, s.kdim) for s in self.convs + [self.fc1]}
        self.params[self.fc2_name] = views(self.fc2_name, self.fc2_out, self.fc2_in)
        # bf16 mirror of the whole slab, kept current by the fused optimizer (fc2 reads it in place); conv / fc1 weights are
        # repacked per step into [cout, (kh, kw, cin)] rows (pitch padded to 8 elements for TMA)
        self.p16 = torch.zeros(lay.total, dtype=torch.bfloat16, device=dev)
        specs = self.convs + [self.fc1]
        self.p16r = torch.zeros(sum(sp.cout * _pad8(sp.kdim) for sp in specs), dtype=torch.bfloat16, device=dev)
        self.w16, table, roff = {}, [], 0
        for sp in specs:
            ew, ld = lay.entries[sp.name + ".weight"], _pad8(sp.kdim)
            self.w16[sp.name] = self.p16r[roff : roff + sp.cout * ld].view(sp.cout, ld)
            table += [ew.offset, roff, sp.cout, sp.cin, sp.k * sp.k, ld]
            roff += sp.cout * ld
        self._repack_table = torch.tensor(table, dtype=torch.int64, device=dev)
        self._repack_max = max(sp.cout * sp.kdim for sp in specs)
        e2 = lay.entries[self.fc2_name + ".weight"]
        self.w16[self.fc2_name] = self.p16[e2.offset : e2.offset + e2.numel].view(self.fc2_out, self.fc2_in)
        if self.fc2_in % 8:
            raise ValueError("NativeCNN: the hidden width must be a multiple of 8")
        self.branch = WgradBranch(dev)  # weight-gradient GEMMs run beside the dgrad chain
        self._bufs: Dict[int, dict] = {}

    # ------------------------------------------------------------------ buffers
    def _buffers(self, B: int) -> dict:
        if B not in self._bufs:
            dev, bf = self.model.device, torch.bfloat16
            z = lambda *shape, dtype=bf: torch.zeros(*shape, dtype=dtype, device=dev)  # noqa: E731
            buf: dict = {"conv": []}
            for s in self.convs:
                m = B * s.oh * s.ow
                buf["conv"].append({
                    "col": z(m, _pad8(s.kdim)), "y": z(m, s.cout), "idx": z(B * (s.oh // 2) * (s.ow // 2), s.cout, dtype=torch.uint8),
                    "dz": z(m, s.cout), "dcol": z(m, _pad8(s.kdim)),
                })
            f = self.fc1
            buf.update(
                col_fc1=z(B, _pad8(f.kdim)), h1=z(B, _pad8(f.cout)), dh1=z(B, _pad8(f.cout)), dcol_fc1=z(B, _pad8(f.kdim)),
                logits=z(B, self.fc2_out, dtype=torch.float32), dz2=z(B, _pad8(self.fc2_out)),
            )
            self._bufs[B] = buf
        return self._bufs[B]

    def refresh_weights(self) -> None:
        """fp32 slab -> bf16 mirror, once per fit / evaluation (the optimizer updates the mirror after every step)."""
        ops.cast_slab(self.model.slab, self.p16)
        self._refresh_padded()

    def _refresh_padded(self) -> None:
        ops.repack_conv_weights(self.model.slab, self.p16r, self._repack_table, len(self.convs) + 1, self._repack_max)

    def _next_geom(self, li: int):
        """Geometry of the layer that consumes conv block ``li``'s pooled output (the next conv, or fc1 as a full-map conv)."""
        nxt = self.convs[li + 1] if li + 1 < len(self.convs) else self.fc1
        return nxt, (nxt.h, nxt.w, nxt.cin, nxt.k, nxt.k, nxt.stride, nxt.pad)

    # ------------------------------------------------------------------ forward
    def _forward(self, x: torch.Tensor, buf: dict) -> int:
        B = x.shape[0]
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        s0 = self.convs[0]
        ops.im2col(x, buf["conv"][0]["col"], B, s0.h, s0.w, s0.cin, s0.k, s0.k, s0.stride, s0.pad, nchw_f32=True)
        for li, s in enumerate(self.convs):
            cb = buf["conv"][li]
            ops.gemm_bf16(cb["col"], self.w16[s.name], B * s.oh * s.ow, s.cout, s.kdim, out_bf16=cb["y"], bias=self.params[s.name][1], relu=True)
            # max-pool + the next layer's patch gather in one pass (the pooled activation is never stored)
            nxt, (h, w, c, kh, kw, st, pd) = self._next_geom(li)
            dst = buf["conv"][li + 1]["col"] if li + 1 < len(self.convs) else buf["col_fc1"]
            ops.pool_im2col(cb["y"], cb["idx"], dst, B, h, w, c, kh, kw, st, pd)
        f = self.fc1
        ops.gemm_bf16(buf["col_fc1"], self.w16[f.name], B, f.cout, f.kdim, out_bf16=buf["h1"], bias=self.params[f.name][1], relu=True)
        ops.gemm_bf16(buf["h1"], self.w16[self.fc2_name], B, self.fc2_out, self.fc2_in, out_f32=buf["logits"], bias=self.params[self.fc2_name][1])
        return B

    # ------------------------------------------------------------------ training step
    def train_step(self, x: torch.Tensor, y: torch.Tensor, loss_out: torch.Tensor) -> None:
        self._refresh_padded()
        B = x.shape[0]
        buf = self._buffers(B)
        self._forward(x, buf)
        # ---- loss + dlogits (bf16) + fc2 bias gradient: one launch
        _, _, gw2, gb2 = self.params[self.fc2_name]
        ops.xent_train(buf["logits"], y, loss_out, buf["dz2"], gb2, B, self.fc2_out)
        self.branch.run(lambda: ops.gemm_bf16(buf["dz2"], buf["h1"], self.fc2_out, self.fc2_in, B, out_f32=gw2, a_mn=True, b_mn=True))
        # ---- fc1 (dgrad of fc2 with the ReLU mask of h1; its column sums are fc1's bias gradient)
        f = self.fc1
        _, _, gw1, gb1 = self.params[f.name]
        ops.gemm_bf16(buf["dz2"], self.w16[self.fc2_name], B, self.fc2_in, self.fc2_out, out_bf16=buf["dh1"], mask=buf["h1"], colsum=gb1, b_mn=True)
        self.branch.run(lambda: ops.gemm_bf16(buf["dh1"], buf["col_fc1"], f.cout, f.kdim, B, out_f32=gw1, a_mn=True, b_mn=True, perm=(f.cin, f.k * f.k)))
        ops.gemm_bf16(buf["dh1"], self.w16[f.name], B, f.kdim, f.cout, out_bf16=buf["dcol_fc1"], b_mn=True)
        # ---- conv blocks, last to first
        for li in range(len(self.convs) - 1, -1, -1):
            s, cb = self.convs[li], buf["conv"][li]
            m = B * s.oh * s.ow
            _, _, gw, gb = self.params[s.name]
            nxt, (h, w, c, kh, kw, st, pd) = self._next_geom(li)
            next_dcol = buf["conv"][li + 1]["dcol"] if li + 1 < len(self.convs) else buf["dcol_fc1"]
            # col2im of the next layer's dcol + unpool + ReLU mask + bias gradient: one launch
            ops.unpool_col2im(next_dcol, cb["idx"], cb["y"], cb["dz"], gb, B, h, w, c, kh, kw, st, pd)
            self.branch.run(lambda s=s, cb=cb, m=m, gw=gw: ops.gemm_bf16(cb["dz"], cb["col"], s.cout, s.kdim, m, out_f32=gw, a_mn=True, b_mn=True, split_k=max(1, min(64, m // 1024)), perm=(s.cin, s.k * s.k)))
            if li > 0:
                ops.gemm_bf16(cb["dz"], self.w16[s.name], m, s.kdim, s.cout, out_bf16=cb["dcol"], b_mn=True)
        self.branch.join()

    @torch.no_grad()
    def eval_step(self, x: torch.Tensor, y: torch.Tensor, acc: torch.Tensor) -> None:
        """``acc[0] += sum NLL``, ``acc[1] += #correct``  (the caller refreshed the bf16 mirror)."""
        buf = self._buffers(x.shape[0])
        B = self._forward(x, buf)
        ops.cross_entropy_eval(buf["logits"][:B], y, acc)

    def logits(self, x: torch.Tensor) -> torch.Tensor:
        self.refresh_weights()
        buf = self._buffers(x.shape[0])
        B = self._forward(x, buf)
        return buf["logits"][:B].clone()


def attach_native_cnn(model, image_size: int = 32) -> NativeCNN:
    if model.device.type != "cuda":
        raise RuntimeError("the native CNN path needs a CUDA device")
    nc = NativeCNN(model, image_size)
    module = model.get_model()
    module.native_train_step = nc.train_step
    module.native_eval = nc.eval_step
    module.native = nc
    return nc
