"""Hand-written training step for the CNN: convolutions as patch gather + tcgen05 GEMM, no autograd, no cuDNN.

Attached to a CUDA ``TorchP2PFLModel`` wrapping :class:`~myfyp_b200.models.cnn.CNN` (conv3x3-ReLU-maxpool2 blocks followed
by linear layers).  Parameters stay the fp32 slab in PyTorch's own layouts - that is what is aggregated and exchanged -
and are only *read* here (bf16 working copies per step, no permutation: the patch index order is (c, kh, kw)).

Activations are NHWC bf16, i.e. the ``[pixels, channels]`` matrices the GEMM reads and writes.  Per conv block:

    forward   col = im2col(P_prev)            Y = relu(col W^T + b)  (GEMM epilogue)        P = maxpool2(Y), argmax kept
    backward  dZ  = unpool(dP) * (Y > 0)      db = colsum(dZ)
              dW  = dZ^T col                  (MN-major operands, split-K over pixels, fp32 atomics into the gradient slab)
              dcol = dZ W  (W MN-major)       dP_prev = col2im(dcol)

The first linear layer after the feature map is a convolution whose kernel covers the whole map, which reproduces
``nn.Flatten``'s NCHW feature order on NHWC data.  Reference path: the same ``nn.Module`` through torch autograd / cuDNN.
"""

from __future__ import annotations

import ctypes
from typing import Dict, List

import torch
from torch import nn

from myfyp_b200 import ops


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
        self.w16 = {n: torch.zeros(p[0].shape[0], _pad8(p[0].shape[1]), dtype=torch.bfloat16, device=dev) for n, p in self.params.items()}
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
                    "col": z(m, _pad8(s.kdim)), "y": z(m, s.cout), "p": z(B * (s.oh // 2) * (s.ow // 2), s.cout), "idx": z(B * (s.oh // 2) * (s.ow // 2), s.cout, dtype=torch.uint8),
                    "dz": z(m, s.cout), "dcol": z(m, _pad8(s.kdim)), "dp": z(B * (s.oh // 2) * (s.ow // 2), s.cout),
                })
            f = self.fc1
            buf.update(
                col_fc1=z(B, _pad8(f.kdim)), h1=z(B, _pad8(f.cout)), dh1=z(B, _pad8(f.cout)), dcol_fc1=z(B, _pad8(f.kdim)),
                logits=z(B, self.fc2_out, dtype=torch.float32), dlogits=z(B, self.fc2_out, dtype=torch.float32), dz2=z(B, _pad8(self.fc2_out)),
            )
            self._bufs[B] = buf
        return self._bufs[B]

    def refresh_weights(self) -> None:
        for n, (w, _, _, _) in self.params.items():
            ops.cast_transpose(w, w.shape[0], w.shape[1], dst=self.w16[n])

    # ------------------------------------------------------------------ forward
    def _forward(self, x: torch.Tensor, buf: dict) -> int:
        B = x.shape[0]
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        prev = x
        for li, s in enumerate(self.convs):
            cb = buf["conv"][li]
            ops.im2col(prev, cb["col"], B, s.h, s.w, s.cin, s.k, s.k, s.stride, s.pad, nchw_f32=(li == 0))
            w, b, _, _ = self.params[s.name]
            ops.gemm_bf16(cb["col"], self.w16[s.name], B * s.oh * s.ow, s.cout, s.kdim, out_bf16=cb["y"], bias=b, relu=True)
            ops.maxpool2_fwd(cb["y"], cb["p"], cb["idx"], B, s.oh, s.ow, s.cout)
            prev = cb["p"]
        f = self.fc1
        ops.im2col(prev, buf["col_fc1"], B, f.h, f.w, f.cin, f.k, f.k, 1, 0)
        ops.gemm_bf16(buf["col_fc1"], self.w16[f.name], B, f.cout, f.kdim, out_bf16=buf["h1"], bias=self.params[f.name][1], relu=True)
        ops.gemm_bf16(buf["h1"], self.w16[self.fc2_name], B, self.fc2_out, self.fc2_in, out_f32=buf["logits"], bias=self.params[self.fc2_name][1])
        return B

    # ------------------------------------------------------------------ training step
    def train_step(self, x: torch.Tensor, y: torch.Tensor, loss_out: torch.Tensor) -> None:
        self.refresh_weights()
        B = x.shape[0]
        buf = self._buffers(B)
        self._forward(x, buf)
        native = ops.native
        lib = native.load()
        loss_out.zero_()
        with native.device_guard(x.device):
            native.check(
                lib.b2_xent_fwd_bwd(native.ptr(buf["logits"]), native.ptr(y), native.ptr(loss_out), native.ptr(buf["dlogits"]), native.ptr(None), B, self.fc2_out, ctypes.c_float(1.0 / B), native.stream_ptr(x.device)),
                "b2_xent_fwd_bwd",
            )
        native.count_launch()
        # ---- fc2
        _, _, gw2, gb2 = self.params[self.fc2_name]
        ops.cast_transpose(buf["dlogits"], B, self.fc2_out, dst=buf["dz2"], colsum=gb2)
        ops.gemm_bf16(buf["dz2"], buf["h1"], self.fc2_out, self.fc2_in, B, out_f32=gw2, a_mn=True, b_mn=True)
        # ---- fc1 (dgrad of fc2 with the ReLU mask of h1; its column sums are fc1's bias gradient)
        f = self.fc1
        _, _, gw1, gb1 = self.params[f.name]
        ops.gemm_bf16(buf["dz2"], self.w16[self.fc2_name], B, self.fc2_in, self.fc2_out, out_bf16=buf["dh1"], mask=buf["h1"], colsum=gb1, b_mn=True)
        ops.gemm_bf16(buf["dh1"], buf["col_fc1"], f.cout, f.kdim, B, out_f32=gw1, a_mn=True, b_mn=True)
        ops.gemm_bf16(buf["dh1"], self.w16[f.name], B, f.kdim, f.cout, out_bf16=buf["dcol_fc1"], b_mn=True)
        last = buf["conv"][-1]
        ops.col2im(buf["dcol_fc1"], last["dp"], B, f.h, f.w, f.cin, f.k, f.k, 1, 0)
        # ---- conv blocks, last to first
        for li in range(len(self.convs) - 1, -1, -1):
            s, cb = self.convs[li], buf["conv"][li]
            m = B * s.oh * s.ow
            _, _, gw, gb = self.params[s.name]
            ops.maxpool2_relu_bwd(cb["dp"], cb["idx"], cb["y"], cb["dz"], B, s.oh, s.ow, s.cout)
            ops.colsum_bf16(cb["dz"], m, s.cout, gb)
            ops.gemm_bf16(cb["dz"], cb["col"], s.cout, s.kdim, m, out_f32=gw, a_mn=True, b_mn=True, split_k=max(1, min(64, m // 1024)))
            if li > 0:
                ops.gemm_bf16(cb["dz"], self.w16[s.name], m, s.kdim, s.cout, out_bf16=cb["dcol"], b_mn=True)
                ops.col2im(cb["dcol"], buf["conv"][li - 1]["dp"], B, s.h, s.w, s.cin, s.k, s.k, s.stride, s.pad)

    @torch.no_grad()
    def eval_step(self, x: torch.Tensor, y: torch.Tensor, acc: torch.Tensor) -> None:
        self.refresh_weights()
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
