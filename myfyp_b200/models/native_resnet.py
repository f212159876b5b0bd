"""Hand-written training step for ResNet-18 (CIFAR variant): no autograd, no cuDNN.

Every convolution is a patch gather + the tcgen05 GEMM (``csrc/gemm.cu``); the GEMM epilogue also accumulates the batch-norm
statistics (column sums and sums of squares of the fp32 accumulators), so batch normalisation costs one elementwise kernel
forward (normalise + scale/shift + residual add + ReLU fused) and two backward (reduce, apply).  Activations are NHWC bf16;
parameters stay the fp32 slab in PyTorch's layouts (bf16 mirror written by the fused optimizer); BatchNorm running statistics
live in the slab too and are updated in place with PyTorch semantics, so what the nodes exchange and aggregate is exactly the
``state_dict`` of the ``nn.Module``.

    conv + bn (+res) (+relu)   col = im2col(x);  z = col W^T  [+ stats];  y = act(bn(z) + res)
    backward                   g = dy * (y > 0);  dz = bn'(g);  dW = dz^T col;  dcol = dz W;  dx = col2im(dcol) (+ second consumer)

Reference path: the same ``nn.Module`` through torch autograd / cuDNN (ResNet-18 is not in the reference itself; BASELINE.json
configs 3-4 name it).
"""

from __future__ import annotations

from typing import Dict, List

import torch

from myfyp_b200 import ops
from myfyp_b200.models._branch import WgradBranch


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


class _ConvBN:
    def __init__(self, conv: str, bn: str, cin: int, cout: int, k: int, stride: int, pad: int, h: int, w: int, relu: bool) -> None:
        self.conv, self.bn, self.cin, self.cout, self.k, self.stride, self.pad, self.h, self.w, self.relu = conv, bn, cin, cout, k, stride, pad, h, w, relu
        self.oh = (h + 2 * pad - k) // stride + 1
        self.ow = (w + 2 * pad - k) // stride + 1
        self.kdim = cin * k * k
        # stride-1 "same" convolutions with 64-channel multiples run as implicit GEMMs (4-D TMA boxes, no patch matrix)
        self.implicit = ops.conv_gemm_supported(h, w, cin, cout, k, stride, pad)


class NativeResNet:
    def __init__(self, model, image_size: int = 32) -> None:
        self.model = model
        m = model.get_model()
        dev, lay = model.device, model.layout
        self.layers: List[_ConvBN] = []
        self.blocks: List[dict] = []
        h = w = image_size
        stem = _ConvBN("conv1", "bn1", m.conv1.in_channels, m.conv1.out_channels, 3, 1, 1, h, w, True)
        self.layers.append(stem)
        self.stem = stem
        for bi, blk in enumerate(m.layers):
            s = blk.conv1.stride[0]
            c1 = _ConvBN(f"layers.{bi}.conv1", f"layers.{bi}.bn1", blk.conv1.in_channels, blk.conv1.out_channels, 3, s, 1, h, w, True)
            c2 = _ConvBN(f"layers.{bi}.conv2", f"layers.{bi}.bn2", blk.conv2.in_channels, blk.conv2.out_channels, 3, 1, 1, c1.oh, c1.ow, True)
            sc = None
            if blk.shortcut is not None:
                sc = _ConvBN(f"layers.{bi}.shortcut.0", f"layers.{bi}.shortcut.1", blk.conv1.in_channels, blk.conv2.out_channels, 1, s, 0, h, w, False)
            self.blocks.append({"c1": c1, "c2": c2, "sc": sc})
            self.layers += [c1, c2] + ([sc] if sc is not None else [])
            h, w = c1.oh, c1.ow
        self.final_hw, self.final_c = h * w, self.blocks[-1]["c2"].cout
        self.fc_in, self.fc_out = m.fc.in_features, m.fc.out_features

        def pv(name: str, shape=None):
            e = lay.entries[name]
            v = model.slab[e.offset : e.offset + e.numel]
            return v.view(shape) if shape is not None else v

        def gv(name: str, shape=None):
            e = lay.entries[name]
            v = model.grad_slab[e.offset : e.offset + e.numel]
            return v.view(shape) if shape is not None else v

        self.p16 = torch.zeros(lay.total, dtype=torch.bfloat16, device=dev)
        self.par: Dict[str, dict] = {}
        # every convolution reads its weights as bf16 [cout, (kh, kw, cin)] (row pitch padded to 8): one repack launch per step
        self.p16r = torch.zeros(sum(L.cout * _pad8(L.kdim) for L in self.layers), dtype=torch.bfloat16, device=dev)
        table, roff = [], 0
        self._repack_max = max(L.cout * L.kdim for L in self.layers)
        for L in self.layers:
            e = lay.entries[L.conv + ".weight"]
            ld = _pad8(L.kdim)
            self.par[L.conv] = {
                "w": pv(L.conv + ".weight", (L.cout, L.kdim)), "gw": gv(L.conv + ".weight", (L.cout, L.kdim)),
                "gamma": pv(L.bn + ".weight"), "beta": pv(L.bn + ".bias"), "dgamma": gv(L.bn + ".weight"), "dbeta": gv(L.bn + ".bias"),
                "rm": pv(L.bn + ".running_mean"), "rv": pv(L.bn + ".running_var"), "nb": pv(L.bn + ".num_batches_tracked"),
                "w16": self.p16r[roff : roff + L.cout * ld].view(L.cout, ld),
            }
            table += [e.offset, roff, L.cout, L.cin, L.k * L.k, ld]
            roff += L.cout * ld
        self._repack_table = torch.tensor(table, dtype=torch.int64, device=dev) if table else None
        self._n_repack = len(table) // 6
        efc = lay.entries["fc.weight"]
        self.fc = {"w": pv("fc.weight", (self.fc_out, self.fc_in)), "b": pv("fc.bias"), "gw": gv("fc.weight", (self.fc_out, self.fc_in)), "gb": gv("fc.bias"),
                   "w16": self.p16[efc.offset : efc.offset + efc.numel].view(self.fc_out, self.fc_in)}
        self.branch = WgradBranch(dev)  # weight-gradient GEMMs run beside the dgrad chain
        self._bufs: Dict[int, dict] = {}

    # ------------------------------------------------------------------ buffers
    def _buffers(self, B: int) -> dict:
        if B in self._bufs:
            return self._bufs[B]
        dev, bf = self.model.device, torch.bfloat16
        z = lambda *shape, dtype=bf: torch.zeros(*shape, dtype=dtype, device=dev)  # noqa: E731
        total_c = sum(L.cout for L in self.layers)
        # one fp32 arena, zeroed once per step: per layer [sum, sumsq] of the forward and [sum g, sum g xhat] of the backward
        stats = z(4 * total_c, dtype=torch.float32)
        buf: dict = {"L": {}, "stats": stats}
        off = 0
        max_dcol = 0
        for L in self.layers:
            M = B * L.oh * L.ow
            buf["L"][L.conv] = {
                "col": None if L.implicit else z(M, _pad8(L.kdim)), "z": z(M, L.cout), "y": z(M, L.cout), "g": z(M, L.cout), "dz": z(M, L.cout), "dy": z(M, L.cout),
                "stats": stats[off : off + 2 * L.cout], "red": stats[off + 2 * L.cout : off + 4 * L.cout], "M": M,
            }
            off += 4 * L.cout
            if not L.implicit:
                max_dcol = max(max_dcol, M * _pad8(L.kdim))
        buf["dcol"] = z(max_dcol)  # transient: one shared buffer
        buf["tmp_dx"] = z(max(B * L.h * L.w * L.cin for L in self.layers))
        buf.update(pool=z(B, _pad8(self.fc_in)), dpool=z(B, _pad8(self.fc_in)), logits=z(B, self.fc_out, dtype=torch.float32), dzfc=z(B, _pad8(self.fc_out)))
        self._bufs[B] = buf
        return buf

    def refresh_weights(self) -> None:
        ops.cast_slab(self.model.slab, self.p16)
        self._refresh_padded()

    def _refresh_padded(self) -> None:
        ops.repack_conv_weights(self.model.slab, self.p16r, self._repack_table, self._n_repack, self._repack_max)

    # ------------------------------------------------------------------ forward
    def _convbn(self, L: _ConvBN, x, B: int, buf: dict, train: bool, residual=None, first: bool = False):
        lb, p = buf["L"][L.conv], self.par[L.conv]
        M = lb["M"]
        st = lb["stats"] if train else None
        cs, cq = (None, None) if st is None else (st[: L.cout], st[L.cout :])
        if L.implicit:
            lb["x"] = x  # the backward weight-gradient GEMM reads the input activation itself
            ops.conv_gemm(ops.CONV_FWD, x, B, L.h, L.w, L.cin, L.cout, L.k, weights=p["w16"], out_bf16=lb["z"], colsum=cs, colsumsq=cq)
        else:
            ops.im2col(x, lb["col"], B, L.h, L.w, L.cin, L.k, L.k, L.stride, L.pad, nchw_f32=first)
            ops.gemm_bf16(lb["col"], p["w16"], M, L.cout, L.kdim, out_bf16=lb["z"], colsum=cs, colsumsq=cq)
        ops.bn_apply(lb["z"], residual, lb["y"], st, p["gamma"], p["beta"], p["rm"], p["rv"], p["nb"] if train else None, M, L.cout, L.relu or residual is not None)
        return lb["y"]

    def _forward(self, x: torch.Tensor, buf: dict, train: bool) -> int:
        B = x.shape[0]
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        if train:
            buf["stats"].zero_()
        a = self._convbn(self.stem, x, B, buf, train, first=True)
        for blk in self.blocks:
            inp = a
            y1 = self._convbn(blk["c1"], inp, B, buf, train)
            sc = self._convbn(blk["sc"], inp, B, buf, train) if blk["sc"] is not None else inp
            a = self._convbn(blk["c2"], y1, B, buf, train, residual=sc)
        ops.avgpool_fwd(a, buf["pool"], B, self.final_hw, self.final_c)
        ops.gemm_bf16(buf["pool"], self.fc["w16"], B, self.fc_out, self.fc_in, out_f32=buf["logits"], bias=self.fc["b"])
        return B

    # ------------------------------------------------------------------ backward helpers
    def _bn_conv_backward(self, L: _ConvBN, dy, B: int, buf: dict, relu: bool, dx=None, add=None):
        """``dy`` = gradient w.r.t. this layer's output.  Accumulates the parameter gradients, writes the gradient of the layer
        input (plus ``add``, the gradient arriving through a second consumer of that input) into ``dx`` when given, and returns
        g, the masked gradient (= the gradient of the residual branch)."""
        lb, p = buf["L"][L.conv], self.par[L.conv]
        M = lb["M"]
        ops.bn_backward(dy, lb["y"], lb["z"], lb["g"], lb["dz"], lb["red"], lb["stats"], p["gamma"], p["dgamma"], p["dbeta"], M, L.cout, relu, red_is_zero=True)
        if L.implicit:
            tiles = ((L.cout + 127) // 128) * (L.kdim // 64)
            self.branch.run(lambda: ops.conv_gemm(ops.CONV_WGRAD, lb["x"], B, L.h, L.w, L.cin, L.cout, L.k, dz=lb["dz"], out_f32=p["gw"], split_k=max(1, min(296 // tiles, M // 128))))
            if dx is not None:
                ops.conv_gemm(ops.CONV_DGRAD, lb["dz"], B, L.h, L.w, L.cin, L.cout, L.k, weights=p["w16"], out_bf16=dx, add=add)
            return lb["g"]
        self.branch.run(lambda: ops.gemm_bf16(lb["dz"], lb["col"], L.cout, L.kdim, M, out_f32=p["gw"], a_mn=True, b_mn=True, split_k=max(1, min(32, M // 2048)), perm=(L.cin, L.k * L.k)))
        if dx is not None:
            dcol = buf["dcol"][: M * _pad8(L.kdim)].view(M, _pad8(L.kdim))
            ops.gemm_bf16(lb["dz"], p["w16"], M, L.kdim, L.cout, out_bf16=dcol, b_mn=True)
            ops.col2im(dcol, dx, B, L.h, L.w, L.cin, L.k, L.k, L.stride, L.pad, add=add)
        return lb["g"]

    # ------------------------------------------------------------------ training step
    def train_step(self, x: torch.Tensor, y: torch.Tensor, loss_out: torch.Tensor) -> None:
        self._refresh_padded()
        B = x.shape[0]
        buf = self._buffers(B)
        self._forward(x, buf, train=True)
        ops.xent_train(buf["logits"], y, loss_out, buf["dzfc"], self.fc["gb"], B, self.fc_out)
        self.branch.run(lambda: ops.gemm_bf16(buf["dzfc"], buf["pool"], self.fc_out, self.fc_in, B, out_f32=self.fc["gw"], a_mn=True, b_mn=True))
        ops.gemm_bf16(buf["dzfc"], self.fc["w16"], B, self.fc_in, self.fc_out, out_bf16=buf["dpool"], b_mn=True)
        last = self.blocks[-1]["c2"]
        ops.avgpool_bwd(buf["dpool"], buf["L"][last.conv]["dy"], B, self.final_hw, self.final_c)
        for bi in range(len(self.blocks) - 1, -1, -1):
            blk = self.blocks[bi]
            c1, c2, sc = blk["c1"], blk["c2"], blk["sc"]
            prev = self.blocks[bi - 1]["c2"] if bi > 0 else self.stem  # the layer whose output is this block's input
            d_in = buf["L"][prev.conv]["dy"]  # gradient w.r.t. the block input goes here
            # conv2 + bn2 (+ residual) + relu
            g2 = self._bn_conv_backward(c2, buf["L"][c2.conv]["dy"], B, buf, True, dx=buf["L"][c1.conv]["dy"])
            # the shortcut branch receives g2 (through its own conv + bn when the block changes shape)
            add = g2
            if sc is not None:
                add = buf["tmp_dx"][: B * sc.h * sc.w * sc.cin]
                self._bn_conv_backward(sc, g2, B, buf, False, dx=add)
            # conv1 + bn1 + relu ; its input gradient plus the shortcut's gradient = gradient of the block input
            self._bn_conv_backward(c1, buf["L"][c1.conv]["dy"], B, buf, True, dx=d_in, add=add)
        self._bn_conv_backward(self.stem, buf["L"][self.stem.conv]["dy"], B, buf, True)
        self.branch.join()

    @torch.no_grad()
    def eval_step(self, x: torch.Tensor, y: torch.Tensor, acc: torch.Tensor) -> None:
        buf = self._buffers(x.shape[0])
        B = self._forward(x, buf, train=False)
        ops.cross_entropy_eval(buf["logits"][:B], y, acc)

    def logits(self, x: torch.Tensor, train: bool = False) -> torch.Tensor:
        self.refresh_weights()
        buf = self._buffers(x.shape[0])
        B = self._forward(x, buf, train)
        return buf["logits"][:B].clone()


def attach_native_resnet(model, image_size: int = 32) -> NativeResNet:
    if model.device.type != "cuda":
        raise RuntimeError("the native ResNet path needs a CUDA device")
    nr = NativeResNet(model, image_size)
    module = model.get_model()
    module.native_train_step = nr.train_step
    module.native_eval = nr.eval_step
    module.native = nr
    return nr
