"""Hand-written training step for the MLP: no autograd, no library GEMM.

Attached to a CUDA ``TorchP2PFLModel`` wrapping :class:`~myfyp_b200.models.mlp.MLP`.  Parameters stay the fp32
slab (that is what gets aggregated / published); every step stages bf16 copies for the tensor cores.

Per step (kernel ids from SURVEY section 8):
  staging   ``cast_transpose``: X -> bf16 (the weights' bf16 mirror is written by the fused optimizer itself)
  forward   T1 ``gemm_bf16`` x L: H_l = relu(H_{l-1} W_l^T + b_l) (bf16) ; last layer -> fp32 logits
  loss      T3 ``xent_train``: loss, bf16 dlogits (already scaled by 1/B), bias gradient of the last layer
  backward  T2 wgrad: dW_l = dZ_l^T H_{l-1}  (fp32 straight into the gradient slab; both operands MN-major)
            T2 dgrad: dZ_{l-1} = (dZ_l W_l) * (H_{l-1} > 0), column sums -> db_{l-1}     (W_l MN-major)
  (the fused optimizer T6 then runs over the whole slab - called by the learner)

The backward GEMMs read the forward pass's buffers as MN-major tensor-core operands, so nothing is ever transposed.

Reference path for the same math: three ``nn.Linear`` + ReLU + ``log_softmax`` + ``cross_entropy`` through
Lightning / autograd in fp32 (p2pfl/learning/frameworks/pytorch/lightning_model.py:169-191).
"""

from __future__ import annotations

from typing import Dict, List

import torch

from myfyp_b200 import ops
from myfyp_b200.models._branch import WgradBranch


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


class NativeMLP:
    def __init__(self, model) -> None:
        self.model = model
        module = model.get_model()
        self.sizes: List[int] = list(module.sizes)
        self.L = len(self.sizes) - 1
        dev = model.device
        lay = model.layout
        self.w, self.b, self.gw, self.gb = [], [], [], []
        for i in range(self.L):
            ew, eb = lay.entries[f"layers.{i}.weight"], lay.entries[f"layers.{i}.bias"]
            self.w.append(model.slab[ew.offset : ew.offset + ew.numel].view(ew.shape))
            self.b.append(model.slab[eb.offset : eb.offset + eb.numel])
            self.gw.append(model.grad_slab[ew.offset : ew.offset + ew.numel].view(ew.shape))
            self.gb.append(model.grad_slab[eb.offset : eb.offset + eb.numel])
        # bf16 mirror of the whole slab (kept current by the fused optimizer); weights whose row length is a multiple of 8
        # are used in place as TMA operands, the others get a padded private copy
        self.p16 = torch.zeros(lay.total, dtype=torch.bfloat16, device=dev)
        self.w16, self._padded = [], []
        for i in range(self.L):
            ew = lay.entries[f"layers.{i}.weight"]
            if self.sizes[i] % 8 == 0:
                self.w16.append(self.p16[ew.offset : ew.offset + ew.numel].view(ew.shape))
            else:
                self.w16.append(torch.zeros(self.sizes[i + 1], _pad8(self.sizes[i]), dtype=torch.bfloat16, device=dev))
                self._padded.append(i)
        self.branch = WgradBranch(model.device)
        self._bufs: Dict[int, dict] = {}

    # ------------------------------------------------------------------ buffers per batch capacity
    def _buffers(self, batch: int) -> dict:
        cap = _pad8(batch)
        if cap not in self._bufs:
            dev, bf, s = self.model.device, torch.bfloat16, self.sizes
            self._bufs[cap] = {
                "h": [torch.zeros(cap, _pad8(s[i]), dtype=bf, device=dev) for i in range(self.L)],  # H_0 (=X) .. H_{L-1}
                "dz": [torch.zeros(cap, _pad8(s[i + 1]), dtype=bf, device=dev) for i in range(self.L)],  # dZ_1 .. dZ_L
                "logits": torch.zeros(cap, s[-1], dtype=torch.float32, device=dev),
                "dlogits": torch.zeros(cap, s[-1], dtype=torch.float32, device=dev),
            }
        return self._bufs[cap]

    def refresh_weights(self) -> None:
        """fp32 slab -> bf16 mirror, once per fit / evaluation (the optimizer updates the mirror after every step)."""
        ops.cast_slab(self.model.slab, self.p16)
        self._refresh_padded()

    def _refresh_padded(self) -> None:
        for i in self._padded:
            ops.cast_transpose(self.w[i], self.sizes[i + 1], self.sizes[i], dst=self.w16[i])

    # ------------------------------------------------------------------ forward
    def _forward(self, x: torch.Tensor, buf: dict) -> int:
        bsz = x.shape[0]
        x2 = x.reshape(bsz, -1)
        if x2.dtype != torch.float32 or not x2.is_contiguous():
            x2 = x2.float().contiguous()
        s = self.sizes
        ops.cast_transpose(x2, bsz, s[0], dst=buf["h"][0])
        for i in range(self.L):
            last = i == self.L - 1
            ops.gemm_bf16(
                buf["h"][i], self.w16[i], bsz, s[i + 1], s[i],
                out_f32=buf["logits"] if last else None, out_bf16=None if last else buf["h"][i + 1], bias=self.b[i], relu=not last,
            )
        return bsz

    # ------------------------------------------------------------------ public hooks used by TorchLearner
    def train_step(self, x: torch.Tensor, y: torch.Tensor, loss_out: torch.Tensor) -> None:
        """Forward + loss + backward; gradients land in the model's gradient slab, mean loss in ``loss_out[0]``."""
        self._refresh_padded()
        buf = self._buffers(x.shape[0])
        bsz = self._forward(x, buf)
        s, L = self.sizes, self.L
        # loss, dZ_L = dlogits/B as bf16, bias gradient of the last layer: one launch
        ops.xent_train(buf["logits"], y, loss_out, buf["dz"][L - 1], self.gb[L - 1], bsz, s[-1])
        for i in range(L - 1, -1, -1):
            # wgrad: dW_i[out,in] = dZ_i^T H_i      - dZ_i [B,out] and H_i [B,in] consumed as MN-major operands (K = batch);
            # off the critical path (side stream / parallel graph branch): only the optimizer needs it
            self.branch.run(lambda i=i: ops.gemm_bf16(buf["dz"][i], buf["h"][i], s[i + 1], s[i], bsz, out_f32=self.gw[i], a_mn=True, b_mn=True))
            if i > 0:
                # dgrad: dZ_{i-1}[B,in] = (dZ_i[B,out] W_i[out,in]) * (H_i > 0) - W_i consumed MN-major; column sums -> db_{i-1}
                ops.gemm_bf16(buf["dz"][i], self.w16[i], bsz, s[i], s[i + 1], out_bf16=buf["dz"][i - 1], mask=buf["h"][i], colsum=self.gb[i - 1], b_mn=True)
        self.branch.join()

    @torch.no_grad()
    def eval_step(self, x: torch.Tensor, y: torch.Tensor, acc: torch.Tensor) -> None:
        """``acc[0] += sum NLL``, ``acc[1] += #correct``  (the caller refreshed the bf16 mirror)."""
        buf = self._buffers(x.shape[0])
        bsz = self._forward(x, buf)
        ops.cross_entropy_eval(buf["logits"][:bsz], y, acc)

    def logits(self, x: torch.Tensor) -> torch.Tensor:
        self.refresh_weights()
        buf = self._buffers(x.shape[0])
        bsz = self._forward(x, buf)
        return buf["logits"][:bsz].clone()


def attach_native_mlp(model) -> NativeMLP:
    """Give the wrapped module ``native_train_step`` / ``native_eval`` so the learner bypasses autograd."""
    if model.device.type != "cuda":
        raise RuntimeError("the native MLP path needs a CUDA device")
    nm = NativeMLP(model)
    module = model.get_model()
    module.native_train_step = nm.train_step
    module.native_eval = nm.eval_step
    module.native = nm
    return nm
