"""Hot ops.  CUDA tensors -> hand-written sm_100a kernels (libmyfyp_b200.so); CPU tensors -> PyTorch.

The CPU branch exists so the control plane and the oracles run on a box without a GPU; it is the
"plain PyTorch fp32 reference of the same op" the kernel tests compare against.  Kernel numbering
(A1..A8, T1..T8) follows SURVEY §8.
"""

from __future__ import annotations

import ctypes
from typing import Any, List, Optional, Sequence, Tuple

import torch

from myfyp_b200.ops import native

MAX_K = 64  # models per K-way kernel launch (kernel-argument pointer table)


def _cuda(t: torch.Tensor) -> bool:
    return t.is_cuda


def pick_device(slabs: Sequence[torch.Tensor], prefer: Optional[torch.device] = None) -> torch.device:
    """Where a K-way reduction runs: the preferred (receiver's) device if it is CUDA, else the first CUDA one."""
    if prefer is not None and prefer.type == "cuda":
        return prefer
    for s in slabs:
        if s.is_cuda:
            return s.device
    return slabs[0].device


def order_after_producers(slabs: Sequence[torch.Tensor], device: torch.device) -> None:
    """A kernel on ``device`` is about to read slabs that live on OTHER GPUs of this process (peer access).  Streams of
    different devices are ordered against nothing, so the consumer's stream is made to wait for everything queued so far on
    each producer GPU's current (default) stream - the stream all aggregation / snapshot / adoption work of a process runs on.
    One event record + one cross-device stream wait per foreign GPU, no host synchronisation."""
    if device.type != "cuda":
        return
    seen = set()
    foreign = {s.device.index for s in slabs if s.is_cuda and s.device != device}
    if foreign:
        from myfyp_b200.parallel.fabric import Fabric

        Fabric.ensure_peers(foreign | {device.index if device.index is not None else torch.cuda.current_device()})
    for s in slabs:
        if s.is_cuda and s.device != device and s.device not in seen:
            seen.add(s.device)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(s.device))
            torch.cuda.current_stream(device).wait_event(ev)


def _same_place(slabs: Sequence[torch.Tensor], device: torch.device) -> List[torch.Tensor]:
    """CPU inputs of a CUDA reduction are staged; CUDA inputs (any GPU) are read in place over NVLink."""
    if device.type == "cuda":
        order_after_producers(slabs, device)
        return [s if s.is_cuda else s.to(device, non_blocking=True) for s in slabs]
    return [s.cpu() if s.is_cuda else s for s in slabs]


# ====================================================================== A4: pull copy
def pull_copy(dst: torch.Tensor, src: torch.Tensor) -> None:
    """``dst <- src`` (flat fp32).  On CUDA: one vectorised pull kernel; ``src`` may be a peer GPU's slab."""
    if dst.data_ptr() == src.data_ptr() and dst.device == src.device:
        return
    n = dst.numel()
    if dst.is_cuda and src.is_cuda and src.dtype == torch.float32 and src.is_contiguous():
        order_after_producers([src], dst.device)
        lib = native.load()
        with native.device_guard(dst.device):
            native.check(lib.b2_pull_copy(native.ptr(dst), native.ptr(src), ctypes.c_int64(n), native.stream_ptr(dst.device)), "b2_pull_copy")
        native.count_launch()
    else:
        dst.copy_(src.reshape(-1)[:n].to(torch.float32), non_blocking=True)


# ====================================================================== A1: K-way weighted sum
def fedavg(slabs: Sequence[torch.Tensor], weights: Sequence[float], device: Optional[torch.device] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out = sum_k w_k * slab_k`` in ONE pass (K pulls fused with the weighted sum)."""
    device = device or pick_device(slabs)
    slabs = _same_place(slabs, device)
    n = slabs[0].numel()
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=device)
    if device.type == "cuda":
        lib = native.load()
        with native.device_guard(device):
            for c0 in range(0, len(slabs), MAX_K):
                chunk, w = slabs[c0 : c0 + MAX_K], weights[c0 : c0 + MAX_K]
                native.check(
                    lib.b2_fedavg(native.ptr(out), native.ptr_array(chunk), native.float_array(w), len(chunk), ctypes.c_int64(n), ctypes.c_int(1 if c0 else 0), native.stream_ptr(device)),
                    "b2_fedavg",
                )
                native.count_launch()
    else:
        acc = torch.zeros(n, dtype=torch.float32)
        for s, w in zip(slabs, weights):
            acc.add_(s.to(torch.float32), alpha=float(w))
        out.copy_(acc)
    return out


# ====================================================================== A3: streaming accumulate
def accumulate(acc: torch.Tensor, src: torch.Tensor, w: float, beta: float = 1.0) -> None:
    """``acc = beta*acc + w*src`` - the partial-model transfer itself when ``src`` is a peer slab."""
    if acc.is_cuda:
        lib = native.load()
        with native.device_guard(acc.device):
            native.check(
                lib.b2_accumulate(native.ptr(acc), native.ptr(src), ctypes.c_float(w), ctypes.c_float(beta), ctypes.c_int64(acc.numel()), native.stream_ptr(acc.device)),
                "b2_accumulate",
            )
        native.count_launch()
    else:
        acc.mul_(beta).add_(src.to(acc.device), alpha=float(w))


# ====================================================================== A5: median / trimmed mean
SHARD_MIN_ELEMS = 1 << 18  # below this a second GPU's launch + event costs more than its share of the reads saves


def shard_devices(slabs: Sequence[torch.Tensor], device: torch.device, force: bool = False) -> List[torch.device]:
    """GPUs of this process that take a coordinate shard of a K-way reduction whose result goes to ``device``: the destination
    plus every other GPU that holds one of the inputs (SURVEY 8, A5/A6 "coordinate-sharded across GPUs").

    Sharding pays when the single launch would pull most of its K inputs through ONE GPU's NVLink ingress: with the inputs spread
    over g GPUs that ingress carries (g-1)/g of K models unsharded, but only 2 (g-1)/g of ONE model sharded (the slices it reads
    plus the slices the others store into it).  With two GPUs the two are about equal and the second launch costs more than it
    saves (measured on 2 x B200, 4 x 64 MB: 280 us in one launch, 320 us sharded; profiles/peer_pull_timing.txt), so the
    automatic choice shards from three GPUs on; ``force`` shards whenever a second GPU holds an input."""
    if device.type != "cuda" or slabs[0].numel() < SHARD_MIN_ELEMS:
        return [device]
    others = sorted({s.device.index for s in slabs if s.is_cuda and s.device != device})
    if not others or (len(others) < 2 and not force):
        return [device]
    return [torch.device("cuda", i) for i in others] + [device]  # the destination launches last: the others are already running


def _shard_bounds(n: int, parts: int, align: int = 1024) -> List[int]:
    per = ((n + parts - 1) // parts + align - 1) // align * align
    return [min(i * per, n) for i in range(parts + 1)]


def _join_shards(workers: Sequence[torch.device], device: torch.device) -> None:
    """``device``'s stream waits for what the other shard workers just queued (one event each, no host synchronisation)."""
    for w in workers:
        if w != device:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(w))
            torch.cuda.current_stream(device).wait_event(ev)


def median(slabs: Sequence[torch.Tensor], trim: Optional[int] = None, device: Optional[torch.device] = None, out: Optional[torch.Tensor] = None,
           shard: Optional[bool] = None) -> torch.Tensor:
    """Coordinate-wise median (``trim=None``; even K -> mean of the two middle values, like NumPy) or trimmed mean.

    When the contributors live on several GPUs of this process the coordinates are sharded over those GPUs: each one sorts ONE
    contiguous slice - reading its K inputs in place over NVLink - and stores the result straight into ``out`` on ``device``
    (peer stores), so the K x model bytes arrive over all the GPUs' links instead of one.  ``shard``: None = automatic (three or
    more GPUs, see :func:`shard_devices`), True = whenever a second GPU holds an input, False = one launch on ``device``."""
    device = device or pick_device(slabs)
    slabs = _same_place(slabs, device)
    k, n = len(slabs), slabs[0].numel()
    if k > MAX_K:
        raise ValueError(f"median over {k} models: at most {MAX_K} per launch")
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=device)
    if device.type == "cuda":
        lib = native.load()
        workers = [device] if shard is False else shard_devices(slabs, device, force=bool(shard))
        bounds = _shard_bounds(n, len(workers))
        for w, lo, hi in zip(workers, bounds[:-1], bounds[1:]):
            if hi <= lo:
                continue
            if w != device:
                order_after_producers(list(slabs) + [out], w)
            ptrs = native.ptr_array(slabs)
            for i, t in enumerate(slabs):
                ptrs[i] = t.data_ptr() + 4 * lo
            with native.device_guard(w):
                native.check(
                    lib.b2_median(ctypes.c_void_p(out.data_ptr() + 4 * lo), ptrs, k, ctypes.c_int(-1 if trim is None else trim), ctypes.c_int64(hi - lo), native.stream_ptr(w)),
                    "b2_median",
                )
            native.count_launch()
        _join_shards(workers, device)
    else:
        st = torch.sort(torch.stack([s.to(torch.float32) for s in slabs]), dim=0).values
        if trim is None:
            out.copy_((st[(k - 1) // 2] + st[k // 2]) * 0.5)
        else:
            out.copy_(st[trim : k - trim].mean(dim=0))
    return out


# ====================================================================== A6/A7: Krum
def pairwise_sqdist(slabs: Sequence[torch.Tensor], device: Optional[torch.device] = None, passes: int = 3, shard: Optional[bool] = None) -> torch.Tensor:
    """K x K matrix of squared L2 distances.  CUDA: centred Gram matrix on tcgen05 (A6) + d = Gii+Gjj-2Gij.

    A Byzantine update may contain NaN / Inf / values whose square overflows.  Such a model is put at distance +inf from every
    other one (never NaN): the Gram matrix is centred on one of the models, so if THAT one is broken every entry is non-finite
    and the product is redone around the next candidate; otherwise only the broken models' rows and columns are."""
    device = device or pick_device(slabs)
    slabs = _same_place(slabs, device)
    k = len(slabs)
    bad: List[int] = []
    for ref in range(k):
        order = [ref] + [i for i in range(k) if i != ref]
        g = _centred_gram([slabs[i] for i in order], device, passes, shard)
        # a broken centre (NaN / Inf, or so large that every difference overflows) shows as a non-finite own entry or as a
        # majority of the other models looking broken
        if bool(torch.isfinite(g[0, 0])) and (k <= 2 or int((~torch.isfinite(torch.diagonal(g)[1:])).sum()) * 2 <= k - 1):
            break
        bad.append(ref)
    else:
        return torch.full((k, k), float("inf"), device=device).fill_diagonal_(0.0)
    d = torch.diagonal(g)
    dist = (d[:, None] + d[None, :] - 2 * g).clamp_min(0)
    broken = ~torch.isfinite(d)
    dist[broken, :] = float("inf")
    dist[:, broken] = float("inf")
    dist = torch.nan_to_num(dist, nan=float("inf"), posinf=float("inf"))
    inv = torch.empty(k, dtype=torch.long, device=dist.device)
    inv[torch.tensor(order, device=dist.device)] = torch.arange(k, device=dist.device)
    dist = dist[inv][:, inv]  # back to the caller's model order
    for b in bad:
        dist[b, :] = float("inf")
        dist[:, b] = float("inf")
    return dist.fill_diagonal_(0.0).float()


def _centred_gram(slabs: Sequence[torch.Tensor], device: torch.device, passes: int, shard: Optional[bool] = None) -> torch.Tensor:
    """fp64 view of G[i][j] = <slab_i - slab_0, slab_j - slab_0>."""
    k, n = len(slabs), slabs[0].numel()
    if device.type == "cuda":
        if k > MAX_K:
            raise ValueError(f"krum over {k} models: at most {MAX_K}")
        lib = native.load()
        # coordinate-sharded over the GPUs that hold the models: a Gram matrix is a sum over coordinates, so every GPU forms the
        # partial K x K matrix of ONE slice (centred on the same model) and the destination adds the K x K partials
        workers = [device] if shard is False else shard_devices(slabs, device, force=bool(shard))
        bounds = _shard_bounds(n, len(workers))
        parts = []
        for w, lo, hi in zip(workers, bounds[:-1], bounds[1:]):
            if hi <= lo:
                continue
            if w != device:
                order_after_producers(slabs, w)
            gram = torch.empty(MAX_K * MAX_K, dtype=torch.float32, device=w)
            ws = _krum_workspace(w)
            ptrs = native.ptr_array(slabs)
            for i, t in enumerate(slabs):
                ptrs[i] = t.data_ptr() + 4 * lo
            with native.device_guard(w):
                native.check(lib.b2_krum_gram(native.ptr(gram), ptrs, k, ctypes.c_int64(hi - lo), native.ptr(ws), ctypes.c_int(passes), native.stream_ptr(w)), "b2_krum_gram")
            native.count_launch()
            parts.append(gram)
        _join_shards(workers, device)
        total = None
        for g in parts:  # K x K partials: the ones born on other GPUs come over as 16 KB copies
            part = g.to(device).view(MAX_K, MAX_K)[:k, :k].double()
            total = part if total is None else total + part
        return total
    m = torch.stack([s.double() for s in slabs])
    m = m - m[0:1]
    return (m @ m.T).float().double()  # through fp32 like the kernel's result: values whose square overflows fp32 count as broken


_krum_ws = {}


def _krum_workspace(device: torch.device) -> torch.Tensor:
    """Split-K partial Grams + completion counter; zeroed once, the kernel leaves the counter at zero."""
    key = str(device)
    if key not in _krum_ws:
        lib = native.load()
        lib.b2_krum_gram_workspace_bytes.restype = ctypes.c_int64
        _krum_ws[key] = torch.zeros(int(lib.b2_krum_gram_workspace_bytes()) // 4, dtype=torch.float32, device=device)
    return _krum_ws[key]


def krum_select(dist: torch.Tensor, f: int, m: int) -> Tuple[List[float], List[int]]:
    """Krum scores (sum of the K-f-2 smallest distances to others) and the indices of the ``m`` best."""
    k = dist.shape[0]
    if dist.is_cuda:
        lib = native.load()
        scores = torch.empty(k, dtype=torch.float32, device=dist.device)
        sel = torch.zeros(k, dtype=torch.int32, device=dist.device)
        d = dist.contiguous()
        with native.device_guard(dist.device):
            native.check(lib.b2_krum_select(native.ptr(d), k, f, native.ptr(scores), native.ptr(sel), native.stream_ptr(dist.device)), "b2_krum_select")
        native.count_launch()
        return scores.tolist(), sel[:m].tolist()
    from myfyp_b200.learning.aggregators.krum import krum_select_reference

    return krum_select_reference(dist.numpy(), f, m)


# ====================================================================== A8 / T7: Scaffold
def scaffold_server_update(x: torch.Tensor, c: torch.Tensor, dys: Sequence[torch.Tensor], dcs: Sequence[torch.Tensor], wy: Sequence[float], wc: float) -> None:
    """``x += sum_k wy_k * dy_k``; ``c += wc * sum_k dc_k`` - one pass over 2K (peer) slabs."""
    if x.is_cuda:
        dys, dcs = _same_place(dys, x.device), _same_place(dcs, x.device)
        lib = native.load()
        with native.device_guard(x.device):
            native.check(
                lib.b2_scaffold_server(native.ptr(x), native.ptr(c), native.ptr_array(dys), native.ptr_array(dcs), native.float_array(wy), ctypes.c_float(wc), len(dys), ctypes.c_int64(x.numel()), native.stream_ptr(x.device)),
                "b2_scaffold_server",
            )
        native.count_launch()
    else:
        for dy, w in zip(dys, wy):
            x.add_(dy.to(x.device), alpha=float(w))
        for dc in dcs:
            c.add_(dc.to(c.device), alpha=float(wc))


def scaffold_finalize(c_i: torch.Tensor, dy: torch.Tensor, dc: torch.Tensor, x: torch.Tensor, y: torch.Tensor, k_lr: float) -> None:
    """``adj = (x-y)/(K*lr)``; ``dc = adj``; ``c_i += adj``; ``dy = y - x`` (reference scaffold_callback.py:124-140)."""
    if c_i.is_cuda:
        lib = native.load()
        with native.device_guard(c_i.device):
            native.check(
                lib.b2_scaffold_finalize(native.ptr(c_i), native.ptr(dy), native.ptr(dc), native.ptr(x), native.ptr(y), ctypes.c_float(1.0 / k_lr), ctypes.c_int64(c_i.numel()), native.stream_ptr(c_i.device)),
                "b2_scaffold_finalize",
            )
        native.count_launch()
    else:
        adj = (x - y) / k_lr
        dc.copy_(adj)
        c_i.add_(adj)
        dy.copy_(y - x)


# ====================================================================== T6: fused optimizers
def adam_step(
    p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int, lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
    weight_decay: float = 0.0, prox_mu: float = 0.0, prox_ref: Optional[torch.Tensor] = None, cv_global: Optional[torch.Tensor] = None,
    cv_local: Optional[torch.Tensor] = None, cv_scale: float = 1.0, zero_grad: bool = True, dev_state: Optional[torch.Tensor] = None,
    step_offset: int = 0, p16: Optional[torch.Tensor] = None,
) -> None:
    """Adam over the flat slab.  ``g' = g + wd*p + mu*(p - ref) + cv_scale*(c - c_i)``; clears ``g`` afterwards.

    ``dev_state`` (floats on the device: step, 1-b1^step, 1-b2^step) replaces the host-side ``step``, which is what makes a
    whole epoch CUDA-graph-capturable.  Two ways: ``step_offset == 0`` - a 1-thread kernel advances the state before the
    update; ``step_offset = k > 0`` - the step is ``dev_state[0] + k`` (k = index inside the captured epoch, baked into
    the launch) and :func:`adam_advance` adds the epoch's length once per replay (one launch per step instead of two).
    ``p16``: bf16 mirror of ``p`` written in the same pass (the tensor-core operands of the next step).
    """
    if p.is_cuda:
        lib = native.load()
        bc1, bc2 = (1.0 - beta1**step, 1.0 - beta2**step) if dev_state is None else (1.0, 1.0)
        with native.device_guard(p.device):
            if dev_state is not None and step_offset == 0:
                native.check(lib.b2_adam_prepare(native.ptr(dev_state), ctypes.c_float(beta1), ctypes.c_float(beta2), native.stream_ptr(p.device)), "b2_adam_prepare")
                native.count_launch()
            native.check(
                lib.b2_adam_step(
                    native.ptr(p), native.ptr(g), native.ptr(m), native.ptr(v), ctypes.c_int64(p.numel()), ctypes.c_float(lr), ctypes.c_float(beta1), ctypes.c_float(beta2),
                    ctypes.c_float(eps), ctypes.c_float(weight_decay), ctypes.c_float(bc1), ctypes.c_float(bc2), native.ptr(dev_state), ctypes.c_float(float(step_offset)), native.ptr(p16),
                    ctypes.c_float(prox_mu if prox_ref is not None else 0.0),
                    native.ptr(prox_ref), native.ptr(cv_global), native.ptr(cv_local), ctypes.c_float(cv_scale), ctypes.c_int(1 if zero_grad else 0), native.stream_ptr(p.device),
                ),
                "b2_adam_step",
            )
        native.count_launch()
        return
    if dev_state is not None:
        if step_offset > 0:
            step = int(dev_state[0].item()) + step_offset
        else:
            dev_state[0] += 1
            step = int(dev_state[0].item())
    gg = _effective_grad(p, g, weight_decay, prox_mu, prox_ref, cv_global, cv_local, cv_scale)
    m.mul_(beta1).add_(gg, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(gg, gg, value=1 - beta2)
    denom = (v / (1 - beta2**step)).sqrt_().add_(eps)
    p.data.addcdiv_(m / (1 - beta1**step), denom, value=-lr)
    if p16 is not None:
        p16.copy_(p.data)
    if zero_grad:
        g.zero_()


def adam_advance(dev_state: torch.Tensor, n_steps: int) -> None:
    """``dev_state[0] += n_steps`` (once per epoch replay, see :func:`adam_step`)."""
    if dev_state.is_cuda:
        lib = native.load()
        with native.device_guard(dev_state.device):
            native.check(lib.b2_adam_advance(native.ptr(dev_state), ctypes.c_float(float(n_steps)), native.stream_ptr(dev_state.device)), "b2_adam_advance")
        native.count_launch()
    else:
        dev_state[0] += n_steps


def cast_slab(p: torch.Tensor, p16: torch.Tensor) -> None:
    """bf16 mirror of a whole fp32 slab in one pass."""
    if p.is_cuda:
        lib = native.load()
        with native.device_guard(p.device):
            native.check(lib.b2_cast_slab(native.ptr(p), native.ptr(p16), ctypes.c_int64(p.numel()), native.stream_ptr(p.device)), "b2_cast_slab")
        native.count_launch()
    else:
        p16.copy_(p)


def sgd_step(
    p: torch.Tensor, g: torch.Tensor, buf: torch.Tensor, lr: float = 1e-2, momentum: float = 0.0, weight_decay: float = 0.0, prox_mu: float = 0.0,
    prox_ref: Optional[torch.Tensor] = None, cv_global: Optional[torch.Tensor] = None, cv_local: Optional[torch.Tensor] = None, cv_scale: float = 1.0, zero_grad: bool = True,
) -> None:
    """SGD(+momentum) over the flat slab with the same fused FedProx / Scaffold terms."""
    if p.is_cuda:
        lib = native.load()
        with native.device_guard(p.device):
            native.check(
                lib.b2_sgd_step(
                    native.ptr(p), native.ptr(g), native.ptr(buf), ctypes.c_int64(p.numel()), ctypes.c_float(lr), ctypes.c_float(momentum), ctypes.c_float(weight_decay),
                    ctypes.c_float(prox_mu if prox_ref is not None else 0.0), native.ptr(prox_ref), native.ptr(cv_global), native.ptr(cv_local), ctypes.c_float(cv_scale),
                    ctypes.c_int(1 if zero_grad else 0), native.stream_ptr(p.device),
                ),
                "b2_sgd_step",
            )
        native.count_launch()
        return
    gg = _effective_grad(p, g, weight_decay, prox_mu, prox_ref, cv_global, cv_local, cv_scale)
    if momentum != 0.0:
        buf.mul_(momentum).add_(gg)
        gg = buf
    p.data.add_(gg, alpha=-lr)
    if zero_grad:
        g.zero_()


def _effective_grad(p, g, wd, mu, ref, c, c_i, cv_scale) -> torch.Tensor:
    gg = g.clone()
    if wd:
        gg.add_(p.data, alpha=wd)
    if ref is not None and mu:
        gg.add_(p.data - ref, alpha=mu)
    if c is not None and c_i is not None:
        gg.add_(c - c_i, alpha=cv_scale)
    return gg


# ====================================================================== T3/T8: fused softmax cross-entropy
class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        b, c = logits.shape
        lg = logits.contiguous().float()
        loss = torch.zeros(1, dtype=torch.float32, device=logits.device)
        dlogits = torch.empty_like(lg)
        lib = native.load()
        with native.device_guard(logits.device):
            native.check(
                lib.b2_xent_fwd_bwd(native.ptr(lg), native.ptr(labels), native.ptr(loss), native.ptr(dlogits), native.ptr(None), b, c, ctypes.c_float(1.0 / b), native.stream_ptr(logits.device)),
                "b2_xent_fwd_bwd",
            )
        native.count_launch()
        ctx.save_for_backward(dlogits)
        return loss[0]

    @staticmethod
    def backward(ctx, grad_out):
        (dlogits,) = ctx.saved_tensors
        return dlogits * grad_out, None


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Mean NLL of log-softmax(logits); forward computes ``dlogits`` too (one kernel, no softmax round trip)."""
    if logits.is_cuda:
        return _XentFn.apply(logits, labels)
    return torch.nn.functional.cross_entropy(logits.float(), labels)


def cross_entropy_eval(logits: torch.Tensor, labels: torch.Tensor, acc: torch.Tensor) -> None:
    """``acc[0] += sum of per-sample NLL``; ``acc[1] += number of correct argmax`` (fused T3+T8)."""
    b, c = logits.shape
    if logits.is_cuda:
        lg = logits.contiguous().float()
        lib = native.load()
        with native.device_guard(logits.device):
            native.check(
                lib.b2_xent_fwd_bwd(native.ptr(lg), native.ptr(labels), native.ptr(acc[0:1]), native.ptr(None), native.ptr(acc[1:2]), b, c, ctypes.c_float(1.0), native.stream_ptr(logits.device)),
                "b2_xent_fwd_bwd",
            )
        native.count_launch()
    else:
        acc[0] += torch.nn.functional.cross_entropy(logits.float(), labels, reduction="sum")
        acc[1] += (logits.argmax(dim=1) == labels).sum()


def confusion(logits: torch.Tensor, labels: torch.Tensor, conf: torch.Tensor) -> None:
    """``conf[label, argmax] += 1`` (``conf`` is a C x C fp32 matrix)."""
    b, c = logits.shape
    if logits.is_cuda:
        lg = logits.contiguous().float()
        lib = native.load()
        with native.device_guard(logits.device):
            native.check(lib.b2_confusion(native.ptr(lg), native.ptr(labels), native.ptr(conf), b, c, native.stream_ptr(logits.device)), "b2_confusion")
        native.count_launch()
    else:
        idx = labels * c + logits.argmax(dim=1)
        conf.view(-1).index_add_(0, idx, torch.ones_like(idx, dtype=conf.dtype))


def prf_from_confusion(conf: torch.Tensor) -> dict:
    """Macro precision / recall / F1 from a confusion matrix (rows = true class, columns = predicted)."""
    conf = conf.double().cpu()
    tp = conf.diag()
    prec = tp / conf.sum(0).clamp_min(1e-12)
    rec = tp / conf.sum(1).clamp_min(1e-12)
    f1 = 2 * prec * rec / (prec + rec).clamp_min(1e-12)
    present = conf.sum(1) > 0
    return {"test_precision": float(prec[present].mean()), "test_recall": float(rec[present].mean()), "test_f1": float(f1[present].mean())}


# ====================================================================== T1/T2: tcgen05 GEMM + staging
class _GemmArgs(ctypes.Structure):
    _fields_ = [
        ("A", ctypes.c_void_p), ("B", ctypes.c_void_p), ("M", ctypes.c_int), ("N", ctypes.c_int), ("K", ctypes.c_int), ("lda", ctypes.c_int), ("ldb", ctypes.c_int),
        ("out_f32", ctypes.c_void_p), ("out_bf16", ctypes.c_void_p), ("out_bf16_t", ctypes.c_void_p), ("ldo", ctypes.c_int), ("ldo16", ctypes.c_int), ("ldt", ctypes.c_int),
        ("bias", ctypes.c_void_p), ("mask", ctypes.c_void_p), ("ldm", ctypes.c_int), ("colsum", ctypes.c_void_p), ("relu", ctypes.c_int), ("accumulate", ctypes.c_int),
        ("a_mn", ctypes.c_int), ("b_mn", ctypes.c_int), ("split_k", ctypes.c_int), ("colsumsq", ctypes.c_void_p), ("perm_cin", ctypes.c_int), ("perm_taps", ctypes.c_int),
    ]


def gemm_bf16(
    a: torch.Tensor, b: torch.Tensor, m: int, n: int, k: int, out_f32: Optional[torch.Tensor] = None, out_bf16: Optional[torch.Tensor] = None,
    out_bf16_t: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None, mask: Optional[torch.Tensor] = None, colsum: Optional[torch.Tensor] = None,
    relu: bool = False, accumulate: bool = False, a_mn: bool = False, b_mn: bool = False, split_k: int = 1, colsumsq: Optional[torch.Tensor] = None,
    perm: Optional[tuple] = None,
) -> None:
    """``D[m,n] = epilogue(A[m,k] @ B[n,k]^T)`` on tcgen05 tensor cores (bf16 in, fp32 accumulate); see csrc/gemm.cu.

    ``a_mn`` / ``b_mn``: the operand is stored transposed (``a`` is ``[k, m]``, ``b`` is ``[k, n]``) and consumed as an
    MN-major UMMA operand - the backward GEMMs read the forward pass's buffers without transposed copies.

    ``perm=(cin, taps)``: ``out_f32`` column ``tap*cin + c`` is stored at ``c*taps + tap`` (weight gradient of a convolution whose
    patch matrix is in (tap, c) order, written in PyTorch's ``[cout, cin, kh, kw]`` layout).

    ``a``/``b`` are 2-D bf16 tensors whose row pitch (``stride(0)``) is the leading dimension; the logical sizes
    ``m, n, k`` may be smaller than the allocations (TMA zero-fills the tails).
    """
    if not a.is_cuda:
        raise RuntimeError("gemm_bf16 is the tcgen05 path; on CPU use torch.matmul")

    def ld(t: Optional[torch.Tensor]) -> int:
        return 0 if t is None else int(t.stride(0))

    args = _GemmArgs(
        a.data_ptr(), b.data_ptr(), m, n, k, ld(a), ld(b), 0 if out_f32 is None else out_f32.data_ptr(), 0 if out_bf16 is None else out_bf16.data_ptr(),
        0 if out_bf16_t is None else out_bf16_t.data_ptr(), ld(out_f32), ld(out_bf16), ld(out_bf16_t), 0 if bias is None else bias.data_ptr(),
        0 if mask is None else mask.data_ptr(), ld(mask), 0 if colsum is None else colsum.data_ptr(), int(relu), int(accumulate),
        int(a_mn), int(b_mn), int(split_k), 0 if colsumsq is None else colsumsq.data_ptr(), 0 if perm is None else int(perm[0]), 0 if perm is None else int(perm[1]),
    )
    lib = native.load()
    with native.device_guard(a.device):
        native.check(lib.b2_gemm_bf16(ctypes.byref(args), native.stream_ptr(a.device)), "b2_gemm_bf16")
    native.count_launch()


class _ConvGemmArgs(ctypes.Structure):
    _fields_ = [
        ("mode", ctypes.c_int), ("act", ctypes.c_void_p), ("weights", ctypes.c_void_p), ("dz", ctypes.c_void_p), ("batch", ctypes.c_int), ("h", ctypes.c_int),
        ("w", ctypes.c_int), ("cin", ctypes.c_int), ("cout", ctypes.c_int), ("ksize", ctypes.c_int), ("out_f32", ctypes.c_void_p), ("out_bf16", ctypes.c_void_p),
        ("bias", ctypes.c_void_p), ("add_bf16", ctypes.c_void_p), ("colsum", ctypes.c_void_p), ("colsumsq", ctypes.c_void_p), ("relu", ctypes.c_int),
        ("accumulate", ctypes.c_int), ("split_k", ctypes.c_int),
    ]


CONV_FWD, CONV_DGRAD, CONV_WGRAD = 0, 1, 2


def conv_gemm_supported(h: int, w: int, cin: int, cout: int, ksize: int, stride: int, pad: int) -> bool:
    """Geometries the implicit-GEMM convolution takes (everything else goes through im2col + gemm_bf16)."""
    if stride != 1 or ksize % 2 == 0 or pad != ksize // 2 or cin % 64 or cout % 64 or w > 128 or 128 % w:
        return False
    for pixels in (128, 64):
        th = pixels // w
        if th <= h:
            if h % th:
                return False
        elif pixels % (h * w):
            return False
    return True


def conv_gemm(
    mode: int, act: torch.Tensor, batch: int, h: int, w: int, cin: int, cout: int, ksize: int, weights: Optional[torch.Tensor] = None, dz: Optional[torch.Tensor] = None,
    out_f32: Optional[torch.Tensor] = None, out_bf16: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None, add: Optional[torch.Tensor] = None,
    colsum: Optional[torch.Tensor] = None, colsumsq: Optional[torch.Tensor] = None, relu: bool = False, accumulate: bool = False, split_k: int = 1,
) -> None:
    """Implicit-GEMM convolution (stride 1, "same" padding) on the tcgen05 GEMM: 4-D TMA boxes of the NHWC activation are the operand
    tiles, no patch matrix is written.  ``weights``: bf16 ``[cout, k*k*cin]`` in (kh, kw, cin) order (:func:`repack_conv_weights`).

    ``CONV_FWD``: ``act`` = x -> ``out[batch*h*w, cout]``; ``CONV_DGRAD``: ``act`` = dz -> ``out[batch*h*w, cin]`` (+ ``add``);
    ``CONV_WGRAD``: ``act`` = x, ``dz`` -> ``out_f32[cout, cin, k, k]``."""
    args = _ConvGemmArgs(
        mode, act.data_ptr(), 0 if weights is None else weights.data_ptr(), 0 if dz is None else dz.data_ptr(), batch, h, w, cin, cout, ksize,
        0 if out_f32 is None else out_f32.data_ptr(), 0 if out_bf16 is None else out_bf16.data_ptr(), 0 if bias is None else bias.data_ptr(),
        0 if add is None else add.data_ptr(), 0 if colsum is None else colsum.data_ptr(), 0 if colsumsq is None else colsumsq.data_ptr(), int(relu), int(accumulate), int(split_k),
    )
    lib = native.load()
    with native.device_guard(act.device):
        native.check(lib.b2_conv_gemm(ctypes.byref(args), native.stream_ptr(act.device)), "b2_conv_gemm")
    native.count_launch()


def repack_conv_weights(slab: torch.Tensor, dst: torch.Tensor, table: torch.Tensor, n_layers: int, max_elems: int) -> None:
    """fp32 conv weights ``[cout, cin, taps]`` inside ``slab`` -> bf16 ``[cout, taps, cin]`` in ``dst`` for all layers of ``table`` (one launch)."""
    lib = native.load()
    with native.device_guard(slab.device):
        native.check(lib.b2_repack_conv_weights(native.ptr(slab), native.ptr(dst), native.ptr(table), n_layers, ctypes.c_int64(max_elems), native.stream_ptr(slab.device)), "b2_repack_conv_weights")
    native.count_launch()


def cast_transpose(src: torch.Tensor, rows: int, cols: int, dst: Optional[torch.Tensor] = None, dst_t: Optional[torch.Tensor] = None, colsum: Optional[torch.Tensor] = None) -> None:
    """fp32 ``src[rows, cols]`` -> bf16 copy ``dst`` and/or transposed bf16 copy ``dst_t`` (+ column sums), one pass."""
    lib = native.load()
    with native.device_guard(src.device):
        native.check(
            lib.b2_cast_transpose(
                native.ptr(src), rows, cols, int(src.stride(0)), native.ptr(dst), 0 if dst is None else int(dst.stride(0)), native.ptr(dst_t),
                0 if dst_t is None else int(dst_t.stride(0)), native.ptr(colsum), native.stream_ptr(src.device),
            ),
            "b2_cast_transpose",
        )
    native.count_launch()


# ====================================================================== T4/T5: convolution building blocks
def _conv_call(name: str, dev, *args) -> None:
    lib = native.load()
    with native.device_guard(dev):
        native.check(getattr(lib, name)(*args, native.stream_ptr(dev)), name)
    native.count_launch()


def im2col(x: torch.Tensor, col: torch.Tensor, B: int, H: int, W: int, C: int, KH: int, KW: int, stride: int, pad: int, nchw_f32: bool = False) -> None:
    """Patch gather ``col[(b,oh,ow), (kh*KW + kw)*C + c]`` (bf16); ``x`` is NHWC bf16, or NCHW fp32 for the input layer."""
    _conv_call("b2_im2col", x.device, native.ptr(x), int(nchw_f32), native.ptr(col), int(col.stride(0)), B, H, W, C, KH, KW, stride, pad)


def col2im(dcol: torch.Tensor, dx: torch.Tensor, B: int, H: int, W: int, C: int, KH: int, KW: int, stride: int, pad: int, add: Optional[torch.Tensor] = None) -> None:
    """``dx = col2im(dcol) (+ add)`` - gather form, no atomics."""
    _conv_call("b2_col2im", dcol.device, native.ptr(dcol), int(dcol.stride(0)), native.ptr(dx), native.ptr(add), B, H, W, C, KH, KW, stride, pad)


def bn_apply(z, residual, y, stats, gamma, beta, running_mean, running_var, num_batches, M: int, C: int, relu: bool, eps: float = 1e-5, momentum: float = 0.1) -> None:
    """Batch-norm apply on ``[M, C]`` bf16 (training when ``stats`` = [sum, sumsq] is given, else running statistics)."""
    _conv_call("b2_bn_apply", z.device, native.ptr(z), native.ptr(residual), native.ptr(y), native.ptr(stats), native.ptr(gamma), native.ptr(beta), native.ptr(running_mean),
               native.ptr(running_var), native.ptr(num_batches), ctypes.c_float(eps), ctypes.c_float(momentum), M, C, int(relu))


def bn_backward(dy, y, z, g, dz, red, stats, gamma, dgamma, dbeta, M: int, C: int, relu: bool, eps: float = 1e-5, red_is_zero: bool = False) -> None:
    """Batch-norm backward (two launches): ``g = dy*(y>0)``, ``dz``, ``dgamma +=``, ``dbeta +=``.  ``red`` (2C floats of scratch) is
    zeroed first unless ``red_is_zero`` (a training step clears one arena for all its layers instead)."""
    _conv_call("b2_bn_backward", dy.device, native.ptr(dy), native.ptr(y), native.ptr(z), native.ptr(g), native.ptr(dz), native.ptr(red), native.ptr(stats), native.ptr(gamma),
               native.ptr(dgamma), native.ptr(dbeta), ctypes.c_float(eps), M, C, int(relu), int(red_is_zero))
    native.count_launch(1)


def avgpool_fwd(x, out, B: int, HW: int, C: int) -> None:
    _conv_call("b2_avgpool_fwd", x.device, native.ptr(x), native.ptr(out), int(out.stride(0)), B, HW, C)


def avgpool_bwd(dout, dx, B: int, HW: int, C: int) -> None:
    _conv_call("b2_avgpool_bwd", dout.device, native.ptr(dout), int(dout.stride(0)), native.ptr(dx), B, HW, C)


def maxpool2_fwd(y: torch.Tensor, p: torch.Tensor, idx: torch.Tensor, B: int, H: int, W: int, C: int) -> None:
    _conv_call("b2_maxpool2_fwd", y.device, native.ptr(y), native.ptr(p), native.ptr(idx), B, H, W, C)


def maxpool2_relu_bwd(dp: torch.Tensor, idx: torch.Tensor, y: torch.Tensor, dz: torch.Tensor, B: int, H: int, W: int, C: int) -> None:
    _conv_call("b2_maxpool2_relu_bwd", dp.device, native.ptr(dp), native.ptr(idx), native.ptr(y), native.ptr(dz), B, H, W, C)


def colsum_bf16(z: torch.Tensor, M: int, C: int, out: torch.Tensor) -> None:
    _conv_call("b2_colsum_bf16", z.device, native.ptr(z), M, C, int(z.stride(0)), native.ptr(out))


def pool_im2col(y: torch.Tensor, idx: torch.Tensor, col: torch.Tensor, B: int, PH: int, PW: int, C: int, KH: int, KW: int, stride: int, pad: int) -> None:
    """Fused 2x2 max-pool of ``y [B,2PH,2PW,C]`` + patch gather of the next layer (geometry over the pooled map)."""
    _conv_call("b2_pool_im2col", y.device, native.ptr(y), native.ptr(idx), native.ptr(col), int(col.stride(0)), B, PH, PW, C, KH, KW, stride, pad)


def unpool_col2im(dcol: torch.Tensor, idx: torch.Tensor, y: torch.Tensor, dz: torch.Tensor, bias_grad: Optional[torch.Tensor], B: int, PH: int, PW: int, C: int, KH: int, KW: int, stride: int, pad: int) -> None:
    """Fused col2im of the next layer's ``dcol`` + unpool to the winners + ReLU mask -> ``dz``; ``bias_grad += colsum(dz)``."""
    _conv_call("b2_unpool_col2im", dcol.device, native.ptr(dcol), int(dcol.stride(0)), native.ptr(idx), native.ptr(y), native.ptr(dz), native.ptr(bias_grad), B, PH, PW, C, KH, KW, stride, pad)


def xent_train(logits: torch.Tensor, labels: torch.Tensor, loss_out: torch.Tensor, dl16: torch.Tensor, bias_grad: Optional[torch.Tensor], B: int, C: int) -> None:
    """One launch: ``loss_out[0] = mean NLL``, ``dl16 = bf16(dlogits / B)``, ``bias_grad += colsum(dlogits / B)``."""
    lib = native.load()
    with native.device_guard(logits.device):
        native.check(
            lib.b2_xent_train(native.ptr(logits), native.ptr(labels), native.ptr(loss_out), native.ptr(dl16), int(dl16.stride(0)), native.ptr(bias_grad), B, C, ctypes.c_float(1.0 / B), native.stream_ptr(logits.device)),
            "b2_xent_train",
        )
    native.count_launch()
