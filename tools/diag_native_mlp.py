"""Diagnostic: native MLP gradients vs (a) fp32 autograd, (b) a torch emulation with bf16 rounding at the same points."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200.models import mlp_model

torch.manual_seed(0)
dev = torch.device("cuda:0")
for batch in (32, 256):
    model = mlp_model(device="cuda:0", seed=3, native=True)
    module = model.get_model()
    x = torch.randn(batch, 28, 28, device=dev)
    y = torch.randint(0, 10, (batch,), device=dev)
    loss = torch.zeros(1, device=dev)
    module.native.refresh_weights()
    module.native_train_step(x, y, loss)
    g_nat = model.grad_slab.clone(); model.grad_slab.zero_()
    l32 = torch.nn.functional.cross_entropy(module(x), y); l32.backward()
    g32 = model.grad_slab.clone(); model.grad_slab.zero_()
    # emulation
    bf = lambda t: t.to(torch.bfloat16).float()
    W = [bf(l.weight.detach()) for l in module.layers]; b = [l.bias.detach() for l in module.layers]
    h = [bf(x.reshape(batch, -1))]
    for i in range(3):
        z = h[i] @ W[i].T + b[i]
        if i < 2: h.append(bf(torch.relu(z)))
        else: logits = z
    p = torch.softmax(logits, 1); dl = p.clone(); dl[torch.arange(batch), y] -= 1; dl /= batch
    gw, gb = [None]*3, [None]*3
    dz = bf(dl); gb[2] = dl.sum(0)
    for i in (2, 1, 0):
        gw[i] = dz.T @ h[i]
        if i > 0:
            dzf = (dz @ W[i]) * (h[i] > 0)
            gb[i-1] = dzf.sum(0)
            dz = bf(dzf)
    print(f"batch {batch}: loss native {loss.item():.5f} fp32 {l32.item():.5f}")
    for i in range(3):
        for nm, ge in ((f"layers.{i}.weight", gw[i]), (f"layers.{i}.bias", gb[i])):
            e = model.layout.entries[nm]
            a = g_nat[e.offset:e.offset+e.numel]; r = g32[e.offset:e.offset+e.numel]; em = ge.reshape(-1)
            print(f"  {nm:18s} |nat-fp32|/|fp32| = {((a-r).norm()/r.norm()).item():.4f}   |nat-emul|/|emul| = {((a-em).norm()/em.norm()).item():.5f}   |emul-fp32|/|fp32| = {((em-r).norm()/r.norm()).item():.4f}")
