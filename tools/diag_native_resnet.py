import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200.models import resnet18_model
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
model = resnet18_model(device="cuda:0", seed=1, native=True); module = model.get_model()
B = 32
x = torch.randn(B, 3, 32, 32, device=dev); y = torch.randint(0, 10, (B,), device=dev)
s0 = model.slab.clone(); loss = torch.zeros(1, device=dev)
module.native.refresh_weights(); module.native_train_step(x, y, loss)
g = model.grad_slab.clone(); model.grad_slab.zero_(); model.slab.copy_(s0)
module.train(); l = torch.nn.functional.cross_entropy(module(x), y); l.backward(); r = model.grad_slab.clone()
print("loss", loss.item(), l.item())
for e in model.layout.in_order():
    if not e.trainable: continue
    a, b = g[e.offset:e.offset+e.numel], r[e.offset:e.offset+e.numel]
    print(f"{e.name:34s} cos {torch.nn.functional.cosine_similarity(a,b,dim=0).item():.4f}  |nat|/|ref| {(a.norm()/(b.norm()+1e-12)).item():.3f}")
model.grad_slab.zero_()
with torch.autocast("cuda", dtype=torch.bfloat16):
    l2 = torch.nn.functional.cross_entropy(module(x).float(), y)
l2.backward(); ac = model.grad_slab.clone()
cs = torch.nn.functional.cosine_similarity
print("overall: native-vs-fp32", cs(g, r, dim=0).item(), "autocast-vs-fp32", cs(ac, r, dim=0).item(), "native-vs-autocast", cs(g, ac, dim=0).item())
for name in ("conv1.weight", "layers.3.conv1.weight", "layers.7.conv1.weight"):
    e = [q for q in model.layout.in_order() if q.name == name][0]
    sl = slice(e.offset, e.offset + e.numel)
    print(name, "autocast-vs-fp32", cs(ac[sl], r[sl], dim=0).item())
