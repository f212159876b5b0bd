"""ResNet-18 (CIFAR) training step, batch 64: the hand-written path (eager launches and as one CUDA graph) against the same
nn.Module through torch autograd / cuDNN (fp32 with TF32 convs, and bf16 autocast channels_last).  CUDA-event timings."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200 import ops
from myfyp_b200.models import resnet18_model
from myfyp_b200.ops import native

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
x = torch.randn(B, 3, 32, 32, device=dev); y = torch.randint(0, 10, (B,), device=dev)

def timed(fn, iters=30, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

out = {"batch": B}
model = resnet18_model(device="cuda:0", seed=0, native=True); module = model.get_model()
n = model.layout.n_trainable
m, v, st = torch.zeros(n, device=dev), torch.zeros(n, device=dev), torch.zeros(4, device=dev)
loss = torch.zeros(1, device=dev)
module.native.refresh_weights()
def native_step():
    module.native_train_step(x, y, loss)
    ops.adam_step(model.slab[:n], model.grad_slab[:n], m, v, 0, dev_state=st, step_offset=1, p16=module.native.p16[:n])
native_step(); torch.cuda.synchronize()
native.reset_launches(); native_step(); out["native_launches_per_step"] = native.launches()
out["native_eager_ms"] = round(timed(native_step), 3)
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    native_step(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        native_step()
out["native_graph_ms"] = round(timed(g.replay), 3)

ref = resnet18_model(device="cuda:0", seed=0, native=False).get_model().train()
opt = torch.optim.Adam(ref.parameters(), lr=1e-3)
def torch_step():
    opt.zero_grad(set_to_none=True)
    torch.nn.functional.cross_entropy(ref(x), y).backward(); opt.step()
out["torch_fp32_tf32_ms"] = round(timed(torch_step), 3)
ref = ref.to(memory_format=torch.channels_last); xc = x.contiguous(memory_format=torch.channels_last)
def torch_bf16_step():
    opt.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        l = torch.nn.functional.cross_entropy(ref(xc).float(), y)
    l.backward(); opt.step()
out["torch_bf16_autocast_ms"] = round(timed(torch_bf16_step), 3)
print(json.dumps(out))
