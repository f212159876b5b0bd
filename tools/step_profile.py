"""One hand-written training step of the flagship model, eagerly (for `ncu --metrics gpu__time_duration.sum` launch lists),
plus fit timings: one node alone vs 8 nodes concurrently on their own streams (CUDA-graph epochs)."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200.learning.dataset import P2PFLDataset
from myfyp_b200.learning.torch_learner import TorchLearner
from myfyp_b200.models import cnn_model, mlp_model, resnet18_model

which = sys.argv[1] if len(sys.argv) > 1 else "cnn"
mode = sys.argv[2] if len(sys.argv) > 2 else "step"
dev = torch.device("cuda:0")
mk = {"cnn": cnn_model, "mlp": mlp_model, "resnet": resnet18_model}[which]
if mode == "step":
    model = mk(device="cuda:0", seed=0, native=True)
    module = model.get_model()
    x = torch.randn(64, *((28, 28) if which == "mlp" else (3, 32, 32)), device=dev)
    y = torch.randint(0, 10, (64,), device=dev)
    loss = torch.zeros(1, device=dev)
    from myfyp_b200 import ops
    n = model.layout.n_trainable
    m, v, st = torch.zeros(n, device=dev), torch.zeros(n, device=dev), torch.zeros(4, device=dev)
    module.native.refresh_weights()
    for it in range(5):
        module.native_train_step(x, y, loss)
        ops.adam_step(model.slab[:n], model.grad_slab[:n], m, v, 0, dev_state=st, step_offset=it + 1, p16=module.native.p16[:n])
    torch.cuda.synchronize()
    print("step ok", loss.item())
else:
    N = 8
    shape_fn = P2PFLDataset.synthetic_mnist if which == "mlp" else P2PFLDataset.synthetic_cifar10
    learners = [TorchLearner(mk(device="cuda:0", seed=i, native=True), shape_fn(6250, 256, seed=i), f"n{i}", batch_size=64) for i in range(N)]
    for l in learners:
        l.fit()  # capture
    torch.cuda.synchronize()
    t0 = time.perf_counter(); learners[0].fit(); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"one node alone: fit (98 steps) {1e3*(t1-t0):.1f} ms  -> {1e6*(t1-t0)/98:.0f} us/step")
    ths = [threading.Thread(target=l.fit) for l in learners]
    t0 = time.perf_counter()
    for t in ths: t.start()
    for t in ths: t.join()
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"8 nodes concurrently: {1e3*(t1-t0):.1f} ms -> {1e6*(t1-t0)/98/8:.0f} us/step/node")
