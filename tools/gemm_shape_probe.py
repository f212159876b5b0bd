"""Time single GEMM launches (warm, back-to-back average) for the small shapes of the CNN step with different knobs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200 import ops
dev = torch.device("cuda:0")
bf = torch.bfloat16
def timed(fn, iters=50):
    """GPU time per launch inside a CUDA graph of `iters` back-to-back (dependent) launches - no CPU launch cost."""
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(iters): fn()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): g.replay()
        e1.record(); torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / (5 * iters)
def empty():
    pass
# wgrad conv2: dW[64, 288] = dz[16384,64]^T col[16384,288]
K, M, N = 16384, 64, 288
dz = torch.randn(K, M, device=dev).to(bf); col = torch.randn(K, N, device=dev).to(bf); out = torch.zeros(M, N, device=dev)
for split in (1, 4, 16, 32, 64):
    print(f"wgrad conv2 split={split:3d}: {timed(lambda: ops.gemm_bf16(dz, col, M, N, K, out_f32=out, a_mn=True, b_mn=True, split_k=split)):.1f} us   perm: "
          f"{timed(lambda: ops.gemm_bf16(dz, col, M, N, K, out_f32=out, a_mn=True, b_mn=True, split_k=split, perm=(32, 9))):.1f} us")
# same product with K-major operands (pre-transposed copies) to see what MN-major costs
dzT = dz.t().contiguous(); colT = col.t().contiguous()
for split in (1, 16, 64):
    print(f"wgrad conv2 K-major split={split:3d}: {timed(lambda: ops.gemm_bf16(dzT, colT, M, N, K, out_f32=out, split_k=split)):.1f} us")
# dcol conv2: dcol[16384, 288] = dz[16384,64] W[64,288]
w = torch.randn(M, N, device=dev).to(bf); dcol = torch.zeros(K, N, device=dev, dtype=bf)
print(f"dcol conv2: {timed(lambda: ops.gemm_bf16(dz, w, K, N, M, out_bf16=dcol, b_mn=True)):.1f} us")
# conv2 fwd: y[16384,64] = col[16384,288] W^T
y = torch.zeros(K, M, device=dev, dtype=bf); wk = torch.randn(M, N, device=dev).to(bf); bias = torch.zeros(M, device=dev)
print(f"conv2 fwd: {timed(lambda: ops.gemm_bf16(col, wk, K, M, N, out_bf16=y, bias=bias, relu=True)):.1f} us")
# fc1 fwd: h[64,256] = x[64,2048] W[256,2048]^T
x = torch.randn(64, 2048, device=dev).to(bf); w1 = torch.randn(256, 2048, device=dev).to(bf); h = torch.zeros(64, 256, device=dev, dtype=bf); b1 = torch.zeros(256, device=dev)
print(f"fc1 fwd: {timed(lambda: ops.gemm_bf16(x, w1, 64, 256, 2048, out_bf16=h, bias=b1, relu=True)):.1f} us")
# fc2 fwd: [64,10] = h[64,256] W[10,256]^T
w2 = torch.randn(10, 256, device=dev).to(bf); lo = torch.zeros(64, 10, device=dev); b2 = torch.zeros(10, device=dev)
print(f"fc2 fwd: {timed(lambda: ops.gemm_bf16(h, w2, 64, 10, 256, out_f32=lo, bias=b2)):.1f} us")
# launch floor: a trivial kernel through the same binding path
t = torch.zeros(4, device=dev)
print(f"adam_advance (1 thread kernel): {timed(lambda: ops.adam_advance(t, 0.0)):.1f} us")
