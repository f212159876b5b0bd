"""Phase timeline (ns, %globaltimer of CTA 0) of single small GEMM launches: where the fixed cost of a tiny tcgen05 GEMM goes."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200 import ops
from myfyp_b200.ops import native
dev = torch.device("cuda:0"); bf = torch.bfloat16
lib = native.load()
trace = torch.zeros(8, dtype=torch.int64, device=dev)
lib.b2_gemm_set_trace(ctypes.c_void_p(trace.data_ptr()))
names = ["entry", "prologue", "tile0_landed", "mma_issued", "acc_ready", "epilogue_done", "exit"]
def run(label, fn):
    rows = []
    for _ in range(6):
        trace.zero_(); torch.cuda.synchronize(); fn(); torch.cuda.synchronize()
        t = trace.tolist(); rows.append([t[i] - t[0] for i in range(7)])
    best = min(rows[1:], key=lambda r: r[6])
    print(f"{label:34s} " + "  ".join(f"{n}={v}" for n, v in zip(names[1:], best[1:])))
h = torch.randn(64, 256, device=dev).to(bf); w2 = torch.randn(10, 256, device=dev).to(bf); lo = torch.zeros(64, 10, device=dev); b2 = torch.zeros(10, device=dev)
run("fc2 fwd 64x10x256 (1 CTA)", lambda: ops.gemm_bf16(h, w2, 64, 10, 256, out_f32=lo, bias=b2))
x = torch.randn(64, 2048, device=dev).to(bf); w1 = torch.randn(256, 2048, device=dev).to(bf); hh = torch.zeros(64, 256, device=dev, dtype=bf); b1 = torch.zeros(256, device=dev)
run("fc1 fwd 64x256x2048 (4 CTAs)", lambda: ops.gemm_bf16(x, w1, 64, 256, 2048, out_bf16=hh, bias=b1, relu=True))
K, M, N = 16384, 64, 288
dz = torch.randn(K, M, device=dev).to(bf); col = torch.randn(K, N, device=dev).to(bf); out = torch.zeros(M, N, device=dev)
run("wgrad conv2 split16", lambda: ops.gemm_bf16(dz, col, M, N, K, out_f32=out, a_mn=True, b_mn=True, split_k=16, perm=(32, 9)))
y = torch.zeros(K, M, device=dev, dtype=bf); wk = torch.randn(M, N, device=dev).to(bf); bias = torch.zeros(M, device=dev)
run("conv2 fwd 16384x64x288", lambda: ops.gemm_bf16(col, wk, K, M, N, out_bf16=y, bias=bias, relu=True))
w = torch.randn(M, N, device=dev).to(bf); dcol = torch.zeros(K, N, device=dev, dtype=bf)
run("dcol conv2 16384x288x64", lambda: ops.gemm_bf16(dz, w, K, N, M, out_bf16=dcol, b_mn=True))
