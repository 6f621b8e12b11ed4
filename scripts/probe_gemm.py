"""One short-K GEMM shape launched a few times (for ncu source-level captures of the store epilogue)."""
import sys
import torch
sys.path.insert(0, ".")
from depth_estimation_b200 import ops
M, K, N = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (262144, 64, 128)
dev = torch.device("cuda:0")
a = torch.randn((M, K), device=dev).bfloat16()
w = torch.randn((N, K), device=dev).bfloat16()
b = torch.randn((N,), device=dev)
out = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
for _ in range(6):
    ops.gemm(a, w, out, bias=b, act=ops.ACT_RELU6)
torch.cuda.synchronize()
