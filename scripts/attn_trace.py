"""Debug: dump the clock64 trace of CTA 0 of attention_ws_kernel (DEB_ATTN_TRACE=1)."""
import os, sys, ctypes, re, io
os.environ["DEB_ATTN_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from depth_estimation_b200 import ops
dev = torch.device("cuda:0")
B, T, H = 32, 577, 16
qkv = torch.randn(B * T, 3 * H * 64, device=dev).bfloat16()
out = torch.empty(B * T, H * 64, device=dev, dtype=torch.bfloat16)
r, w = os.pipe(); old = os.dup(2); os.dup2(w, 2)
for _ in range(3):
    ops.attention(qkv, out, B, T, H, 0.125)
torch.cuda.synchronize()
os.dup2(old, 2); os.close(w)
txt = os.read(r, 4096).decode()
ptr = int(re.search(r"buffer (0x[0-9a-f]+)", txt).group(1), 16)
host = (ctypes.c_longlong * (2 * 64 * 16))()
cudart = ctypes.CDLL("/usr/local/cuda/lib64/libcudart.so")
cudart.cudaMemcpy(host, ctypes.c_void_p(ptr), ctypes.sizeof(host), 2)
import numpy as np
a = np.frombuffer(host, dtype=np.int64).reshape(2, 64, 16)
t0 = a[a > 0].min()
names = ["sm:pre_wait_S", "sm:S_ready", "sm:max_done", "sm:exp_done", "sm:P_arrived", "mma:pre_wait_P", "mma:P_ready", "mma:Sfree_seen", "mma:Snext_issued", "mma:PV_issued"]
for g in range(0, 12):
    for s in range(2):
        print(f"g={g} s={s} " + " ".join(f"{names[k].split(':')[1]}={a[s,g,k]-t0 if a[s,g,k] else -1}" for k in range(10)))
