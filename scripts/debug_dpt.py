"""Stage-by-stage comparison of the CUDA DPT path against the CPU oracle (diagnostics, run on the GPU box)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from depth_estimation_b200 import ops
from depth_estimation_b200.depth_midas import DepthEstimator
from oracle.predict import build_model, synthetic_frame
from oracle import transforms as T

mt = sys.argv[1] if len(sys.argv) > 1 else "DPT_Large"
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (384, 384)
model = build_model(mt, 0)
est = DepthEstimator(mt, weights=model.state_dict(), use_cuda_graph=False)
net = est.estimator
frame = synthetic_frame(W, H, 3, "S")
x_ref = torch.from_numpy(T.midas_transform_bgr(frame, mt))
def rel(a, b):
    a = a.float().cpu(); b = b.float().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))
with torch.no_grad():
    nw, nh = ops.get_size(W, H, mt)
    gh, gw = nh // 16, nw // 16
    dev = est.device
    fr = torch.from_numpy(frame[None]).to(dev)
    patches = ops.preprocess(fr, nh, nw, (0.5,)*3, (0.5,)*3, net.preprocess_layout)
    feats_holder = {}
    orig = net.dec.forward
    def cap(feats, B):
        feats_holder["f"] = [f.clone() for f in feats]
        return orig(feats, B)
    net.dec.forward = cap
    pred = net.forward(patches, 1, nh, nw)
    torch.cuda.synchronize()
    acts, _ = model.pretrained.model.forward_flex(x_ref, set(model.hooks))
    xfin = list(net._x.values())[0]
    print("tokens after last block:", rel(xfin.view(1, -1, net.dim), acts[max(model.hooks)]))
    ls = model.forward_vit(x_ref)
    for i, (f, l) in enumerate(zip(feats_holder["f"], ls)):
        print(f"reassemble {i+1}:", tuple(f.shape), rel(f.permute(0, 3, 1, 2), l))
    ref = model.decode(*ls)
    print("pred:", rel(pred, ref), "range", float(ref.min()), float(ref.max()),
          "dyn", float((pred.cpu() - ref).abs().max() / (ref.max() - ref.min())))
