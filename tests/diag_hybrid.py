"""Developer diagnostic (not collected by pytest): where does DPT_Hybrid's bf16 path drift from the fp32 oracle?
Runs both on the same bf16-rounded network input and prints max|a-b|/max|b| per stage.  GPU box only:
    python tests/diag_hybrid.py
Lives under tests/ because it imports the oracle (test infrastructure)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from depth_estimation_b200.depth_midas import DepthEstimator
from oracle import transforms as T
from oracle.predict import build_model, synthetic_frame


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / b.abs().max()), float((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt())


def main():
    mt = "DPT_Hybrid"
    model = build_model(mt, seed=0).eval()
    est = DepthEstimator(mt, weights=model.state_dict())
    net = est.estimator
    frame = synthetic_frame(384, 384, 3, "S")
    x = torch.from_numpy(np.asarray(T.midas_transform_bgr(frame, mt))).reshape(1, 3, 384, 384).float()
    xb = x                                        # K1 feeds the hybrid stem in fp32
    ref = {}
    vit = model.pretrained.model
    bb = vit.patch_embed.backbone
    with torch.no_grad():
        xin = xb
        h = bb.stem(xin)
        ref["stem+pool"] = h
        for si in range(3):
            for bi, blk in enumerate(bb.stages[si].blocks):
                h = blk(h)
                ref[f"s{si}b{bi}"] = h
        acts, _ = vit.forward_flex(xin, {8, 11})
        ref["tok8"], ref["tok11"] = acts[8], acts[11]
        l1, l2, l3, l4 = model.forward_vit(xin)
        ref["l3"], ref["l4"] = l3, l4
        ref["out"] = model(xin)
    got = {}
    orig_b = net._bottleneck

    def bneck(tag, blk, x_, x32, B, H, W, **kw):
        r = orig_b(tag, blk, x_, x32, B, H, W, **kw)
        got[f"s{tag[0]}b{tag[1]}"] = r[1].clone().permute(0, 3, 1, 2)
        return r
    net._bottleneck = bneck
    for j, r in enumerate(net.reasm):
        def fwd(xs, B, gh, gw, _r=r, _j=j, _o=r.forward):
            got["tok8" if _j == 0 else "tok11"] = xs.clone().reshape(1, -1, 768)
            o = _o(xs, B, gh, gw)
            got["l3" if _j == 0 else "l4"] = o.clone().permute(0, 3, 1, 2)
            return o
        r.forward = fwd
    out = net.forward(xb.cuda().contiguous(), 1, 384, 384)
    torch.cuda.synchronize()
    got["out"] = out.clone()
    for k in ref:
        if k in got:
            m, r2 = rel(got[k], ref[k])
            print(f"{k:10s} max-rel {m:.3e}   rms-rel {r2:.3e}   ref absmax {float(ref[k].abs().max()):.3f}")


if __name__ == "__main__":
    main()
