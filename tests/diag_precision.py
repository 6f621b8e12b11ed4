"""Developer diagnostic (not collected by pytest; CPU only): which parts of a DPT model must run above plain bf16
operand precision for the final map to stay within north_star's 1e-3 of the fp32 result?

bf16 operand rounding (inputs + weights of every Linear / Conv / ConvTranspose whose module name matches a filter) is
injected into the fp32 oracle, group by group, and the distance of the final map from the all-fp32 run is printed as
max|y - y*| / max|y*| (`rel`) and per pixel where y* > 1e-2 max (`pix`).

    python tests/diag_precision.py DPT_Hybrid [wide]      # `wide`: HF-style wide-range maps (positive final filter)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from oracle import transforms as T
from oracle.predict import build_model, synthetic_frame

GROUPS = {
    "backbone": ("pretrained.model.patch_embed",),
    "vit": ("pretrained.model.blocks",),
    "reassemble": ("pretrained.act_postprocess",),
    "layer_rn": ("scratch.layer",),
    "refinenet4": ("scratch.refinenet4",),
    "refinenet3": ("scratch.refinenet3",),
    "refinenet2": ("scratch.refinenet2",),
    "refinenet1": ("scratch.refinenet1",),
    "head0": ("scratch.output_conv.0",),
    "head2": ("scratch.output_conv.2",),
    "head4": ("scratch.output_conv.4",),
}
HF_GROUPS = {
    "backbone": ("hf.dpt.embeddings",),
    "vit": ("hf.dpt.encoder",),
    "reassemble": ("hf.neck.reassemble_stage",),
    "layer_rn": ("hf.neck.convs",),
    "refinenet4": ("hf.neck.fusion_stage.layers.0",),
    "refinenet3": ("hf.neck.fusion_stage.layers.1",),
    "refinenet2": ("hf.neck.fusion_stage.layers.2",),
    "refinenet1": ("hf.neck.fusion_stage.layers.3",),
    "head0": ("hf.head.head.0",),
    "head2": ("hf.head.head.2",),
    "head4": ("hf.head.head.4",),
}


def rb(t):
    return t.bfloat16().float()


class Inject:
    def __init__(self, model, prefixes, attention=False):
        self.model, self.prefixes, self.attention = model, tuple(prefixes), attention
        self.hooks, self.saved = [], {}

    def __enter__(self):
        for name, m in self.model.named_modules():
            if isinstance(m, (torch.nn.Linear, torch.nn.Conv2d, torch.nn.ConvTranspose2d)) and name.startswith(self.prefixes):
                self.saved[name] = m.weight.data
                m.weight.data = rb(m.weight.data)
                self.hooks.append(m.register_forward_pre_hook(lambda mod, inp: (rb(inp[0]),) + tuple(inp[1:])))
        if self.attention:
            self._mm = torch.Tensor.__matmul__
            mm = self._mm
            torch.Tensor.__matmul__ = lambda a, b: mm(rb(a), rb(b))
            self._matmul = torch.matmul
            mm2 = self._matmul
            torch.matmul = lambda a, b, *k, **kw: mm2(rb(a), rb(b), *k, **kw)
        return self

    def __exit__(self, *exc):
        if self.attention:
            torch.Tensor.__matmul__ = self._mm
            torch.matmul = self._matmul
        for h in self.hooks:
            h.remove()
        mods = dict(self.model.named_modules())
        for name, w in self.saved.items():
            mods[name].weight.data = w
        return False


def errs(y, ref):
    d = (y - ref).abs()
    rel = float(d.max() / ref.abs().max())
    m = ref > 1e-2 * ref.max()
    pix = float((d[m] / ref[m]).max())
    dyn = float(d.max() / (ref.max() - ref.min()))
    return rel, pix, dyn


def main():
    mt = sys.argv[1] if len(sys.argv) > 1 else "DPT_Hybrid"
    hf_mode = len(sys.argv) > 2 and sys.argv[2] == "hf"
    side = 384
    frame = synthetic_frame(side, side, 7 if hf_mode else 3, "S")
    x = torch.from_numpy(np.asarray(T.midas_transform_bgr(frame, mt))).reshape(1, 3, side, side).float()
    groups = GROUPS
    if hf_mode:                      # the wide-range maps of tests/test_oracle_hf_crosscheck.py::test_product_matches_hf
        from tests.test_oracle_hf_crosscheck import _hf, _HFNet
        hf, _ = _hf(mt, 1 if mt == "DPT_Hybrid" else 0)
        model = _HFNet(hf).eval()
        groups = HF_GROUPS
    else:
        model = build_model(mt, seed=0).eval()
    with torch.no_grad():
        ref = model(x)
    print(f"{mt} hf={hf_mode}: ref max {float(ref.max()):.3f} min {float(ref.min()):.3f}")
    allp = tuple(p for g in groups.values() for p in g)

    def run(label, names, attention=False):
        prefixes = tuple(p for g in names for p in groups[g])
        with torch.no_grad(), Inject(model, prefixes, attention):
            y = model(x)
        rel, pix, dyn = errs(y, ref)
        print(f"{label:40s} rel {rel:.2e}  pix {pix:.2e}  dyn {dyn:.2e}", flush=True)

    G = list(groups)
    run("all bf16 (+attention)", G, True)
    run("CUDA-path model: all but head4", [g for g in G if g != "head4"], True)
    run("attention matmuls only", (), True)
    for g in G:
        run("only " + g, [g])
    dec = ("layer_rn", "refinenet4", "refinenet3", "refinenet2", "refinenet1", "head0", "head2")
    run("decoder (rn+refinenets+head0+head2)", dec)
    run("vit + attention + reassemble", ("vit", "reassemble"), True)
    nb = [g for g in G if g not in ("backbone", "head4")]
    run("all but backbone/head4", nb, True)
    run("all but backbone/head4, head0+2", [g for g in nb if g not in ("head0", "head2")], True)
    run("all but backbone/head4, head0+2, rn1", [g for g in nb if g not in ("head0", "head2", "refinenet1")], True)
    run("all but backbone/head4, decoder", [g for g in nb if g not in dec], True)


if __name__ == "__main__":
    main()
