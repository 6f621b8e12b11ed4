"""TEST INFRASTRUCTURE.  An independent second implementation of the DPT models as a parity pin.

The reference (/root/reference/depth_midas.py:10) loads ``DPTDepthModel`` from intel-isl/MiDaS over torch.hub;
that code is not vendored and cannot be fetched here.  The HF ``transformers`` package in
this image carries an independently written ``DPTForDepthEstimation`` whose weights are converted from the very
same MiDaS checkpoints (``convert_dpt_to_pytorch.py`` renames upstream keys to HF keys).  Here the renaming
is inverted: a randomly initialised HF model's state_dict is rewritten into upstream MiDaS key names, so the same
numbers can be fed to (a) the HF model, (b) the oracle restatement and (c) the CUDA product path.

Nothing in this file is imported by the product package.
"""
from __future__ import annotations

import torch


def hf_large_config(**over):
    from transformers import DPTConfig
    kw = dict(hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, intermediate_size=4096,
              image_size=384, patch_size=16, backbone_out_indices=[5, 11, 17, 23],
              neck_hidden_sizes=[256, 512, 1024, 1024], fusion_hidden_size=256, reassemble_factors=[4, 2, 1, 0.5],
              readout_type="project", hidden_act="gelu", layer_norm_eps=1e-6, head_in_index=-1,
              use_batch_norm_in_fusion_residual=False, add_projection=False, qkv_bias=True,
              hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, is_hybrid=False)
    kw.update(over)
    return DPTConfig(**kw)


def hf_hybrid_config(**over):
    from transformers import DPTConfig
    kw = dict(hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072,
              image_size=384, patch_size=16, backbone_out_indices=[2, 5, 8, 11],
              neck_hidden_sizes=[256, 512, 768, 768], fusion_hidden_size=256, reassemble_factors=[1, 1, 1, 0.5],
              neck_ignore_stages=[0, 1], backbone_featmap_shape=[1, 1024, 24, 24],
              readout_type="project", hidden_act="gelu", layer_norm_eps=1e-6, head_in_index=-1,
              use_batch_norm_in_fusion_residual=False, add_projection=False, qkv_bias=True,
              hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, is_hybrid=True)
    kw.update(over)
    return DPTConfig(**kw)


def randomize_(model, seed=0, positive_head_bias=True):
    """Every parameter gets a non-trivial seeded value (HF leaves cls/pos-embed/biases at zero), drawn from the same
    distribution as ``depth_estimation_b200.init.random_state_dict`` - the init every other parity case runs on:
    Linear 0.02, conv 1/sqrt(fan_in), norm gains 1 +- 0.1, biases / pos-embed 0.02 - with the head biases of SURVEY 8d
    D2's degeneracy guard (the final ReLU must not zero the map).  An attenuating init (all gains 0.7/sqrt(fan_in),
    the first version of this file) shrinks the signal by 0.7 per decoder conv until the map is bias-dominated
    (max / dynamic range = 6); bf16 activation rounding (2^-9 of the MAGNITUDE) then reads as 2e-3 of the maximum and
    1.3e-2 of the range, for the fp32-accumulating CUDA path and for the oracle with bf16 rounding injected alike - a
    property of that init, not of either implementation."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "embeddings.position" in name or "cls_token" in name:
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
            elif p.ndim == 2:
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
            elif p.ndim > 2:
                fan_in = p[0].numel()
                p.copy_(torch.randn(p.shape, generator=g) * (1.0 / fan_in ** 0.5))
            elif name.endswith("weight"):                                  # norm gains
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            else:                                                          # biases
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
        if positive_head_bias:
            sd = dict(model.named_parameters())
            sd["head.head.4.bias"].fill_(1.0)
            sd["head.head.2.bias"].fill_(0.5)
            sd["head.head.4.weight"].abs_()          # 32 post-ReLU inputs x positive weights: the map cannot clip to 0
    return model


class bf16_operands:
    """Context manager: ``model`` computes with bf16-rounded operands (inputs and weights of every Linear / Conv /
    ConvTranspose, both sides of every explicit matmul) and fp32 accumulation - the arithmetic contract of the CUDA
    path (north_star: "bf16 ... fp32 accumulate") applied to an independent implementation.  Its distance from the
    fp32 run of the same model is the precision floor of that contract for the given weights and input."""

    def __init__(self, model):
        self.model, self.hooks, self.saved = model, [], {}

    @staticmethod
    def _rb(t):
        return t.bfloat16().float() if torch.is_floating_point(t) else t

    def __enter__(self):
        rb = self._rb
        for name, m in self.model.named_modules():
            if isinstance(m, (torch.nn.Linear, torch.nn.Conv2d, torch.nn.ConvTranspose2d)):
                self.saved[name] = m.weight.data
                m.weight.data = rb(m.weight.data)
                self.hooks.append(m.register_forward_pre_hook(lambda mod, inp: (rb(inp[0]),) + tuple(inp[1:])))
        self._matmul = torch.matmul
        torch.matmul = lambda a, b, *k, **kw: self._matmul(rb(a), rb(b), *k, **kw)
        return self.model

    def __exit__(self, *exc):
        torch.matmul = self._matmul
        for h in self.hooks:
            h.remove()
        mods = dict(self.model.named_modules())
        for name, w in self.saved.items():
            mods[name].weight.data = w
        return False


_BLOCK = {
    "layernorm_before": "norm1", "layernorm_after": "norm2", "attention.output.dense": "attn.proj",
    "intermediate.dense": "mlp.fc1", "output.dense": "mlp.fc2",
}


def hf_to_upstream(hf_sd: dict, hybrid: bool = False) -> dict:
    """HF DPTForDepthEstimation state_dict -> intel-isl/MiDaS ``DPTDepthModel`` state_dict (upstream key names;
    inverse of HF's conversion script's renames, q/k/v re-fused into ``attn.qkv``)."""
    out = {}
    nlayers = 1 + max(int(k.split(".")[3]) for k in hf_sd if k.startswith("dpt.encoder.layer."))
    pm = "pretrained.model."
    out[pm + "cls_token"] = hf_sd["dpt.embeddings.cls_token"]
    out[pm + "pos_embed"] = hf_sd["dpt.embeddings.position_embeddings"]
    for wb in ("weight", "bias"):
        pe = "dpt.embeddings.projection." if hybrid else "dpt.embeddings.patch_embeddings.projection."
        out[pm + "patch_embed.proj." + wb] = hf_sd[pe + wb]
        if "dpt.layernorm." + wb in hf_sd:
            out[pm + "norm." + wb] = hf_sd["dpt.layernorm." + wb]
    for i in range(nlayers):
        src = f"dpt.encoder.layer.{i}."
        dst = f"{pm}blocks.{i}."
        for wb in ("weight", "bias"):
            out[dst + "attn.qkv." + wb] = torch.cat(
                [hf_sd[f"{src}attention.attention.{n}.{wb}"] for n in ("query", "key", "value")], 0)
            for a, b in _BLOCK.items():
                out[dst + b + "." + wb] = hf_sd[src + a + "." + wb]
    # reassemble: HF index j -> act_postprocess{j+1}
    for j in range(4):
        ap = f"pretrained.act_postprocess{j + 1}."
        for wb in ("weight", "bias"):
            k = f"neck.reassemble_stage.readout_projects.{j}.0.{wb}"
            if k in hf_sd:
                out[ap + "0.project.0." + wb] = hf_sd[k]
            k = f"neck.reassemble_stage.layers.{j}.projection.{wb}"
            if k in hf_sd:
                out[ap + "3." + wb] = hf_sd[k]
            k = f"neck.reassemble_stage.layers.{j}.resize.{wb}"
            if k in hf_sd:
                out[ap + "4." + wb] = hf_sd[k]
        out[f"scratch.layer{j + 1}_rn.weight"] = hf_sd[f"neck.convs.{j}.weight"]
        rn = f"scratch.refinenet{4 - j}."                       # HF fusion layer 0 is refinenet4
        fl = f"neck.fusion_stage.layers.{j}."
        for wb in ("weight", "bias"):
            out[rn + "out_conv." + wb] = hf_sd[fl + "projection." + wb]
            for u in (1, 2):
                for c in (1, 2):
                    out[f"{rn}resConfUnit{u}.conv{c}.{wb}"] = hf_sd[f"{fl}residual_layer{u}.convolution{c}.{wb}"]
    for i in (0, 2, 4):
        for wb in ("weight", "bias"):
            out[f"scratch.output_conv.{i}.{wb}"] = hf_sd[f"head.head.{i}.{wb}"]
    if hybrid:
        bb = "dpt.embeddings.backbone.bit."
        ub = pm + "patch_embed.backbone."
        out[ub + "stem.conv.weight"] = hf_sd[bb + "embedder.convolution.weight"]
        for wb in ("weight", "bias"):
            out[ub + "stem.norm." + wb] = hf_sd[bb + "embedder.norm." + wb]
        for k, v in hf_sd.items():
            if not k.startswith(bb + "encoder.stages."):
                continue
            parts = k[len(bb + "encoder.stages."):].split(".")          # s, layers, b, <module...>
            s, b, rest = parts[0], parts[2], ".".join(parts[3:])
            out[f"{ub}stages.{s}.blocks.{b}.{rest}"] = v
    return {k: v.detach().clone() for k, v in out.items()}


def load_into_oracle(oracle_model, upstream_sd: dict):
    """strict on everything the HF model carries; the oracle's extra (unused) tensors keep their init."""
    own = oracle_model.state_dict()
    missing = [k for k in upstream_sd if k not in own]
    assert not missing, f"keys unknown to the oracle: {missing[:5]}"
    for k, v in upstream_sd.items():
        assert own[k].shape == v.shape, (k, tuple(own[k].shape), tuple(v.shape))
    oracle_model.load_state_dict(upstream_sd, strict=False)
    return [k for k in own if k not in upstream_sd]


def complete_for_product(upstream_sd: dict, model_type: str) -> dict:
    """The product's checkpoint loader validates the full upstream key inventory; add the tensors a real MiDaS
    checkpoint has but HF drops (unused 1000-way head, refinenet4.resConfUnit1, final norm) as zeros."""
    from depth_estimation_b200.init import param_shapes
    sd = dict(upstream_sd)
    for k, shp in param_shapes(model_type).items():
        if k not in sd:
            sd[k] = torch.zeros(shp)
    return sd
