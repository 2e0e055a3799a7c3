"""Generates the committed golden vectors from the REFERENCE'S OWN CODE.  Runs only in the build
container (needs /root/reference); the GPU box uses the .npz files.

    python tests/golden/make_golden.py

* flow_cosine.npz      — dataset_construction/step4_flow_check.py:cosine_similarity, imported as is.
* attn_processor.npz   — src/models/attention_processor.py: the reference's `Attention` module driven
  through `AttnProcessor` (eager) and `AttnProcessor2_0` (SDPA) in fp32 on CPU.  `diffusers` is not
  installed, so the four leaf imports at :21-24 are satisfied by a shim (USE_PEFT_BACKEND=True ->
  `linear_cls = nn.Linear`, reference :176-177; `maybe_allow_in_graph` = identity; logging/deprecate
  stubs).  No reference arithmetic is replaced.
* temporal_patched.npz — src/train_decouples.py:377-495 (the monkey-patched temporal forward) cannot be
  imported (it is a closure inside `main`), so its source lines are extracted with `ast`/text slicing
  from the reference file at generation time and exec'd against the reference's `Attention`.
* blocks.npz           — src/models/attention.py: the reference's own `FeedForward`, `BasicTransformerBlock` and
  `TemporalBasicTransformerBlock` classes (fp32, CPU).  Their leaf `GEGLU` lives in the un-vendored diffusers 0.24
  (`models/activations.py`); the shim below restates its published forward (`proj` -> chunk(2) -> hidden * F.gelu(gate))
  — every other line that runs (the block forwards, the FeedForward container, Attention + processor) is the reference's.
* spatio_temporal.npz  — src/models/transformer_temporal.py: the reference's own `TransformerSpatioTemporalModel` (forward
  :253-360) on its own blocks; the un-vendored diffusers leaves Timesteps / TimestepEmbedding / AlphaBlender are the
  restatements in timelapse_video_interpolation_b200/modules.py.  NOTE (circularity): for those three leaves this golden
  pins the package's OWN mirrors against themselves — it proves the reference's module code around them (GroupNorm,
  proj_in / proj_out, block order, `+ emb`, blend call, residual, layouts), not the leaves' arithmetic, which stays
  "restated from the published diffusers 0.24 source, unpinned" (DESIGN.md §2).
* step_glue.npz        — src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py:542-546 (CFG duplication,
  scale_model_input, concat with the image latents over dim 2) and :563-566 (per-frame guidance combine): those source
  lines are sliced out of the reference file at generation time and exec'd, the way temporal_patched.npz is made.  The one
  leaf they call, `scheduler.scale_model_input` (un-vendored diffusers EulerDiscreteScheduler), is the published
  `sample / sqrt(sigma^2 + 1)`; `scheduler.step` is NOT part of this golden (declared-unpinned leaf, oracle/step_ref.py).
* farneback.npz        — dataset_construction/step4_flow_check.py:compute_optical_flow_farneback imported as is, i.e. the
  output of cv2.cvtColor + cv2.calcOpticalFlowFarneback (third-party OpenCV, 4.13.0 here) on synthetic frames.
"""
from __future__ import annotations

import importlib.util
import logging as pylogging
import sys
import textwrap
import types
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent


def _shim_diffusers():
    d = types.ModuleType("diffusers")
    utils = types.ModuleType("diffusers.utils")
    utils.USE_PEFT_BACKEND = True
    utils.deprecate = lambda *a, **k: None
    lg = types.SimpleNamespace(get_logger=lambda name: pylogging.getLogger(name))
    utils.logging = lg
    iu = types.ModuleType("diffusers.utils.import_utils")
    iu.is_xformers_available = lambda: False
    tu = types.ModuleType("diffusers.utils.torch_utils")
    tu.maybe_allow_in_graph = lambda cls: cls
    models = types.ModuleType("diffusers.models")
    lora = types.ModuleType("diffusers.models.lora")

    class LoRACompatibleLinear(torch.nn.Linear):
        def forward(self, x, scale: float = 1.0):
            return super().forward(x)

    lora.LoRACompatibleLinear = LoRACompatibleLinear
    lora.LoRALinearLayer = type("LoRALinearLayer", (torch.nn.Module,), {})
    for name, mod in {"diffusers": d, "diffusers.utils": utils, "diffusers.utils.import_utils": iu,
                      "diffusers.utils.torch_utils": tu, "diffusers.models": models, "diffusers.models.lora": lora}.items():
        sys.modules[name] = mod


def _load(path: Path, name: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_flow():
    ref = _load(REF / "dataset_construction/step4_flow_check.py", "ref_step4")
    rng = np.random.default_rng(0)
    cases = {}
    for i, (h, w) in enumerate([(4, 6), (33, 47), (120, 160), (135, 240)]):
        a = rng.normal(0, 3, (h, w, 2)).astype(np.float32)
        b = (0.7 * a + rng.normal(0, 2, (h, w, 2))).astype(np.float32)
        if i == 1:  # zero vectors exercise the +1e-6 terms
            a[::3, ::5] = 0
            b[1::4, ::2] = 0
        cases[f"a{i}"], cases[f"b{i}"] = a, b
        cases[f"out{i}"] = np.asarray(ref.cosine_similarity(a, b), dtype=np.float32)
    # identical and opposite fields: known answers ~ +1 / -1
    a = rng.normal(0, 1, (16, 16, 2)).astype(np.float32)
    cases["a4"], cases["b4"], cases["out4"] = a, a.copy(), np.asarray(ref.cosine_similarity(a, a), dtype=np.float32)
    cases["a5"], cases["b5"], cases["out5"] = a, -a, np.asarray(ref.cosine_similarity(a, -a), dtype=np.float32)
    np.savez_compressed(OUT / "flow_cosine.npz", **cases)
    print("flow_cosine.npz", {k: float(v) for k, v in cases.items() if k.startswith("out")})


def _attn_cases():
    # (name, query_dim, cross_dim, heads, B', S, L)  — small SVD-shaped calls; d = 64 everywhere
    return [
        ("spatial_self", 128, None, 2, 2, 80, None),
        ("temporal_self", 128, None, 2, 6, 14, None),
        ("temporal_self_t25", 64, None, 1, 3, 25, None),
        ("cross_l1", 128, 96, 2, 3, 14, 1),
        ("cross_l2", 128, 96, 2, 3, 14, 2),
        ("spatial_cross_l2", 128, 96, 2, 2, 40, 2),
    ]


def make_attention():
    _shim_diffusers()
    ap = _load(REF / "src/models/attention_processor.py", "ref_attention_processor")
    out = {}
    for name, qd, cd, heads, B, S, L in _attn_cases():
        torch.manual_seed(hash(name) % 1000 + 1)
        g = torch.Generator().manual_seed(len(name) * 7 + 3)
        attn = ap.Attention(query_dim=qd, cross_attention_dim=cd, heads=heads, dim_head=64, processor=ap.AttnProcessor())
        for p in attn.parameters():
            p.data = torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5)
        x = torch.randn(B, S, qd, generator=g)
        e = torch.randn(B, L, cd, generator=g) if L else None
        with torch.no_grad():
            y_eager = attn(x, encoder_hidden_states=e)
            attn.set_processor(ap.AttnProcessor2_0())
            y_sdpa = attn(x, encoder_hidden_states=e)
        assert torch.allclose(y_eager, y_sdpa, atol=2e-5, rtol=1e-4), name
        out[f"{name}.x"] = x.numpy()
        if e is not None:
            out[f"{name}.e"] = e.numpy()
        out[f"{name}.wq"] = attn.to_q.weight.detach().numpy()
        out[f"{name}.wk"] = attn.to_k.weight.detach().numpy()
        out[f"{name}.wv"] = attn.to_v.weight.detach().numpy()
        out[f"{name}.wo"] = attn.to_out[0].weight.detach().numpy()
        out[f"{name}.bo"] = attn.to_out[0].bias.detach().numpy()
        out[f"{name}.heads"] = np.int64(heads)
        out[f"{name}.y"] = y_eager.numpy()
        out[f"{name}.y_sdpa"] = y_sdpa.numpy()
    np.savez_compressed(OUT / "attn_processor.npz", **out)
    print("attn_processor.npz", sorted({k.split(".")[0] for k in out}))
    return ap


def make_temporal_patched(ap):
    src = (REF / "src/train_decouples.py").read_text().splitlines()
    # lines 377..495 (1-based) hold `def ca_temporal_forward(...)` up to `return forward`
    start = next(i for i, l in enumerate(src) if "def ca_temporal_forward(self, controller, module_name)" in l)
    end = next(i for i in range(start, len(src)) if src[i].strip() == "return forward")
    code = textwrap.dedent("\n".join(src[start:end + 1]))
    ns = {"torch": torch, "Optional": __import__("typing").Optional, "USE_PEFT_BACKEND": True}
    exec(code, ns)  # defines ca_temporal_forward exactly as the reference wrote it
    ca_temporal_forward = ns["ca_temporal_forward"]

    class Controller:  # same recording interface as AttnController, src/train_decouples.py:518-545
        def __init__(self, use_attentionmask):
            self.use_attentionmask = use_attentionmask
            self.temporal_cross_repeat = 0
            self.start, self.end = {}, {}

        def set_startframe_attention(self, f, v):
            self.start[f] = v

        def set_endframe_attention(self, f, v):
            self.end[f] = v

    out = {}
    for name, use_mask, L in [("self", False, None), ("cross_l2", False, 2), ("cross_l2_mask", True, 2)]:
        g = torch.Generator().manual_seed(11 + len(name))
        attn = ap.Attention(query_dim=128, cross_attention_dim=96 if L else None, heads=2, dim_head=64)
        for p in attn.parameters():
            p.data = torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5)
        T = 14
        x = torch.randn(5, T, 128, generator=g)
        e = torch.randn(5, L, 96, generator=g) if L else None
        ctl = Controller(use_mask)
        fwd = ca_temporal_forward(attn, ctl, "temporal")
        with torch.no_grad():
            y = fwd(x, encoder_hidden_states=e)
        out[f"{name}.x"] = x.numpy()
        if e is not None:
            out[f"{name}.e"] = e.numpy()
            out[f"{name}.stats"] = np.array([[ctl.start[f], ctl.end[f]] for f in range(T)], dtype=np.float64)
        for k, lin in (("wq", attn.to_q), ("wk", attn.to_k), ("wv", attn.to_v), ("wo", attn.to_out[0])):
            out[f"{name}.{k}"] = lin.weight.detach().numpy()
        out[f"{name}.bo"] = attn.to_out[0].bias.detach().numpy()
        out[f"{name}.y"] = y.numpy()
    np.savez_compressed(OUT / "temporal_patched.npz", **out)
    print("temporal_patched.npz", sorted({k.split(".")[0] for k in out}))


def _shim_block_leaves(ap):
    """The leaf classes src/models/attention.py:6-12 imports from diffusers; only GEGLU is ever instantiated here."""
    act = types.ModuleType("diffusers.models.activations")

    class GEGLU(torch.nn.Module):  # diffusers 0.24 models/activations.py (restated; USE_PEFT_BACKEND -> nn.Linear)
        def __init__(self, dim_in, dim_out):
            super().__init__()
            self.proj = torch.nn.Linear(dim_in, dim_out * 2)

        def forward(self, hidden_states, scale: float = 1.0):
            hidden_states, gate = self.proj(hidden_states).chunk(2, dim=-1)
            return hidden_states * torch.nn.functional.gelu(gate)

    act.GEGLU = GEGLU
    act.GELU = act.ApproximateGELU = type("Unused", (torch.nn.Module,), {})
    emb = types.ModuleType("diffusers.models.embeddings")
    emb.SinusoidalPositionalEmbedding = type("SinusoidalPositionalEmbedding", (torch.nn.Module,), {})
    nrm = types.ModuleType("diffusers.models.normalization")
    nrm.AdaLayerNorm = nrm.AdaLayerNormZero = type("Unused", (torch.nn.Module,), {})
    pkg = types.ModuleType("models")
    pkg.attention_processor = ap
    for name, mod in {"diffusers.models.activations": act, "diffusers.models.embeddings": emb,
                      "diffusers.models.normalization": nrm, "models": pkg, "models.attention_processor": ap}.items():
        sys.modules[name] = mod


def make_blocks(ap):
    _shim_block_leaves(ap)
    ra = _load(REF / "src/models/attention.py", "ref_attention")
    out = {}

    def randomise(m, g):
        for name, p in m.named_parameters():
            if "norm" in name:  # affine LayerNorm parameters away from (1, 0) so that they are exercised
                p.data = (1.0 + 0.2 * torch.randn(p.shape, generator=g)) if name.endswith("weight") else 0.1 * torch.randn(p.shape, generator=g)
            else:
                p.data = torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.1)

    def save(tag, m, **tensors):
        for k, v in m.state_dict().items():
            out[f"{tag}.sd.{k}"] = v.numpy()
        for k, v in tensors.items():
            out[f"{tag}.{k}"] = v.numpy()

    g = torch.Generator().manual_seed(101)
    ff = ra.FeedForward(64, dim_out=32, activation_fn="geglu")
    randomise(ff, g)
    x = torch.randn(3, 10, 64, generator=g)
    with torch.no_grad():
        save("ff", ff, x=x, y=ff(x))

    C, heads, cross = 64, 1, 48
    sb = ra.BasicTransformerBlock(C, heads, 64, cross_attention_dim=cross)
    sb.attn1.set_processor(ap.AttnProcessor2_0())
    sb.attn2.set_processor(ap.AttnProcessor2_0())
    randomise(sb, g)
    x = torch.randn(4, 48, C, generator=g)          # [B*F, S, C]
    e = torch.randn(4, 2, cross, generator=g)
    with torch.no_grad():
        save("spatial", sb, x=x, e=e, y=sb(x, encoder_hidden_states=e))

    # time_mix_inner_dim != dim cannot run in the reference (:474 reshapes the output with the INPUT channel count)
    for tag, inner in (("temporal", C),):
        tb = ra.TemporalBasicTransformerBlock(C, inner, inner // 64, 64, cross_attention_dim=cross)
        tb.attn1.set_processor(ap.AttnProcessor2_0())
        tb.attn2.set_processor(ap.AttnProcessor2_0())
        randomise(tb, g)
        T, S, B = 7, 12, 2
        x = torch.randn(B * T, S, C, generator=g)
        e = torch.randn(B * S, 1, cross, generator=g)
        with torch.no_grad():
            save(tag, tb, x=x, e=e, y=tb(x, num_frames=T, encoder_hidden_states=e))
        out[f"{tag}.num_frames"] = np.int64(T)
    np.savez_compressed(OUT / "blocks.npz", **out)
    print("blocks.npz", sorted({k.split(".")[0] for k in out}), {k: tuple(v.shape) for k, v in out.items() if k.endswith(".y")})


def make_spatio_temporal(ap):
    """src/models/transformer_temporal.py:TransformerSpatioTemporalModel (the reference's own class and forward) on top of
    the reference's block classes; restated leaves: GEGLU (above), Timesteps / TimestepEmbedding (diffusers 0.24
    models/embeddings.py), AlphaBlender (models/resnet.py) — taken from the package's diffusers-free mirrors."""
    sys.path.insert(0, str(OUT.parents[1]))
    from timelapse_video_interpolation_b200 import modules as mirrors
    ra = sys.modules.get("ref_attention") or _load(REF / "src/models/attention.py", "ref_attention")
    cfgu = types.ModuleType("diffusers.configuration_utils")
    cfgu.ConfigMixin = type("ConfigMixin", (), {})
    cfgu.register_to_config = lambda f: f
    sys.modules["diffusers.utils"].BaseOutput = type("BaseOutput", (), {})
    datt = types.ModuleType("diffusers.models.attention")
    datt.BasicTransformerBlock, datt.TemporalBasicTransformerBlock = ra.BasicTransformerBlock, ra.TemporalBasicTransformerBlock
    emb = sys.modules["diffusers.models.embeddings"]
    emb.TimestepEmbedding, emb.Timesteps = mirrors.TimestepEmbedding, mirrors.Timesteps
    mu = types.ModuleType("diffusers.models.modeling_utils")
    mu.ModelMixin = torch.nn.Module
    rn = types.ModuleType("diffusers.models.resnet")
    rn.AlphaBlender = mirrors.AlphaBlender
    sys.modules["models"].attention = ra
    for name, mod in {"diffusers.configuration_utils": cfgu, "diffusers.models.attention": datt, "models.attention": ra,
                      "diffusers.models.modeling_utils": mu, "diffusers.models.resnet": rn}.items():
        sys.modules[name] = mod
    rt = _load(REF / "src/models/transformer_temporal.py", "ref_transformer_temporal")
    g = torch.Generator().manual_seed(77)
    C, heads, cross = 64, 1, 48
    m = rt.TransformerSpatioTemporalModel(heads, 64, in_channels=C, num_layers=1, cross_attention_dim=cross)
    for mod in m.modules():
        if isinstance(mod, ap.Attention):
            mod.set_processor(ap.AttnProcessor2_0())
    for name, p in m.named_parameters():
        if name.endswith("mix_factor"):
            p.data = torch.tensor([0.3])
        elif "norm" in name:
            p.data = (1.0 + 0.2 * torch.randn(p.shape, generator=g)) if name.endswith("weight") else 0.1 * torch.randn(p.shape, generator=g)
        else:
            p.data = torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.1)
    B, T, H, W = 2, 5, 4, 6
    x = torch.randn(B * T, C, H, W, generator=g)
    e = torch.randn(B * T, 1, cross, generator=g)
    ind = torch.zeros(B, T)
    ind[1, 3] = 1.0  # one image-only frame: alpha == 1 there
    out = {}
    with torch.no_grad():
        y = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
    for k, v in m.state_dict().items():
        out[f"sd.{k}"] = v.numpy()
    out.update(x=x.numpy(), e=e.numpy(), ind=ind.numpy(), y=y.numpy())
    np.savez_compressed(OUT / "spatio_temporal.npz", **out)
    print("spatio_temporal.npz", tuple(y.shape), float(y.abs().max()))


def make_step():
    """Pins oracle/step_ref.py:model_input_ref / cfg_ref to the reference's own loop-body lines."""
    src = (REF / "src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py").read_text().splitlines()

    def block(first_marker, n):
        i = next(k for k, l in enumerate(src) if first_marker in l)
        return i, "\n".join(src[i:i + n])

    i_in, _ = block("latent_model_input = torch.cat([latents] * 2) if do_classifier_free_guidance else latents", 1)
    i_cat = next(k for k in range(i_in, i_in + 8) if "latent_model_input = torch.cat([latent_model_input, image_latents], dim=2)" in src[k])
    code_in = textwrap.dedent("\n".join(l for l in src[i_in:i_cat + 1] if l.strip() and not l.strip().startswith("#")))
    i_cfg = next(k for k, l in enumerate(src) if "noise_pred_uncond, noise_pred_cond = noise_pred.chunk(2)" in l)
    code_cfg = textwrap.dedent("\n".join(src[i_cfg:i_cfg + 2]))
    assert "self.guidance_scale * (noise_pred_cond - noise_pred_uncond)" in code_cfg
    print("step glue lines:", i_in + 1, "-", i_cat + 1, "and", i_cfg + 1, "-", i_cfg + 2)

    class Scheduler:  # diffusers EulerDiscreteScheduler.scale_model_input (published formula; the only leaf these lines call)
        def __init__(self, sigma):
            self.sigma = sigma

        def scale_model_input(self, sample, t):
            return sample / ((self.sigma ** 2 + 1) ** 0.5)

    out = {}
    g = torch.Generator().manual_seed(31)
    for tag, B, T, C, H, W, sigma in (("a", 1, 14, 4, 9, 16, 7.5), ("b", 2, 5, 4, 6, 8, 0.31)):
        latents = torch.randn(B, T, C, H, W, generator=g) * sigma
        image_latents = torch.randn(2 * B, T, C, H, W, generator=g)
        noise_pred = torch.randn(2 * B, T, C, H, W, generator=g)
        # per-frame guidance (:518-530): linspace(min, max, T)[None] -> repeated per clip -> _append_dims to 5-D
        gs = torch.linspace(1.0, 3.0, T)[None].repeat(B, 1)[:, :, None, None, None]
        this = types.SimpleNamespace(scheduler=Scheduler(sigma), guidance_scale=gs)
        ns = {"torch": torch, "self": this, "latents": latents, "image_latents": image_latents, "t": None,
              "do_classifier_free_guidance": True, "noise_pred": noise_pred.clone()}
        exec(code_in, ns)
        exec(code_cfg, ns)
        out.update({f"{tag}.latents": latents.numpy(), f"{tag}.image_latents": image_latents.numpy(), f"{tag}.noise_pred": noise_pred.numpy(),
                    f"{tag}.guidance_scale": gs.numpy(), f"{tag}.sigma": np.float64(sigma),
                    f"{tag}.latent_model_input": ns["latent_model_input"].numpy(), f"{tag}.guided": ns["noise_pred"].numpy()})
    np.savez_compressed(OUT / "step_glue.npz", **out)
    print("step_glue.npz", {k: tuple(v.shape) for k, v in out.items() if k.endswith("latent_model_input") or k.endswith("guided")})


def make_farneback():
    """cv2.calcOpticalFlowFarneback itself (OpenCV 4.13 in this image) with the reference's parameters
    (dataset_construction/step4_flow_check.py:13-22) on synthetic textured BGR frames related by a small affine motion."""
    import cv2
    ref = _load(REF / "dataset_construction/step4_flow_check.py", "ref_step4b")
    out = {"opencv_version": np.array(cv2.__version__)}
    for i, (h, w) in enumerate([(96, 128), (135, 240), (70, 81), (40, 200), (6, 60), (40, 7)]):  # the last two: a side < 10 (border test wraps)
        r = np.random.default_rng(10 + i)
        base = cv2.resize(r.random((h // 8 + 2, w // 8 + 2, 3)).astype(np.float32), (w, h), interpolation=cv2.INTER_CUBIC)
        a = np.clip((base + 0.15 * r.random((h, w, 3)).astype(np.float32)) * 255, 0, 255).astype(np.uint8)
        A = np.float32([[1, 0.02, 1.7 - i], [-0.01, 1, -1.2 + 0.5 * i]])
        b = cv2.warpAffine(a, A, (w, h), borderMode=cv2.BORDER_REFLECT)
        out[f"a{i}"], out[f"b{i}"] = a, b
        out[f"gray{i}"] = cv2.cvtColor(a, cv2.COLOR_BGR2GRAY)
        out[f"flow{i}"] = ref.compute_optical_flow_farneback(a, b)      # the reference's own function
    np.savez_compressed(OUT / "farneback.npz", **out)
    print("farneback.npz", {k: tuple(v.shape) for k, v in out.items() if k.startswith("flow")})


if __name__ == "__main__":
    torch.set_num_threads(4)
    make_flow()
    ap = make_attention()
    make_temporal_patched(ap)
    make_blocks(ap)
    make_spatio_temporal(ap)
    make_step()
    make_farneback()
