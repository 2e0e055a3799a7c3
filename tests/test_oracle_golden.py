"""CPU: the oracle restatement reproduces the golden vectors generated from the reference's own code
(tests/golden/make_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest
import torch

from oracle import attention_ref, flow_ref, slerp_ref
from util import attn_from_golden, max_abs


def test_flow_cosine_matches_reference_function(golden_dir):
    z = np.load(golden_dir / "flow_cosine.npz")
    for i in range(6):
        got = flow_ref.cosine_similarity(z[f"a{i}"], z[f"b{i}"])
        assert got.dtype == np.float32
        assert got == z[f"out{i}"], (i, got, z[f"out{i}"])  # same numpy ops -> bit-identical


ATTN_CASES = ["spatial_self", "temporal_self", "temporal_self_t25", "cross_l1", "cross_l2", "spatial_cross_l2"]


@pytest.mark.parametrize("name", ATTN_CASES)
def test_attention_oracle_matches_reference_processors(golden_dir, name):
    z = np.load(golden_dir / "attn_processor.npz")
    attn = attn_from_golden(z, name)
    x = torch.from_numpy(z[f"{name}.x"])
    e = torch.from_numpy(z[f"{name}.e"]) if f"{name}.e" in z else None
    y = attention_ref.attn_processor_ref(attn, x, e)
    assert y.shape == x.shape
    # fp32 on both sides; only the summation order inside bmm/softmax may differ
    assert max_abs(y, z[f"{name}.y"]) < 2e-5
    assert max_abs(y, z[f"{name}.y_sdpa"]) < 2e-5
    y64 = attention_ref.attn_processor_ref(attn, x, e, dtype=torch.float64)
    assert max_abs(y64, z[f"{name}.y"]) < 2e-5


@pytest.mark.parametrize("name,mask", [("self", False), ("cross_l2", False), ("cross_l2_mask", True)])
def test_patched_temporal_forward_oracle(golden_dir, name, mask):
    z = np.load(golden_dir / "temporal_patched.npz")
    attn = attn_from_golden(z, name)
    x = torch.from_numpy(z[f"{name}.x"])
    e = torch.from_numpy(z[f"{name}.e"]) if f"{name}.e" in z else None
    y, stats = attention_ref.ca_temporal_forward_ref(attn, x, e, use_attentionmask=mask)
    assert max_abs(y, z[f"{name}.y"]) < 2e-5
    if e is None:
        assert stats is None
    else:
        assert max_abs(stats, z[f"{name}.stats"]) < 2e-6
    if not mask:  # without the ramp bias the patched forward equals the stock processor
        assert max_abs(y, attention_ref.attn_processor_ref(attn, x, e)) < 2e-5


def test_one_key_cross_attention_is_query_independent():
    # with one context token softmax == 1 and the output ignores Q
    from util import make_attn
    attn = make_attn(128, 96, 2, seed=5)
    g = torch.Generator().manual_seed(0)
    e = torch.randn(3, 1, 96, generator=g)
    y1 = attention_ref.attn_processor_ref(attn, torch.randn(3, 14, 128, generator=g), e)
    y2 = attention_ref.attn_processor_ref(attn, torch.randn(3, 14, 128, generator=g), e)
    assert torch.equal(y1, y2)
    assert torch.equal(y1[:, 0], y1[:, 5])


# ---------------------------------------------------------------- slerp spec (known answers)
def _rand(shape, seed):
    return np.random.default_rng(seed).standard_normal(shape).astype(np.float32)


def test_slerp_endpoints_and_shape():
    v0, v1 = _rand((3, 4, 8, 16), 0), _rand((3, 4, 8, 16), 1)
    out = slerp_ref.slerp_ref(v0, v1, 14)
    assert out.shape == (3, 14, 4, 8, 16) and out.dtype == np.float32
    np.testing.assert_allclose(out[:, 0], v0, rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(out[:, -1], v1, rtol=1e-6, atol=1e-6)


def test_slerp_identical_inputs_take_lerp_branch():
    v0 = _rand((2, 64), 2)
    out = slerp_ref.slerp_ref(v0, v0.copy(), 7)
    for i in range(7):
        np.testing.assert_allclose(out[:, i], v0, rtol=2e-7, atol=0)


def test_slerp_orthogonal_equal_norm_keeps_norm():
    v0 = np.zeros((1, 64), np.float32)
    v1 = np.zeros((1, 64), np.float32)
    v0[0, :32] = 1.0
    v1[0, 32:] = 1.0
    out = slerp_ref.slerp_ref(v0, v1, 25)
    np.testing.assert_allclose(np.linalg.norm(out[0].astype(np.float64), axis=1), np.sqrt(32.0), rtol=1e-6)


def test_slerp_matches_closed_form_fp64():
    v0, v1 = _rand((2, 4, 72, 128), 3), _rand((2, 4, 72, 128), 4)
    out = slerp_ref.slerp_ref(v0, v1, 14)
    for c in range(2):
        x0, x1 = v0[c].astype(np.float64).ravel(), v1[c].astype(np.float64).ravel()
        th = np.arccos(x0 @ x1 / np.linalg.norm(x0) / np.linalg.norm(x1))
        for i in (0, 5, 13):
            t = i / 13
            ref = (np.sin((1 - t) * th) * x0 + np.sin(t * th) * x1) / np.sin(th)
            np.testing.assert_allclose(out[c, i].ravel(), ref, rtol=1e-5, atol=1e-6)


def test_slerp_nearly_opposite_is_finite():
    v0 = _rand((1, 256), 5)
    v1 = -v0 + 1e-3 * _rand((1, 256), 6)
    out = slerp_ref.slerp_ref(v0, v1, 14)
    assert np.isfinite(out).all()


# ---------------------------------------------------------------- warp spec vs grid_sample semantics
def test_warp_oracle_identity_and_shift():
    g = torch.Generator().manual_seed(0)
    img = torch.rand(2, 3, 12, 20, generator=g)
    zero = torch.zeros(2, 2, 12, 20)
    warped, occ, valid, _ = flow_ref.backward_warp_check(img, zero, zero)
    assert max_abs(warped, img) < 1e-12 and not occ.any() and valid.all()
    f = torch.zeros(2, 2, 12, 20)
    f[:, 0] = 1.0  # sample one pixel to the right
    warped, occ, valid, _ = flow_ref.backward_warp_check(img, f, -f)
    assert max_abs(warped[..., :-1], img[..., 1:].double()) < 1e-12
    assert (warped[..., -1] == 0).all() and not valid[..., -1].any() and valid[..., :-1].all()
    assert not occ[..., :-2].any()  # consistent flows are not occluded


def _block_from_golden(z, tag, module):
    sd = {k[len(tag) + 4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith(f"{tag}.sd.")}
    missing, unexpected = module.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    return module.eval()


def test_feed_forward_oracle_matches_reference_container(golden_dir):
    # the reference's FeedForward class (src/models/attention.py:481-535) over the restated GEGLU leaf
    from oracle import feedforward_ref
    from timelapse_video_interpolation_b200.modules import FeedForward
    z = np.load(golden_dir / "blocks.npz")
    ff = _block_from_golden(z, "ff", FeedForward(64, dim_out=32))
    x = torch.from_numpy(z["ff.x"])
    assert max_abs(feedforward_ref.feed_forward_ref(ff, x), z["ff.y"]) < 2e-5
    with torch.no_grad():  # the mirror's own eager forward (host-logic tests rely on it)
        assert max_abs(ff(x), z["ff.y"]) < 2e-5


def test_block_oracles_match_reference_block_forwards(golden_dir):
    # the reference's BasicTransformerBlock / TemporalBasicTransformerBlock forwards, state_dict keys included
    from oracle import blocks_ref
    from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, TemporalBasicTransformerBlock
    z = np.load(golden_dir / "blocks.npz")
    sb = _block_from_golden(z, "spatial", BasicTransformerBlock(64, 1, 48))
    x, e = torch.from_numpy(z["spatial.x"]), torch.from_numpy(z["spatial.e"])
    y = blocks_ref.basic_transformer_block_ref(sb, x, e)
    assert max_abs(y, z["spatial.y"]) < 5e-5
    tb = _block_from_golden(z, "temporal", TemporalBasicTransformerBlock(64, 64, 1, 48))
    x, e, T = torch.from_numpy(z["temporal.x"]), torch.from_numpy(z["temporal.e"]), int(z["temporal.num_frames"])
    y = blocks_ref.temporal_basic_transformer_block_ref(tb, x, T, e)
    assert max_abs(y, z["temporal.y"]) < 5e-5


def test_spatio_temporal_oracle_matches_reference_model(golden_dir):
    # the reference's own TransformerSpatioTemporalModel.forward (src/models/transformer_temporal.py:253-360)
    from oracle import spatio_temporal_ref
    from timelapse_video_interpolation_b200.modules import TransformerSpatioTemporalModel
    z = np.load(golden_dir / "spatio_temporal.npz")
    m = TransformerSpatioTemporalModel(1, 64, in_channels=64, num_layers=1, cross_attention_dim=48)
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd.")}
    missing, unexpected = m.load_state_dict(sd, strict=True)   # same state_dict keys as the reference module
    assert not missing and not unexpected
    x, e, ind = (torch.from_numpy(z[k]) for k in ("x", "e", "ind"))
    y = spatio_temporal_ref.transformer_spatio_temporal_ref(m.eval(), x, e, ind)
    assert y.shape == x.shape
    assert max_abs(y, z["y"]) < 1e-4
    # the image-only frame (alpha == 1) takes the spatial branch only: perturbing the temporal blocks leaves it unchanged
    with torch.no_grad():
        for p in m.temporal_transformer_blocks.parameters():
            p.add_(0.5)
    y2 = spatio_temporal_ref.transformer_spatio_temporal_ref(m, x, e, ind)
    frame = 1 * 5 + 3
    assert max_abs(y2[frame], z["y"][frame]) < 1e-4 and max_abs(y2, z["y"]) > 1e-2


@pytest.mark.parametrize("i", range(6))
def test_farneback_oracle_matches_opencv(golden_dir, i):
    # golden = the reference's compute_optical_flow_farneback (cv2.cvtColor + cv2.calcOpticalFlowFarneback) itself
    from oracle import farneback_ref
    z = np.load(golden_dir / "farneback.npz")
    a, b = z[f"a{i}"], z[f"b{i}"]
    gray = farneback_ref.gray_from_bgr(a)
    assert np.array_equal(gray, z[f"gray{i}"])                      # integer work: bit-exact
    flow = farneback_ref.calc_optical_flow_farneback(gray, farneback_ref.gray_from_bgr(b))
    ref = z[f"flow{i}"]
    assert flow.shape == ref.shape and flow.dtype == np.float32
    err = np.abs(flow - ref)
    # OpenCV accumulates the blur / polynomial passes in float SIMD, the restatement in double: tolerance 1e-3 px
    mag = max(1.0, float(np.abs(ref).max()))   # cases 4, 5 have a side < 10: OpenCV's unsigned border test wraps, flows are large
    assert err.max() < 1e-3 * mag and err.mean() < 1e-5 * mag, (err.max(), err.mean())
    assert np.abs(ref).max() > 1.0                                   # the fixture really moves


def test_step_glue_oracle_matches_reference_loop_lines(golden_dir):
    """oracle/step_ref.py:model_input_ref / cfg_ref against the reference's own loop-body lines
    (src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py:542-546, :565-566; tests/golden/make_golden.py:make_step).
    The Euler scheduler step stays a declared-unpinned leaf."""
    from oracle import step_ref
    z = np.load(golden_dir / "step_glue.npz")
    for tag in ("a", "b"):
        lat, img = torch.from_numpy(z[f"{tag}.latents"]), torch.from_numpy(z[f"{tag}.image_latents"])
        sigma = float(z[f"{tag}.sigma"])
        got = step_ref.model_input_ref(lat, img, sigma)
        want = torch.from_numpy(z[f"{tag}.latent_model_input"])
        assert got.shape == want.shape
        assert float((got.double() - want.double()).abs().max()) <= 1e-6 * max(1.0, float(want.abs().max()))
        guided = step_ref.cfg_ref(torch.from_numpy(z[f"{tag}.noise_pred"]), torch.from_numpy(z[f"{tag}.guidance_scale"]))
        want = torch.from_numpy(z[f"{tag}.guided"])
        assert guided.shape == want.shape
        assert float((guided - want.double()).abs().max()) <= 1e-6 * max(1.0, float(want.abs().max()))


def test_step_oracle_with_model_dtype_rounding_tracks_the_fp64_one():
    """oracle/step_ref.loop_step_model_dtype_ref (the reference loop's arithmetic types, bf16) against the fp64
    restatement of the same step: they differ by the roundings of the CFG combine and of prev_sample only."""
    import torch
    from oracle import step_ref
    g = torch.Generator().manual_seed(5)
    B, T, C, H, W = 2, 5, 4, 6, 8
    for pred in ("epsilon", "v_prediction"):
        for sigma, sigma_next in ((50.0, 20.0), (1.5, 0.4), (0.4, 0.0)):
            lat = (torch.randn(B, T, C, H, W, generator=g) * (sigma ** 2 + 1) ** 0.5).to(torch.bfloat16)
            noise = torch.randn(2 * B, T, C, H, W, generator=g).to(torch.bfloat16)
            gs = torch.linspace(1.0, 3.0, T).reshape(1, T, 1, 1, 1)
            prev, scaled = step_ref.loop_step_model_dtype_ref(noise, lat, gs, sigma, sigma_next, pred)
            assert prev.dtype == torch.bfloat16 and scaled.dtype == torch.bfloat16
            want = step_ref.euler_step_ref(step_ref.cfg_ref(noise, gs), lat, sigma, sigma_next, pred)
            # bf16 CFG: three roundings of O(|e|) values, scaled by |sigma_next - sigma| for epsilon and by up to sigma for v
            tol = 2 ** -7 * (want.abs() + 4.0 * max(sigma, 1.0) * 3.0)
            assert bool(((prev.double() - want).abs() <= tol).all())
            assert torch.equal(scaled, step_ref.scale_model_input_ref(prev, sigma_next))
