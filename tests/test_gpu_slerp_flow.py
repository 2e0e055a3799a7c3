"""GPU: slerp, flow cosine similarity and backward-warp/consistency kernels vs the CPU oracle.
Tolerance from BASELINE.json: 1e-5 relative for fp32 slerp and warp."""
import numpy as np
import pytest
import torch

import timelapse_video_interpolation_b200 as tvi
from timelapse_video_interpolation_b200 import _lib, ops
from oracle import flow_ref, slerp_ref

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _rand(shape, seed):
    return np.random.default_rng(seed).standard_normal(shape).astype(np.float32)


@pytest.mark.parametrize("clips,T", [(1, 14), (3, 25), (64, 14), (2, 2)])
def test_slerp_matches_oracle(cuda, clips, T):
    v0, v1 = _rand((clips, 4, 72, 128), clips), _rand((clips, 4, 72, 128), clips + 100)
    out = tvi.slerp(torch.from_numpy(v0).to(cuda), torch.from_numpy(v1).to(cuda), T).cpu().numpy()
    ref = slerp_ref.slerp_ref(v0, v1, T)
    assert out.shape == ref.shape
    scale = np.abs(ref).max()
    assert np.abs(out - ref).max() <= RTOL * scale
    # the arithmetic is specified down to the rounding: expect (near) bit equality
    assert (out == ref).mean() > 0.999


def test_slerp_known_answers(cuda):
    v0 = _rand((2, 4, 8, 16), 1)
    dev = lambda a: torch.from_numpy(a).to(cuda)
    out = tvi.slerp(dev(v0), dev(v0.copy()), 9).cpu().numpy()          # identical -> lerp branch, all frames == v0
    np.testing.assert_allclose(out, np.repeat(v0[:, None], 9, 1), rtol=2e-7)
    a = np.zeros((1, 64), np.float32); b = np.zeros((1, 64), np.float32)
    a[0, :32] = 1; b[0, 32:] = 1                                       # orthogonal, equal norm -> norm preserved
    out = tvi.slerp(dev(a), dev(b), 25).cpu().numpy()
    np.testing.assert_allclose(np.linalg.norm(out[0].astype(np.float64), axis=1), np.sqrt(32.0), rtol=1e-6)
    np.testing.assert_array_equal(out[0, 0], a[0]); np.testing.assert_array_equal(out[0, -1], b[0])
    v1 = -v0 + 1e-3 * _rand(v0.shape, 2)                               # nearly opposite: conditioning
    out = tvi.slerp(dev(v0), dev(v1), 14).cpu().numpy()
    ref = slerp_ref.slerp_ref(v0, v1, 14)
    assert np.isfinite(out).all() and np.abs(out - ref).max() <= 1e-4 * np.abs(ref).max()
    z = np.zeros((1, 64), np.float32)                                  # zero vector -> lerp
    out = tvi.slerp(dev(z), dev(b), 5).cpu().numpy()
    np.testing.assert_allclose(out[0, 2], 0.5 * b[0])
    assert tvi.slerp(dev(v0[:0]), dev(v0[:0]), 14).shape == (0, 14, 4, 8, 16)   # empty batch
    with pytest.raises(tvi.TviError):
        tvi.slerp(dev(_rand((1, 6), 0)), dev(_rand((1, 6), 1)), 14)     # n % 4 != 0


def test_slerp_many_clips_endpoints(cuda):
    # throughput-sized batch (512 clips, 1.2 GB out): endpoints reproduce the inputs, midpoint symmetric
    g = torch.Generator(device="cuda").manual_seed(0)
    v0 = torch.randn(512, 4, 72, 128, device=cuda, generator=g)
    v1 = torch.randn(512, 4, 72, 128, device=cuda, generator=g)
    out = tvi.slerp(v0, v1, 15)
    rev = tvi.slerp(v1, v0, 15)
    assert torch.allclose(out[:, 0], v0, rtol=1e-6, atol=1e-6) and torch.allclose(out[:, -1], v1, rtol=1e-6, atol=1e-6)
    assert torch.allclose(out, rev.flip(1), rtol=1e-5, atol=1e-5)


def test_flow_cosine_matches_reference_golden(cuda, golden_dir):
    z = np.load(golden_dir / "flow_cosine.npz")
    for i in range(6):
        got = tvi.cosine_similarity(z[f"a{i}"], z[f"b{i}"])
        assert isinstance(got, np.float32)
        assert abs(float(got) - float(z[f"out{i}"])) <= RTOL * max(1.0, abs(float(z[f"out{i}"]))), i


def test_flow_cosine_batched_layouts_and_1080p(cuda):
    rng = np.random.default_rng(0)
    a = rng.normal(0, 3, (3, 270, 480, 2)).astype(np.float32)
    b = (0.5 * a + rng.normal(0, 2, a.shape)).astype(np.float32)
    ref = np.array([flow_ref.cosine_similarity(a[i], b[i]) for i in range(3)])
    ta, tb = torch.from_numpy(a).to(cuda), torch.from_numpy(b).to(cuda)
    got = tvi.cosine_similarity(ta, tb).cpu().numpy()
    np.testing.assert_allclose(got, ref, rtol=RTOL)
    got_chw = tvi.cosine_similarity_chw(ta.permute(0, 3, 1, 2).contiguous(), tb.permute(0, 3, 1, 2).contiguous()).cpu().numpy()
    np.testing.assert_allclose(got_chw, ref, rtol=RTOL)
    a = rng.normal(0, 3, (1080, 1920, 2)).astype(np.float32)
    b = (0.5 * a + rng.normal(0, 2, a.shape)).astype(np.float32)
    np.testing.assert_allclose(tvi.cosine_similarity(a, b), flow_ref.cosine_similarity(a, b), rtol=RTOL)
    odd = rng.normal(0, 1, (7, 9, 2)).astype(np.float32)              # ragged: npix odd
    np.testing.assert_allclose(tvi.cosine_similarity(odd, odd[::-1].copy()), flow_ref.cosine_similarity(odd, odd[::-1].copy()), rtol=RTOL, atol=1e-7)


def _smooth(shape, seed, sigma, amp):
    g = torch.Generator().manual_seed(seed)
    t = torch.randn(*shape, generator=g)
    k = int(4 * sigma) | 1
    ax = torch.arange(k) - k // 2
    w = torch.exp(-ax.float() ** 2 / (2 * sigma ** 2)); w /= w.sum()
    c = shape[1]
    t = torch.nn.functional.conv2d(t, w.view(1, 1, 1, k).repeat(c, 1, 1, 1), padding=(0, k // 2), groups=c)
    t = torch.nn.functional.conv2d(t, w.view(1, 1, k, 1).repeat(c, 1, 1, 1), padding=(k // 2, 0), groups=c)
    return t / t.std() * amp


@pytest.mark.parametrize("B,H,W", [(2, 64, 96), (1, 135, 241), (1, 270, 480)])
def test_backward_warp_check_matches_grid_sample_oracle(cuda, B, H, W):
    img = _smooth((B, 3, H, W), 1, 2.0, 1.0)
    f_fwd = _smooth((B, 2, H, W), 2, 8.0, 5.0)
    f_bwd = -f_fwd + _smooth((B, 2, H, W), 3, 3.0, 0.4)
    f_fwd[:, :, :4, :4] = 300.0  # far out of bounds
    warped, occ, valid = tvi.backward_warp_check(img.to(cuda), f_fwd.to(cuda), f_bwd.to(cuda))
    torch.cuda.synchronize()
    w_ref, occ_ref, valid_ref, margin = flow_ref.backward_warp_check(img, f_fwd, f_bwd)
    scale = float(w_ref.abs().max())
    assert float((warped.cpu().double() - w_ref).abs().max()) <= RTOL * scale
    assert torch.equal(valid.cpu(), valid_ref)
    decided = margin.abs() > 1e-4  # pixels within rounding of the threshold may flip (counted)
    assert torch.equal(occ.cpu()[decided], occ_ref[decided])
    assert float((~decided).float().mean()) < 1e-3
    assert 0.0 < float(occ_ref.float().mean()) < 1.0


def test_backward_warp_identity_and_1080p_roundtrip(cuda):
    g = torch.Generator(device="cuda").manual_seed(0)
    img = torch.rand(2, 3, 1080, 1920, device=cuda, generator=g)
    zero = torch.zeros(2, 2, 1080, 1920, device=cuda)
    warped, occ, valid = tvi.backward_warp_check(img, zero, zero)
    assert torch.equal(warped, img) and not bool(occ.any()) and bool(valid.all())
    shift = zero.clone(); shift[:, 0] = 3.0; shift[:, 1] = -2.0   # integer shift: exact gather
    warped, occ, valid = tvi.backward_warp_check(img, shift, -shift)
    assert torch.equal(warped[:, :, 2:, :-3], img[:, :, :-2, 3:])
    assert not bool(valid[:, :2].any()) and not bool(valid[:, :, -3:].any()) and bool(valid[:, 2:, :-3].all())
    assert not bool(occ[:, 4:, :-6].any())


@pytest.mark.parametrize("i", range(6))
def test_farneback_matches_opencv_golden_and_oracle(cuda, golden_dir, i):
    import numpy as np
    from oracle import farneback_ref
    from timelapse_video_interpolation_b200 import compute_optical_flow_farneback
    z = np.load(golden_dir / "farneback.npz")
    a, b = z[f"a{i}"], z[f"b{i}"]
    flow = compute_optical_flow_farneback(a, b)                      # numpy BGR in -> numpy (H, W, 2) out
    ref = z[f"flow{i}"]
    assert isinstance(flow, np.ndarray) and flow.shape == ref.shape and flow.dtype == np.float32
    err = np.abs(flow - ref)
    print(f"farneback {ref.shape}: vs cv2 max {err.max():.2e} mean {err.mean():.2e} (max |flow| {np.abs(ref).max():.2f})")
    mag = max(1.0, float(np.abs(ref).max()))
    assert err.max() < 5e-3 * mag and err.mean() < 5e-5 * mag
    ora = farneback_ref.calc_optical_flow_farneback(farneback_ref.gray_from_bgr(a), farneback_ref.gray_from_bgr(b))
    assert np.abs(flow - ora).max() < 5e-3 * mag
    # gray input path and CUDA-tensor path
    ga = torch.from_numpy(z[f"gray{i}"]).to(cuda)
    gb = torch.from_numpy(farneback_ref.gray_from_bgr(b)).to(cuda)
    f2 = ops.farneback(ga, gb)
    assert torch.equal(f2.cpu(), torch.from_numpy(flow))


def test_farneback_full_hd_and_edge_cases(cuda):
    import numpy as np
    from oracle import farneback_ref
    rng = np.random.default_rng(0)
    # 1080p: identical frames -> zero flow; shifted frame -> the shift (size-independent properties)
    base = rng.random((1080 // 8 + 2, 1920 // 8 + 2)).astype(np.float32)
    t = torch.nn.functional.interpolate(torch.from_numpy(base)[None, None], size=(1080 + 8, 1920 + 8), mode="bicubic")[0, 0]
    img = (t.clamp(0, 1) * 255).to(torch.uint8)
    a, b = img[4:-4, 4:-4].contiguous().to(cuda), img[2:-6, 7:-1].contiguous().to(cuda)  # b(y, x) = a(y - 2, x + 3)
    f0 = ops.farneback(a, a)   # identical frames: zero flow away from the image border (the border terms are OpenCV's)
    assert float(f0[300:-300, 300:-300].abs().max()) < 1e-3
    f = ops.farneback(a, b)
    core = f[100:-100, 100:-100]
    assert abs(float(core[..., 0].median()) + 3.0) < 0.1 and abs(float(core[..., 1].median()) - 2.0) < 0.1
    # tiny image (pyramid collapses to one level) against the oracle
    sa = (rng.random((33, 40)) * 255).astype(np.uint8)
    sb = np.roll(sa, 1, axis=1)
    got = ops.farneback(torch.from_numpy(sa).to(cuda), torch.from_numpy(sb).to(cuda)).cpu().numpy()
    want = farneback_ref.calc_optical_flow_farneback(sa, sb)
    assert np.abs(got - want).max() < 5e-3 * max(1.0, np.abs(want).max())
    with pytest.raises(_lib.TviError):
        ops.farneback(torch.zeros(8, 8, 2, dtype=torch.uint8, device=cuda), torch.zeros(8, 8, 2, dtype=torch.uint8, device=cuda))
    with pytest.raises(_lib.TviError):
        ops.farneback(torch.zeros(8, 8, dtype=torch.float32, device=cuda), torch.zeros(8, 8, dtype=torch.float32, device=cuda))


def test_check_clip_flows_matches_the_reference_loop(cuda):
    # dataset_construction/step4_flow_check.py:55-81 on a 5-frame clip: steady shift, then a jump backwards (an outlier)
    import numpy as np
    from oracle import farneback_ref
    rng = np.random.default_rng(3)
    base = np.kron(rng.random((12, 14, 3)), np.ones((8, 8, 1)))
    base = (base * 200 + 25 * rng.random(base.shape)).astype(np.uint8)
    shifts = [0, 2, 4, 6, 1]
    frames = np.stack([np.roll(base, s, axis=1)[8:-8, 8:-8] for s in shifts])          # [5, 80, 96, 3]
    flows, sims, flagged = tvi.check_clip_flows(frames)
    assert flows.shape == (4, 80, 96, 2) and sims.shape == (3,) and flagged.dtype == torch.bool
    want_flows = [farneback_ref.calc_optical_flow_farneback(farneback_ref.gray_from_bgr(frames[i]), farneback_ref.gray_from_bgr(frames[i + 1]))
                  for i in range(4)]
    for i in range(4):
        assert np.abs(flows[i].cpu().numpy() - want_flows[i]).max() < 5e-3
    want_sims = [flow_ref.cosine_similarity(want_flows[i], want_flows[i + 1]) for i in range(3)]
    assert np.abs(sims.cpu().numpy() - np.array(want_sims)).max() < 1e-4
    assert flagged.cpu().tolist() == [bool(s < 0.5) for s in want_sims] == [False, False, True]
    with pytest.raises(ValueError):
        tvi.check_clip_flows(frames[:2])


def test_farneback_clip_equals_pairwise_calls(cuda):
    import numpy as np
    rng = np.random.default_rng(5)
    base = (np.kron(rng.random((14, 18)), np.ones((8, 8))) * 255).astype(np.uint8)
    clip = torch.from_numpy(np.stack([np.roll(base, (i, 2 * i), axis=(0, 1)) for i in range(5)])).to(cuda)   # [5, 112, 144] gray
    flows = ops.farneback_clip(clip)
    assert flows.shape == (4, 112, 144, 2)
    for i in range(4):
        assert torch.equal(flows[i], ops.farneback(clip[i], clip[i + 1]))   # the shared pyramids change nothing
    with pytest.raises(_lib.TviError):
        ops.farneback_clip(clip[:1])
