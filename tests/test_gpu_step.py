"""GPU parity of the fused per-step glue (CFG + Euler step + next model input) against the oracle restatement."""
import pytest
import torch

from oracle import step_ref
from timelapse_video_interpolation_b200 import _lib, ops
from timelapse_video_interpolation_b200.step import EulerCfgStepper

pytestmark = pytest.mark.gpu


def _sigmas(n, smin=0.002, smax=700.0, rho=7.0):
    ramp = torch.linspace(0, 1, n, dtype=torch.float64)
    s = (smax ** (1 / rho) + ramp * (smin ** (1 / rho) - smax ** (1 / rho))) ** rho
    return torch.cat([s, torch.zeros(1, dtype=torch.float64)])


@pytest.mark.parametrize("B,T,H,W,pred,per_batch", [(1, 14, 8, 16, "v_prediction", False), (2, 5, 6, 10, "epsilon", True),
                                                     (1, 25, 72, 128, "v_prediction", False)])
def test_stepper_matches_oracle_over_a_loop(cuda, B, T, H, W, pred, per_batch):
    g = torch.Generator().manual_seed(T)
    C, Ci, steps = 4, 5, 6
    sig = _sigmas(steps)
    guidance = torch.linspace(1.0, 3.0, T).repeat(B, 1) * (torch.arange(B).view(B, 1) * 0.5 + 1) if per_batch else torch.linspace(1.0, 3.0, T)
    lat = torch.randn(B, T, C, H, W, generator=g) * float((sig[0] ** 2 + 1) ** 0.5)
    img = torch.randn(2 * B, T, Ci, H, W, generator=g).to(torch.bfloat16)
    st = EulerCfgStepper(sig, guidance, prediction_type=pred)
    assert abs(st.init_noise_sigma - float((sig[0] ** 2 + 1) ** 0.5)) < 1e-3
    launches0 = ops.LAUNCHES
    mi = st.begin(lat.to(cuda), img.to(cuda))
    ref_lat = lat.to(torch.float64)
    want = step_ref.model_input_ref(ref_lat, img, float(sig[0]))
    assert mi.shape == (2 * B, T, C + Ci, H, W)
    assert float((mi.float().cpu() - want).abs().max()) <= 2 ** -8 * float(want.abs().max())
    assert torch.equal(mi[:, :, C:].cpu(), img)
    gs5 = guidance.reshape(-1, T)[..., None, None, None]
    for i in range(steps):
        noise = torch.randn(2 * B, T, C, H, W, generator=g).to(torch.bfloat16)
        mi = st.step(noise.to(cuda), i)
        ref_lat = step_ref.euler_step_ref(step_ref.cfg_ref(noise, gs5), ref_lat, float(sig[i]), float(sig[i + 1]), pred)
        got = st.latents.cpu().to(torch.float64)
        assert float((got - ref_lat).abs().max()) <= 1e-5 * max(1.0, float(ref_lat.abs().max())), i  # fp32 vs fp64
        if i < steps - 1:
            want = step_ref.model_input_ref(ref_lat, img, float(sig[i + 1]))
            assert float((mi.float().cpu() - want).abs().max()) <= 2 ** -8 * float(want.abs().max()) + 1e-6
            assert torch.equal(mi[:B, :, :C], mi[B:, :, :C])   # both CFG halves see the same latents
            assert torch.equal(mi[:, :, C:].cpu(), img)  # the constant channels are never rewritten
    assert ops.LAUNCHES - launches0 == steps + 1        # one launch per denoising step


def test_stepper_argument_checks(cuda):
    sig = _sigmas(3)
    st = EulerCfgStepper(sig, torch.ones(4))
    with pytest.raises(_lib.TviError):
        st.step(torch.zeros(2, 4, 4, 8, 8, dtype=torch.bfloat16, device=cuda), 0)            # before begin
    with pytest.raises(_lib.TviError):
        st.begin(torch.zeros(1, 4, 4, 8, 8), torch.zeros(2, 4, 5, 8, 8, dtype=torch.bfloat16))  # CPU tensors
    st.begin(torch.zeros(1, 4, 4, 8, 8, device=cuda), torch.zeros(2, 4, 5, 8, 8, dtype=torch.bfloat16, device=cuda))
    with pytest.raises(_lib.TviError):
        st.step(torch.zeros(2, 4, 4, 8, 8, dtype=torch.float32, device=cuda), 0)             # not bf16
    with pytest.raises(_lib.TviError):
        st.step(torch.zeros(2, 4, 4, 8, 8, dtype=torch.bfloat16, device=cuda), 3)            # past the schedule
    with pytest.raises(_lib.TviError):
        EulerCfgStepper([1.0], torch.ones(4))
    with pytest.raises(_lib.TviError):
        EulerCfgStepper(sig, torch.ones(4), prediction_type="sample")


@pytest.mark.parametrize("B,T,H,W,pred", [(1, 14, 8, 16, "v_prediction"), (2, 5, 6, 10, "epsilon"), (1, 25, 72, 128, "v_prediction")])
def test_stepper_with_the_reference_loops_rounding(cuda, B, T, H, W, pred):
    """latent_rounding="bf16": CFG in the model dtype, prev_sample cast back to it every step, model input divided in it —
    the arithmetic types of the reference loop (pipeline :563-569 + diffusers EulerDiscreteScheduler), against the torch
    bf16 restatement step by step (each step starts from the kernel's own latents, so one-ulp differences — the kernel's
    fused multiply-adds against torch's separate ops — cannot compound)."""
    g = torch.Generator().manual_seed(T + 1)
    C, Ci, steps = 4, 5, 6
    sig = _sigmas(steps)
    guidance = torch.linspace(1.0, 3.0, T)
    lat = (torch.randn(B, T, C, H, W, generator=g) * float((sig[0] ** 2 + 1) ** 0.5)).to(torch.bfloat16)
    img = torch.randn(2 * B, T, Ci, H, W, generator=g).to(torch.bfloat16)
    st = EulerCfgStepper(sig, guidance, prediction_type=pred, latent_rounding="bf16")
    st.begin(lat.to(cuda), img.to(cuda))
    assert torch.equal(st.latents.cpu(), lat.float())
    gs5 = guidance.reshape(-1, T)[..., None, None, None]
    for i in range(steps):
        before = st.latents.cpu().to(torch.bfloat16)
        assert torch.equal(before.float(), st.latents.cpu())            # the fp32 array holds bf16-representable values
        noise = torch.randn(2 * B, T, C, H, W, generator=g).to(torch.bfloat16)
        mi = st.step(noise.to(cuda), i)
        want_lat, want_in = step_ref.loop_step_model_dtype_ref(noise, before, gs5, float(sig[i]), float(sig[i + 1]), pred)
        got = st.latents.cpu()
        diff = (got - want_lat.float()).abs()
        # one bf16 step of the result, plus the fp32 rounding of the operands where the update cancels (x and sigma * e are
        # O(sigma) while the result may be small)
        ulp = want_lat.float().abs() * 2 ** -7 + 2e-6 * (float(before.float().abs().max()) + 20.0 * float(sig[i]))
        assert bool((diff <= ulp).all()), i                              # never more than one bf16 step apart
        assert float((diff > 0).float().mean()) <= 0.01, i               # and identical for all but a few elements
        if i < steps - 1:
            want_in = step_ref.scale_model_input_ref(got.to(torch.bfloat16), float(sig[i + 1]))  # from the kernel's own latents
            d2 = (mi[:B, :, :C].float().cpu() - want_in.float()).abs()
            assert bool((d2 <= want_in.float().abs().clamp_min(1e-30) * 2 ** -7).all()), i
            assert float((d2 > 0).float().mean()) <= 0.02, i
    with pytest.raises(_lib.TviError):
        EulerCfgStepper(sig, guidance, latent_rounding="fp16")
