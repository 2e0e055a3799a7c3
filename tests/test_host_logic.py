"""CPU: host-side behaviour of the plugin — validation, install/uninstall, no silent CPU fallback."""
import types

import pytest
import torch
from torch import nn

import timelapse_video_interpolation_b200 as tvi
from timelapse_video_interpolation_b200.modules import Attention
from timelapse_video_interpolation_b200.processor import TviAttnProcessor, install, uninstall


class TinyBlock(nn.Module):
    def __init__(self):
        super().__init__()
        self.attn1 = Attention(128, None, heads=2)
        self.attn2 = Attention(128, 96, heads=2)


class TinyUNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.transformer_blocks = nn.ModuleList([TinyBlock()])
        self.temporal_transformer_blocks = nn.ModuleList([TinyBlock()])
        self.other = nn.Linear(4, 4)


def test_state_dict_keys_match_reference_layout():
    keys = set(Attention(320, 1024, heads=5).state_dict())
    assert keys == {"to_q.weight", "to_k.weight", "to_v.weight", "to_out.0.weight", "to_out.0.bias"}
    a = Attention(320, 1024, heads=5)
    assert a.to_k.weight.shape == (320, 1024) and a.scale == 64 ** -0.5


def test_cpu_tensors_raise_not_fallback():
    attn = Attention(128, None, heads=2).to(torch.bfloat16)
    with pytest.raises(tvi.TviError, match="no CPU fallback"):
        attn(torch.zeros(2, 14, 128, dtype=torch.bfloat16))
    with pytest.raises(tvi.TviError):
        tvi.slerp(torch.zeros(1, 16), torch.zeros(1, 16), 14)
    with pytest.raises(tvi.TviError):
        tvi.backward_warp_check(torch.zeros(1, 3, 4, 4), torch.zeros(1, 2, 4, 4), torch.zeros(1, 2, 4, 4))


def test_unsupported_configurations_raise():
    p = TviAttnProcessor()
    attn = Attention(128, None, heads=2)
    x = torch.zeros(2, 14, 128, dtype=torch.bfloat16)
    with pytest.raises(tvi.TviError, match="attention_mask"):
        p(attn, x, attention_mask=torch.zeros(1))
    attn.residual_connection = True
    with pytest.raises(tvi.TviError, match="residual_connection"):
        p(attn, x)
    attn.residual_connection = False
    attn.group_norm = nn.GroupNorm(1, 128)
    with pytest.raises(tvi.TviError, match="group_norm"):
        p(attn, x)
    attn.group_norm = None
    with pytest.raises(tvi.TviError, match="batch, tokens, channels"):
        p(attn, torch.zeros(2, 128, 4, 4, dtype=torch.bfloat16))


def test_install_and_uninstall_walk_all_attention_modules():
    net = TinyUNet()
    old = [m.processor for m in net.modules() if isinstance(m, Attention)]
    shared = TviAttnProcessor()
    assert install(net, shared) == 4
    assert all(m.processor is shared for m in net.modules() if isinstance(m, Attention))
    assert uninstall(net) == 4
    assert [m.processor for m in net.modules() if isinstance(m, Attention)] == old


def test_install_reowns_monkey_patched_temporal_forward():
    # src/train_decouples.py:511-515 assigns an instance-level forward on every *temporal* Attention
    net = TinyUNet()
    calls = []
    for name, m in net.named_modules():
        if isinstance(m, Attention) and "temporal" in name:
            m.forward = lambda *a, **k: calls.append("patched")
    sink = lambda stats, attn: None
    n = install(net, temporal_ramp_bias=True, temporal_stats_sink=sink)
    assert n == 4
    t = net.temporal_transformer_blocks[0].attn2
    s = net.transformer_blocks[0].attn2
    assert isinstance(t.forward, types.MethodType) and t.processor.ramp_bias and t.processor.stats_sink is sink
    assert not s.processor.ramp_bias and "forward" not in s.__dict__
    # the re-owned forward dispatches to the processor (which refuses CPU tensors), not to the patch
    with pytest.raises(tvi.TviError):
        t(torch.zeros(1, 14, 128, dtype=torch.bfloat16), encoder_hidden_states=torch.zeros(1, 2, 96, dtype=torch.bfloat16))
    assert calls == []
    uninstall(net)
    t(torch.zeros(1, 14, 128))
    assert calls == ["patched"]


def test_fused_weight_cache_tracks_parameter_updates():
    p = TviAttnProcessor()
    attn = Attention(128, None, heads=2)
    w1 = p._fused(attn, ("to_q", "to_k", "to_v"))
    assert w1.shape == (384, 128) and w1.dtype == torch.bfloat16
    assert p._fused(attn, ("to_q", "to_k", "to_v")) is w1
    with torch.no_grad():
        attn.to_k.weight.add_(1.0)  # what optimizer.step() does: bumps _version
    w2 = p._fused(attn, ("to_q", "to_k", "to_v"))
    assert w2 is not w1 and torch.equal(w2[128:256], attn.to_k.weight.to(torch.bfloat16))
    assert "fused" not in "".join(attn.state_dict())


def test_step_oracle_known_answers():
    # N4 oracle: stepping to sigma_next = 0 lands on the predicted clean sample; CFG with scale 1 is the cond branch
    import torch
    from oracle import step_ref
    g = torch.Generator().manual_seed(0)
    x, e = torch.randn(1, 3, 4, 2, 2, generator=g, dtype=torch.float64), torch.randn(1, 3, 4, 2, 2, generator=g, dtype=torch.float64)
    sigma = 2.5
    assert torch.allclose(step_ref.euler_step_ref(e, x, sigma, 0.0, "epsilon"), x - sigma * e)
    v0 = e * (-sigma / (sigma ** 2 + 1) ** 0.5) + x / (sigma ** 2 + 1)
    assert torch.allclose(step_ref.euler_step_ref(e, x, sigma, 0.0, "v_prediction"), v0)
    assert torch.allclose(step_ref.euler_step_ref(e, x, sigma, sigma, "v_prediction"), x)  # dt = 0
    pair = torch.cat([x, e])
    assert torch.allclose(step_ref.cfg_ref(pair, torch.ones(1, 3, 1, 1, 1)), e)
    assert torch.allclose(step_ref.cfg_ref(pair, torch.zeros(1, 3, 1, 1, 1)), x)
    mi = step_ref.model_input_ref(x, torch.ones(2, 3, 5, 2, 2), 1.0)
    assert mi.shape == (2, 3, 9, 2, 2) and torch.allclose(mi[:, :, :4], torch.cat([x, x]) / 2 ** 0.5)


def test_weight_caches_are_keyed_by_live_modules():
    # regression: an id()-keyed cache served a dead module's fused weights to a new module that recycled its id and
    # parameter addresses; the caches are WeakKeyDictionaries now and drop their entries with the module
    import gc
    import weakref
    import torch
    from timelapse_video_interpolation_b200 import feedforward
    from timelapse_video_interpolation_b200.modules import Attention
    from timelapse_video_interpolation_b200.processor import TviAttnProcessor
    proc = TviAttnProcessor()
    assert isinstance(proc._wcache, weakref.WeakKeyDictionary) and isinstance(feedforward._cache, weakref.WeakKeyDictionary)
    for seed in range(3):
        torch.manual_seed(seed)
        attn = Attention(64, None, heads=1, processor=proc).to(torch.bfloat16)
        fused = proc._fused(attn, ("to_q", "to_k", "to_v"))
        want = torch.cat([attn.to_q.weight, attn.to_k.weight, attn.to_v.weight]).detach()
        assert torch.equal(fused, want)
        assert proc._fused(attn, ("to_q", "to_k", "to_v")) is fused          # cache hit
        with torch.no_grad():
            attn.to_k.weight.add_(1.0)                                        # in-place update bumps _version
        assert torch.equal(proc._fused(attn, ("to_q", "to_k", "to_v"))[64:128], attn.to_k.weight.detach())
        del attn, fused, want
        gc.collect()
        assert len(proc._wcache) == 0


def test_cache_generation_tracks_data_copy():
    """`param.data.copy_()` (diffusers EMAModel.copy_to / restore, reference
    src/train_decouples.py:1122-1124, :1199) does not bump `_version`; the derived-weight caches are keyed on a
    generation that `invalidate_caches()`, every install / uninstall and the train -> no-grad transition advance."""
    import torch
    from timelapse_video_interpolation_b200 import caches, invalidate_caches
    from timelapse_video_interpolation_b200.modules import Attention
    from timelapse_video_interpolation_b200.processor import TviAttnProcessor

    proc = TviAttnProcessor()
    attn = Attention(64, None, heads=1, dim_head=64, processor=proc).to(torch.bfloat16)
    w1 = proc._fused(attn, ("to_q", "to_k", "to_v"))
    assert proc._fused(attn, ("to_q", "to_k", "to_v")) is w1            # cache hit
    v0 = attn.to_q.weight._version
    attn.to_q.weight.data.copy_(torch.ones_like(attn.to_q.weight))     # invisible to _version ...
    assert attn.to_q.weight._version == v0
    assert proc._fused(attn, ("to_q", "to_k", "to_v")) is w1            # ... so the stale copy is still served
    invalidate_caches(attn)
    w2 = proc._fused(attn, ("to_q", "to_k", "to_v"))
    assert w2 is not w1 and torch.equal(w2[:64], torch.ones(64, 64, dtype=torch.bfloat16))
    # train -> no-grad transition (the reference's train / EMA copy_to / validate sequence) starts a new generation
    g0 = caches.generation()
    caches.note_call(True)
    caches.note_call(True)
    assert caches.generation() == g0
    caches.note_call(False)
    assert caches.generation() == g0 + 1
    caches.note_call(False)
    assert caches.generation() == g0 + 1
    # an optimizer-style in-place update (bumps _version) is seen without any of that
    with torch.no_grad():
        attn.to_k.weight.add_(1.0)
    assert proc._fused(attn, ("to_q", "to_k", "to_v")) is not w2


def test_stepper_and_feed_forward_reject_bad_arguments_without_a_gpu():
    """Argument checks that run before any CUDA call: they must raise TviError on a CPU-only machine (no silent fallback)."""
    import torch
    from timelapse_video_interpolation_b200 import _lib, feedforward
    from timelapse_video_interpolation_b200.step import EulerCfgStepper
    sig = [10.0, 5.0, 1.0, 0.0]
    with pytest.raises(_lib.TviError):
        EulerCfgStepper(sig, torch.ones(4), latent_rounding="fp16")            # only None / "bf16" are built
    with pytest.raises(_lib.TviError):
        EulerCfgStepper(sig, torch.ones(4), prediction_type="sample")
    with pytest.raises(_lib.TviError):
        EulerCfgStepper([1.0], torch.ones(4))                                   # needs at least two sigmas
    st = EulerCfgStepper(sig, torch.ones(4), latent_rounding="bf16")
    assert st.rounding == _lib.LATENT_ROUND_BF16 and abs(st.init_noise_sigma - 101.0 ** 0.5) < 1e-6
    with pytest.raises(_lib.TviError):
        st.begin(torch.zeros(1, 4, 4, 8, 8), torch.zeros(2, 4, 5, 8, 8, dtype=torch.bfloat16))  # CPU latents
    with pytest.raises(_lib.TviError):
        st.step(torch.zeros(2, 4, 4, 8, 8, dtype=torch.bfloat16), 0)           # step before begin
    with pytest.raises(_lib.TviError):
        feedforward.layer_norm_residual(torch.nn.LayerNorm(16), torch.zeros(2, 16), pad_ones=True)  # CPU fp32 rows
