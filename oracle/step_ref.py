"""CPU restatement of the per-step glue of the denoising loop (test infrastructure).

  per-frame CFG          src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py:563-566
  model input            :542-546 (cat([latents] * 2) -> scale_model_input -> cat with image_latents over dim 2)
  EulerDiscreteScheduler scale_model_input / step (s_churn = 0): upstream diffusers 0.24
                         schedulers/scheduling_euler_discrete.py — un-vendored, restated from the published code;
                         parity for the scheduler arithmetic is UNPINNED by the reference (the CFG and concat lines are the
                         reference's own).
"""
from __future__ import annotations

import torch


def scale_model_input_ref(sample: torch.Tensor, sigma: float) -> torch.Tensor:
    return sample / ((sigma ** 2 + 1) ** 0.5)


def euler_step_ref(model_output: torch.Tensor, sample: torch.Tensor, sigma: float, sigma_next: float, prediction_type: str):
    sample = sample.to(torch.float64)
    model_output = model_output.to(torch.float64)
    if prediction_type == "epsilon":
        pred_original = sample - sigma * model_output
    elif prediction_type == "v_prediction":
        pred_original = model_output * (-sigma / (sigma ** 2 + 1) ** 0.5) + (sample / (sigma ** 2 + 1))
    else:
        raise ValueError(prediction_type)
    derivative = (sample - pred_original) / sigma
    return sample + derivative * (sigma_next - sigma)


def cfg_ref(noise_pred: torch.Tensor, guidance_scale: torch.Tensor) -> torch.Tensor:
    """noise_pred [2B, T, C, H, W]; guidance_scale broadcastable [B or 1, T, 1, 1, 1] (:530 `_append_dims`)."""
    uncond, cond = noise_pred.to(torch.float64).chunk(2)
    return uncond + guidance_scale.to(torch.float64) * (cond - uncond)


def model_input_ref(latents: torch.Tensor, image_latents: torch.Tensor, sigma: float) -> torch.Tensor:
    x = torch.cat([latents] * 2)
    x = scale_model_input_ref(x, sigma)
    return torch.cat([x, image_latents.to(x.dtype)], dim=2)


def loop_step_model_dtype_ref(noise_pred: torch.Tensor, latents: torch.Tensor, guidance_scale: torch.Tensor, sigma: float,
                              sigma_next: float, prediction_type: str):
    """One loop iteration with the arithmetic TYPES of the reference loop (model dtype = that of `noise_pred` / `latents`,
    e.g. bf16): the CFG lines :563-566 on model-dtype tensors (torch rounds every op), EulerDiscreteScheduler.step's fp32
    upcast and `prev_sample.to(model_output.dtype)`, then `scale_model_input` in the model dtype for the next UNet call.
    Returns (latents, scaled latents), both in the model dtype."""
    dt = noise_pred.dtype
    uncond, cond = noise_pred.chunk(2)
    e = uncond + guidance_scale.to(dt) * (cond - uncond)
    sample = latents.to(torch.float32)
    mo = e.to(torch.float32)
    if prediction_type == "epsilon":
        pred_original = sample - sigma * mo
    elif prediction_type == "v_prediction":
        pred_original = mo * (-sigma / (sigma ** 2 + 1) ** 0.5) + (sample / (sigma ** 2 + 1))
    else:
        raise ValueError(prediction_type)
    derivative = (sample - pred_original) / sigma
    prev = (sample + derivative * (sigma_next - sigma)).to(dt)
    return prev, scale_model_input_ref(prev, sigma_next)
