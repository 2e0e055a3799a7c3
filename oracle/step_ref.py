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
