"""CPU oracle for the tvi hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  Nothing under timelapse_video_interpolation_b200/ imports it: the product path has
no CPU fallback.

Pinning status:
  attention_ref   pinned: checked against golden vectors produced by the reference's OWN
                  src/models/attention_processor.py (Attention + AttnProcessor / AttnProcessor2_0),
                  imported in the build container behind a 30-line shim for the absent diffusers
                  leaf imports (tests/golden/make_golden.py).
  flow_ref.cosine_similarity  pinned: golden vectors from the reference's own
                  dataset_construction/step4_flow_check.py:25-34.
  flow_ref.backward_warp_check / slerp_ref   PARITY UNPINNED: the reference has no such code
                  (0-byte frame_slerp_interp.py; no warp in step4_flow_check.py) — the spec is the
                  build's own; warp is cross-checked against
                  torch.nn.functional.grid_sample.
"""
