"""Validity of the derived weight caches (fused QKV / KV weights, block-interleaved GEGLU weights, the frame-index
embedding, the blender alpha).

Those caches hold bf16 COPIES of module parameters, keyed on `(data_ptr, _version)` of the sources.  That key sees
optimizer steps, `load_state_dict`, `param.copy_()` and `param.data = new` — but not an in-place write through
`.data` (`param.data.copy_(x)` leaves `_version` untouched), which is exactly what diffusers' `EMAModel.copy_to /
restore` do around validation (reference src/train_decouples.py:1122-1124, :1199).  Every cache key therefore also
carries a process-wide GENERATION number, and the generation advances
  * on `invalidate_caches()` (public; call it after any `.data` surgery on the weights),
  * on every `install*` / `uninstall*`,
  * automatically at the first no-grad libtvi call that follows a call on the autograd tape — the
    train -> EMA copy_to -> validate sequence of the reference's loop.
Rebuilding is a concat / cast of a few MB, so a spurious invalidation costs microseconds.
"""
from __future__ import annotations

__all__ = ["invalidate_caches", "generation", "note_call"]

_generation = 0
_saw_tape_call = False


def generation() -> int:
    return _generation


def invalidate_caches(model=None) -> int:
    """Drops every derived weight cache (lazily: the next call rebuilds what it needs).  `model` is accepted for
    symmetry with `install(model)`; the invalidation is process-wide.  Returns the new generation."""
    global _generation
    _generation += 1
    return _generation


def note_call(on_tape: bool) -> None:
    """Called by the attention / feed-forward entry points: a no-grad call that follows a training call starts a new
    generation (the weights may have been swapped through `.data` in between, e.g. EMA validation)."""
    global _saw_tape_call
    if on_tape:
        _saw_tape_call = True
    elif _saw_tape_call:
        _saw_tape_call = False
        invalidate_caches()
