"""Change tracking for the modules whose forward is a call into the CUDA library: the library holds a packed (folded,
bf16) copy of the weights, which must be rebuilt when the module's tensors change - without walking the module tree on
every forward.  PyTorch is plumbing here: the key is built from `Tensor._version` (bumped by every in-place op that goes
through autograd's bookkeeping: `copy_`, optimizer steps, `load_state_dict`) plus an epoch bumped by `invalidate()`."""
from __future__ import annotations

from typing import Tuple

import torch


class PackedWeightsMixin:
    """Mix into an `nn.Module` (before `nn.Module` in the MRO).  `_skip_in_key` names state-dict entries that never
    reach the library (e.g. `num_batches_tracked`, the vocoder's `step`)."""

    _skip_in_key: Tuple[str, ...] = ("num_batches_tracked",)

    def _init_tracking(self) -> None:
        self._epoch = 0
        self._tensor_cache = None

    def _tracked(self):
        """(parameters + affine terms, BatchNorm running statistics), cached until `invalidate()`."""
        if self._tensor_cache is None:
            params, stats = [], []
            for name, t in self.state_dict(keep_vars=True).items():
                if any(s in name for s in self._skip_in_key):
                    continue
                (stats if "running_" in name else params).append(t)
            self._tensor_cache = (tuple(params), tuple(stats))
        return self._tensor_cache

    def _weights_key(self):
        """((epoch, version sum of all tensors), (epoch, version sum without the running statistics)).  Writes through
        `.data` (`p.data.copy_(...)`, what the reference's own `load_flattened_params` does) do not bump `_version`:
        call `invalidate()` after such a write."""
        params, stats = self._tracked()
        p = 0
        for t in params:
            p += t._version
        k = p
        for t in stats:
            k += t._version
        return (self._epoch, k), (self._epoch, p)

    def invalidate(self) -> None:
        """Mark the library's packed copy of the weights stale: the next forward re-packs from the module's tensors.
        Needed only after writes that bypass the version counter (`tensor.data`) or after replacing a parameter object;
        `load_state_dict`, `.to()`/`.cuda()` and in-place ops on the parameters are picked up automatically."""
        self._epoch = getattr(self, "_epoch", 0) + 1
        self._tensor_cache = None

    refresh_weights = invalidate

    def _apply(self, fn, *a, **k):      # .to() / .cuda() / .half(): parameter storage may be replaced
        r = super()._apply(fn, *a, **k)
        self.invalidate()
        return r

    def load_state_dict(self, *a, **k):
        r = super().load_state_dict(*a, **k)
        self.invalidate()
        return r
