"""Post-VFE perturbation variant: the state and arithmetic of
``StrongAdversarialVoxelNet.update_adversarial_strength`` / ``apply_enhanced_perturbations``
(/root/reference/models/detectors/strong_adversarial_voxelnet.py:114-192) without the detector around them.

The adversary perturbs the ``[M, C]`` features that come out of the voxel encoder; its perturbation is scaled by a
schedule (epoch, weak-attack boost, curriculum), a momentum term ``alpha * last`` is added and the L2 norm of the
result is tracked.  The elementwise part -- five ATen kernels and a norm in the reference -- is one fused kernel here
(``rpc_momentum_apply_fwd/bwd``); the schedule is host arithmetic on Python scalars exactly as in the reference.
Dropped: prints."""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from . import ops


class _MomentumApply(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, last, scaling: float, alpha: float, y_is_delta: bool):
        out, sp, l2 = ops.momentum_apply_fwd(x, y, last, scaling, alpha, y_is_delta)
        ctx.save_for_backward(sp, l2)
        ctx.scaling, ctx.y_is_delta = scaling, y_is_delta
        ctx.mark_non_differentiable(sp)
        return out, l2, sp

    @staticmethod
    def backward(ctx, g_out, g_l2, _g_sp):
        sp, l2 = ctx.saved_tensors
        gx, gy = ops.momentum_apply_bwd(g_out, g_l2, sp, l2, ctx.scaling, ctx.needs_input_grad[0],
                                        ctx.needs_input_grad[1], ctx.y_is_delta)
        return gx, gy, None, None, None, None


class EnhancedPerturbation:
    """Same parameters, state and return values as the reference methods (:31-36, :95-100, :114-192)."""

    def __init__(self, adversary, dynamic_scaling: bool = True, curriculum_learning: bool = True,
                 scaling_factor: float = 1.5, max_scaling: float = 5.0, momentum_alpha: float = 0.9):
        self.adversary = adversary
        self.dynamic_scaling = dynamic_scaling
        self.curriculum_learning = curriculum_learning
        self.scaling_factor = scaling_factor
        self.max_scaling = max_scaling
        self.momentum_alpha = momentum_alpha
        self._epoch = 0
        self._iteration = 0
        self._attack_history: List[float] = []
        self._current_scaling = 1.0
        self._last_perturbations: Optional[torch.Tensor] = None

    def update_adversarial_strength(self) -> float:
        if not self.dynamic_scaling:
            return 1.0
        epoch_scaling = min(1.0 + (self._epoch * 0.1), self.max_scaling)
        if len(self._attack_history) > 50:
            avg_strength = np.mean([abs(v) for v in self._attack_history[-50:]])
            boost = 2.0 if avg_strength < 0.1 else 1.5 if avg_strength < 0.3 else 1.0
            epoch_scaling *= boost
        if self.curriculum_learning:
            epoch_scaling *= min(1.0 + (self._iteration / 10000.0), 2.0)
        self._current_scaling = min(epoch_scaling, self.max_scaling)
        return self._current_scaling

    def apply_enhanced_perturbations(self, voxel_features: torch.Tensor, training: bool = True):
        if not training or self.adversary is None:
            return voxel_features, 0.0
        scaling = self.update_adversarial_strength()
        output = self.adversary(voxel_features)
        is_delta = not isinstance(output, tuple)                   # :152-161: (perturbed, l2) or the perturbation
        y = output if is_delta else output[0]
        last = self._last_perturbations
        if last is not None and last.shape != voxel_features.shape:
            last = None                                             # momentum reset on a size change (:170-173)
        out, l2, sp = _MomentumApply.apply(voxel_features, y, last, float(scaling), float(self.momentum_alpha),
                                           is_delta)
        self._last_perturbations = sp
        self._attack_history.append(l2.item())                      # the reference's one host sync per call (:183)
        if len(self._attack_history) > 1000:
            self._attack_history = self._attack_history[-500:]
        return out, l2
