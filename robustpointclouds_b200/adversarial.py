"""Host code of the adversarial input path, both compositions:

* reference order -- ``perturb_voxels`` / ``extract_voxel_features``: voxelize first, then perturb the kept
  points inside the padded voxel tensor, exactly the sequence of
  /root/reference/models/detectors/adversarial_voxelnet.py:66-137 (``extract_feat``) and
  adversarial_centerpoint.py:62-102 (``extract_pts_feat``): ``view(-1, C)`` -> padding mask -> adversary on
  the valid rows -> gradient-connected copy + masked scatter-back -> ``voxel_encoder``.
* north-star order -- ``adversarial_step``: perturb every input point, hard-voxelize the perturbed cloud and
  take the HardSimpleVFE mean in ONE persistent kernel; the backward carries the gradient of the voxel
  means to the raw per-point perturbation (``rpc_voxelize_hard`` with ``raw`` + ``rpc_step_bwd``).
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Sequence, Tuple

import torch

from . import ops
from ._lib import PerturbCfg, VoxelCfg


class _ScatterRows(torch.autograd.Function):
    """out = voxels; out_2d[valid] = rows  -- adversarial_voxelnet.py:113-117"""

    @staticmethod
    def forward(ctx, voxels: torch.Tensor, rows: torch.Tensor, row_idx: torch.Tensor, row2k: torch.Tensor):
        ctx.save_for_backward(row_idx)
        ctx.K, ctx.C = rows.shape
        return ops.scatter_rows(voxels, row2k, rows)

    @staticmethod
    def backward(ctx, grad_out):
        (row_idx,) = ctx.saved_tensors
        g_rows = ops.gather_rows(grad_out.contiguous(), row_idx, ctx.K, ctx.C) if ctx.needs_input_grad[1] else None
        # the copy is gradient-connected to `voxels` as well (out = voxels + 0 on the untouched rows);
        # voxels come from a no_grad preprocessor in the reference, so this branch is normally unused
        g_vox = None
        if ctx.needs_input_grad[0]:
            g_vox = grad_out.clone().view(-1, ctx.C)
            g_vox[row_idx[: ctx.K].long()] = 0
            g_vox = g_vox.view_as(grad_out)
        return g_vox, g_rows, None, None


LOSS_KEYS = ("l2_norm", "intensity_loss", "bias_loss", "imbalance_loss")


def zero_loss_dict(device: torch.device) -> Dict[str, torch.Tensor]:
    """The dict the reference's ``extract_feat`` falls back to when there is no valid row or the adversary fails
    (adversarial_voxelnet.py:119-131): the four keys ``AdversarialVoxelNet.loss`` indexes, all zero.  (The reference
    builds three of them on the CPU; here all four live on the voxels' device.)"""
    return {k: torch.zeros((), dtype=torch.float32, device=device) for k in LOSS_KEYS}


def perturb_voxels(voxels: torch.Tensor, num_points: torch.Tensor, adversary: Callable,
                   mask_mode: str = "rowsum"):
    """-> (perturbed_voxels [M,P,C], loss_dict, valid_count).  ``mask_mode='rowsum'`` reproduces the
    reference's ``voxels_2d.sum(dim=1) != 0`` (which also skips real points whose features sum to zero);
    ``'count'`` uses ``slot < num_points``.  One host sync (the size of the valid set), where the
    reference's boolean index has one too.  No valid row: the clean voxels and the zero-valued loss dict
    (adversarial_voxelnet.py:118-121)."""
    M, P, C = voxels.shape
    vr = ops.valid_rows(voxels, num_points, mask_mode)
    K = vr.k()
    if K == 0:
        return voxels, zero_loss_dict(voxels.device), 0
    valid = vr.rows[:K]
    out = adversary(valid)
    perturbed_valid, loss_dict = out if isinstance(out, tuple) else (out, None)
    perturbed = _ScatterRows.apply(voxels, perturbed_valid, vr.row_idx, vr.row2k)
    return perturbed, loss_dict, K


def extract_voxel_features(voxel_dict: Dict[str, torch.Tensor], voxel_encoder: Callable,
                           adversary: Optional[Callable] = None, training: bool = True, epoch: int = 3,
                           disabled: bool = False, mask_mode: str = "rowsum"):
    """The part of ``AdversarialVoxelNet.extract_feat`` in front of the middle encoder
    (adversarial_voxelnet.py:66-137): gate ``training and adversary and not disabled and epoch >= 3``
    (:77-78), perturb the valid rows, then ``voxel_encoder(voxels, num_points, coors)``.  Failure paths as the
    reference has them (:118-132): no valid row, ANY exception raised by the adversary, or its NaN fall-back
    (which returns ``(input, tensor(0.))`` -- indexing that tensor with ``'l2_norm'`` raises in the reference and
    lands in the same ``except Exception``) all continue with the CLEAN voxels and the zero-valued four-key loss
    dict that ``AdversarialVoxelNet.loss`` indexes.  Returns (voxel_features, loss_dict); loss_dict is None only
    when the gate is closed (the reference leaves ``_current_loss_dict = None`` then)."""
    voxels, num_points, coors = voxel_dict["voxels"], voxel_dict["num_points"], voxel_dict["coors"]
    loss_dict = None
    if training and adversary is not None and not disabled and epoch >= 3:
        clean = voxels
        try:
            perturbed, loss_dict, _ = perturb_voxels(voxels, num_points, adversary, mask_mode)
            loss_dict["l2_norm"]                         # a tensor / None here raises like the reference's lookup
            voxels = perturbed
        except Exception as e:                           # noqa: BLE001 -- the reference catches Exception (:123)
            print(f"Warning: Adversary failed with error: {e}")
            loss_dict = zero_loss_dict(clean.device)
            voxels = clean
    feats = voxel_encoder(voxels, num_points, coors)
    return feats, loss_dict


# ------------------------------------------------------------------------------------------------
class _AdversarialStep(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, raw, n_per_sample, cfg: VoxelCfg, bounds: PerturbCfg, out: Optional[ops.HardVoxels]):
        hv = ops.hard_voxelize(points, n_per_sample, cfg, raw=raw, bounds=bounds, want_feats=True,
                               want_maps=True, want_slots=False, out=out)
        # a caller-owned `out` is overwritten through raw pointers by the next forward: the backward keeps its own
        # copies of the two small integer maps then (one backward per forward stays valid after `out` is reused)
        num_points, p2v = (hv.num_points.clone(), hv.point2voxel.clone()) if out is not None else (hv.num_points, hv.point2voxel)
        ctx.save_for_backward(raw, num_points, p2v)
        ctx.cfg, ctx.bounds = cfg, bounds
        ctx.mark_non_differentiable(hv.voxels, hv.coors, hv.num_points, hv.voxel_num)
        return hv.feats, hv.voxels, hv.coors, hv.num_points, hv.voxel_num

    @staticmethod
    def backward(ctx, grad_feats, *unused):
        raw, num_points, p2v = ctx.saved_tensors
        grad_raw = ops.step_bwd(grad_feats.contiguous(), num_points, p2v, ctx.cfg, raw=raw, bounds=ctx.bounds)
        return None, grad_raw, None, None, None, None


def adversarial_step(points: torch.Tensor, raw: torch.Tensor, n_per_sample: Sequence[int], cfg: VoxelCfg,
                     bounds: PerturbCfg, out: Optional[ops.HardVoxels] = None):
    """Fused north-star step for a batch of sweeps concatenated in ``points`` [sum n, C]; ``raw`` [sum n, C]
    is the perturber network's output in [-1, 1] (requires grad).  Returns capacity-sized
    ``(feats, voxels, coors, num_points, voxel_num)``; only the first ``voxel_num[-1]`` rows are defined and
    nothing here synchronises the host.  ``feats.backward(g)`` fills ``raw.grad`` with
    ``(-fb <= raw*eb <= fb) ? g[v]/n[v] * eb : 0`` for the stored points and 0 elsewhere."""
    return _AdversarialStep.apply(points, raw, list(n_per_sample), cfg, bounds, out)


class _DynamicStep(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, raw, n_per_sample, cfg: VoxelCfg, bounds: PerturbCfg, out: Optional[ops.DynVoxels]):
        dv = ops.dynamic_scatter_mean(points, n_per_sample, cfg, raw=raw, bounds=bounds, out=out)
        num_points, p2v = (dv.num_points.clone(), dv.point2voxel.clone()) if out is not None else (dv.num_points, dv.point2voxel)
        ctx.save_for_backward(raw, num_points, p2v)
        ctx.cfg, ctx.bounds = cfg, bounds
        ctx.mark_non_differentiable(dv.voxel_coors, dv.num_points, dv.voxel_num, dv.point2voxel)
        return dv.feats, dv.voxel_coors, dv.num_points, dv.voxel_num, dv.point2voxel

    @staticmethod
    def backward(ctx, grad_feats, *unused):
        raw, num_points, p2v = ctx.saved_tensors
        grad_raw = ops.step_bwd(grad_feats.contiguous(), num_points, p2v, ctx.cfg, raw=raw, bounds=ctx.bounds)
        return None, grad_raw, None, None, None, None


def dynamic_step(points: torch.Tensor, raw: torch.Tensor, n_per_sample: Sequence[int], cfg: VoxelCfg,
                 bounds: PerturbCfg, out: Optional[ops.DynVoxels] = None):
    """Fused dynamic configuration (BASELINE.json config 4): perturb every point, dynamic-voxelize, mean of ALL
    points of a voxel (mmcv DynamicScatter(mean): voxels per sample in sorted (z, y, x) order, no caps).
    Returns ``(feats, voxel_coors, num_points, voxel_num, point2voxel)`` sized for one voxel per point;
    ``feats.backward(g)`` fills ``raw.grad`` with ``(-fb <= raw*eb <= fb) ? g[v]/n[v] * eb : 0``
    (DynamicScatter's autograd rule through the clamp and scale of voxel_perturber.py:256,357-359)."""
    return _DynamicStep.apply(points, raw, list(n_per_sample), cfg, bounds, out)
