"""Tensor-level entry points over the C-ABI: PyTorch owns device memory and streams, the kernels in
``librpc_b200.so`` do the work.  No op here has a CPU or eager-PyTorch fallback.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import PerturbCfg, VoxelCfg


def _need_cuda(*tensors: Optional[torch.Tensor]) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("robustpointclouds_b200 ops run on CUDA tensors only (sm_100a kernels; "
                               "there is no CPU fallback)")
        dev = t.device
    return dev


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream(dev: torch.device) -> int:
    # (the raw handle of the current stream; torch.cuda.current_stream(dev).cuda_stream builds a Stream object first and
    # costs ~7 us a call -- 23 calls per training step)
    return torch._C._cuda_getCurrentRawStream(dev.index if dev.index is not None else torch.cuda.current_device())


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def grid_size(cfg: VoxelCfg) -> Tuple[int, int, int]:
    g = (ctypes.c_int32 * 3)()
    _lib.check(_lib.lib().rpc_grid_size(ctypes.byref(cfg), g), "rpc_grid_size")
    return g[0], g[1], g[2]


@dataclass
class HardVoxels:
    """Outputs of one batched hard voxelization, sized at capacity (sum_b min(n_b, V)) rows; only the
    first ``voxel_num[-1]`` rows are defined.  ``voxel_num`` lives on the device ([B+1]: per sweep,
    then the total) so that nothing here forces a host sync."""
    voxels: torch.Tensor
    coors: torch.Tensor
    num_points: torch.Tensor
    voxel_num: torch.Tensor
    point2voxel: Optional[torch.Tensor]
    point2slot: Optional[torch.Tensor]
    feats: Optional[torch.Tensor]
    capacity: int

    def total(self) -> int:
        """number of voxels (host sync)"""
        return int(self.voxel_num[-1].item())


class Workspace:
    """Grow-only scratch buffers, one per (device, stream) and purpose, reused across calls.  Two ops that
    run on different streams never share a buffer, and a buffer that is replaced by a larger one stays
    alive (``_retired``): a captured CUDA graph may still hold its address."""

    def __init__(self, headroom: float = 1.0) -> None:
        self._buf = {}
        self._retired = []
        self.headroom = headroom

    def _key(self, dev: torch.device):
        i = dev.index if dev.index is not None else torch.cuda.current_device()
        return (i, torch._C._cuda_getCurrentRawStream(i))

    def get(self, nbytes: int, dev: torch.device) -> torch.Tensor:
        k = self._key(dev)
        b = self._buf.get(k)
        if b is None or b.numel() < nbytes:
            if b is not None:
                self._retired.append(b)
            b = torch.empty(max(int(nbytes * self.headroom) + 256, 1), dtype=torch.uint8, device=dev)
            self._buf[k] = b
        return b

    def current(self, dev: torch.device) -> torch.Tensor:
        return self._buf[self._key(dev)]


_default_ws = Workspace()
_scratch_ws = Workspace()        # the small fp64 accumulators of the reduction kernels (zeroed by every call that uses them)


def _scratch64(n: int, dev: torch.device) -> torch.Tensor:
    """n doubles of per-(device, stream) scratch: the kernels that use it are stream-ordered and zero it themselves, so
    one cached buffer replaces a torch.empty per call (part of the Python overhead around the small kernels)"""
    return _scratch_ws.get(8 * n, dev)[:8 * n].view(torch.float64)
# rpc_voxelize_hard has a workspace of its own per (device, stream): a captured CUDA graph bakes the address in, and
# nothing else should be scribbling there while a step is in flight.  25 % headroom: ragged batches of a training run
# change the layout a little every step without outgrowing the buffer.
_hard_ws = Workspace(headroom=1.25)


def hard_workspace_bytes(cfg: VoxelCfg, n_per_sample: Sequence[int]) -> int:
    """``rpc_workspace_bytes``: smallest workspace ``hard_voxelize`` accepts for this batch.  A caller that keeps
    its own workspace (a captured CUDA graph must) allocates it once with ``torch.empty``; every call initialises what
    it reads."""
    n = _lib.lib().rpc_workspace_bytes(ctypes.byref(cfg), len(n_per_sample), _lib.counts(n_per_sample))
    if n == 0:
        raise _lib.RpcError(-2, "rpc_workspace_bytes")
    return int(n)


def dynamic_workspace_bytes(cfg: VoxelCfg, n_per_sample: Sequence[int]) -> int:
    """``rpc_dynamic_workspace_bytes``: scratch of ``dynamic_scatter_mean`` for this batch"""
    n = _lib.lib().rpc_dynamic_workspace_bytes(ctypes.byref(cfg), len(n_per_sample), _lib.counts(n_per_sample))
    if n == 0:
        raise _lib.RpcError(-2, "rpc_dynamic_workspace_bytes")
    return int(n)


def hard_voxelize(points: torch.Tensor, n_per_sample: Sequence[int], cfg: VoxelCfg,
                  raw: Optional[torch.Tensor] = None, bounds: Optional[PerturbCfg] = None,
                  want_feats: bool = False, want_maps: bool = False, want_slots: Optional[bool] = None,
                  out: Optional[HardVoxels] = None, workspace: Optional[torch.Tensor] = None) -> HardVoxels:
    """``rpc_voxelize_hard``: B sweeps concatenated in ``points`` [sum n_b, C]; ``raw`` [sum n_b, C] +
    ``bounds`` voxelize the perturbed points x + clamp(raw*eb, +-fb) instead.  ``want_maps`` adds point2voxel
    (what the backward needs) and, unless ``want_slots=False``, point2slot."""
    dev = _need_cuda(points, raw)
    points = _f32c(points)
    if points.dim() != 2 or points.shape[1] != cfg.C:
        raise ValueError(f"points must be [N, {cfg.C}], got {tuple(points.shape)}")
    if sum(n_per_sample) != points.shape[0]:
        raise ValueError("n_per_sample does not add up to the number of rows")
    if raw is not None:
        raw = _f32c(raw)
        if raw.shape != points.shape or bounds is None:
            raise ValueError("raw must match points and come with bounds")
    L = _lib.lib()
    B = len(n_per_sample)
    n_arr = _lib.counts(n_per_sample)
    cap = L.rpc_voxel_capacity(ctypes.byref(cfg), B, n_arr)
    if cap < 0:
        raise _lib.RpcError(-2, "rpc_voxel_capacity")
    nbytes = L.rpc_workspace_bytes(ctypes.byref(cfg), B, n_arr)
    ws = workspace if workspace is not None else _hard_ws.get(nbytes, dev)
    N, P, C, F = points.shape[0], cfg.max_points, cfg.C, cfg.F
    if out is None:
        i32 = dict(dtype=torch.int32, device=dev)
        out = HardVoxels(
            voxels=torch.empty((cap, P, C), dtype=torch.float32, device=dev),
            coors=torch.empty((cap, 4), **i32), num_points=torch.empty((cap,), **i32),
            voxel_num=torch.empty((B + 1,), **i32),
            point2voxel=torch.empty((N,), **i32) if want_maps else None,
            point2slot=torch.empty((N,), **i32) if (want_maps if want_slots is None else want_slots) else None,
            feats=torch.empty((cap, F), dtype=torch.float32, device=dev) if want_feats else None,
            capacity=int(cap))
    st = L.rpc_voxelize_hard(ctypes.byref(cfg), _ptr(points), n_arr, B, _ptr(raw),
                             ctypes.byref(bounds) if bounds is not None else None,
                             _ptr(out.voxels), _ptr(out.coors), _ptr(out.num_points), _ptr(out.voxel_num),
                             _ptr(out.point2voxel), _ptr(out.point2slot), _ptr(out.feats),
                             ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(st, "rpc_voxelize_hard")
    return out


def step_bwd(grad_feats: torch.Tensor, num_points: torch.Tensor, point2voxel: torch.Tensor,
             cfg: VoxelCfg, raw: Optional[torch.Tensor] = None, bounds: Optional[PerturbCfg] = None,
             out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``rpc_step_bwd``: gradient of the voxel means w.r.t. the raw per-point perturbation [N, C]
    (or w.r.t. the point features when ``raw`` is None)."""
    dev = _need_cuda(grad_feats, num_points, point2voxel, raw)
    grad_feats = _f32c(grad_feats)
    N = point2voxel.shape[0]
    if raw is not None:
        raw = _f32c(raw)
    if out is None:
        out = torch.empty((N, cfg.C), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_step_bwd(ctypes.byref(cfg), _ptr(grad_feats), _ptr(num_points), _ptr(point2voxel),
                                 _ptr(raw), ctypes.byref(bounds) if bounds is not None else None,
                                 N, _ptr(out), _stream(dev))
    _lib.check(st, "rpc_step_bwd")
    return out


def dynamic_voxelize(points: torch.Tensor, n_per_sample: Sequence[int], cfg: VoxelCfg,
                     raw: Optional[torch.Tensor] = None, bounds: Optional[PerturbCfg] = None,
                     ncol: int = 4) -> torch.Tensor:
    """``rpc_voxelize_dynamic``: per-point (batch, z, y, x) (ncol 4) or (z, y, x) (ncol 3); dropped
    points get -1 in all three cell columns."""
    dev = _need_cuda(points, raw)
    points = _f32c(points)
    if raw is not None:
        raw = _f32c(raw)
    coors = torch.empty((points.shape[0], ncol), dtype=torch.int32, device=dev)
    st = _lib.lib().rpc_voxelize_dynamic(ctypes.byref(cfg), _ptr(points), _lib.counts(n_per_sample),
                                         len(n_per_sample), _ptr(raw),
                                         ctypes.byref(bounds) if bounds is not None else None,
                                         _ptr(coors), ncol, _stream(dev))
    _lib.check(st, "rpc_voxelize_dynamic")
    return coors


@dataclass
class DynVoxels:
    """Outputs of one batched dynamic voxelization + scatter-mean, sized for one voxel per point; only the
    first ``voxel_num[-1]`` rows of ``feats`` / ``voxel_coors`` / ``num_points`` are defined."""
    feats: torch.Tensor            # [N, F] f32
    voxel_coors: torch.Tensor      # [N, 4] i32 (b, z, y, x)
    num_points: torch.Tensor       # [N] i32
    voxel_num: torch.Tensor        # [B + 1] i32 on the device
    point2voxel: torch.Tensor      # [N] i32
    coors: Optional[torch.Tensor]  # [N, 4] i32 per point, or None

    def total(self) -> int:
        return int(self.voxel_num[-1].item())


def dynamic_scatter_mean(points: torch.Tensor, n_per_sample: Sequence[int], cfg: VoxelCfg,
                         raw: Optional[torch.Tensor] = None, bounds: Optional[PerturbCfg] = None,
                         coors_in: Optional[torch.Tensor] = None, want_coors: bool = True,
                         out: Optional[DynVoxels] = None, workspace: Optional[torch.Tensor] = None) -> DynVoxels:
    """``rpc_dynamic_scatter_mean_fwd``: dynamic voxelization of the (perturbed) points and mmcv
    DynamicScatter(mean) in one launch; with ``coors_in`` [N,4] the points are scattered by those rows."""
    dev = _need_cuda(points, raw, coors_in)
    points = _f32c(points)
    if points.dim() != 2 or points.shape[1] != cfg.C:
        raise ValueError(f"points must be [N, {cfg.C}], got {tuple(points.shape)}")
    if sum(n_per_sample) != points.shape[0]:
        raise ValueError("n_per_sample does not add up to the number of rows")
    if raw is not None:
        raw = _f32c(raw)
        if raw.shape != points.shape or bounds is None:
            raise ValueError("raw must match points and come with bounds")
    if coors_in is not None:
        coors_in = coors_in.to(torch.int32).contiguous()
        if coors_in.shape != (points.shape[0], 4):
            raise ValueError("coors_in must be [N, 4] (batch, z, y, x)")
    L = _lib.lib()
    B = len(n_per_sample)
    n_arr = _lib.counts(n_per_sample)
    nbytes = L.rpc_dynamic_workspace_bytes(ctypes.byref(cfg), B, n_arr)
    if nbytes == 0:
        raise _lib.RpcError(-2, "rpc_dynamic_workspace_bytes")
    ws = workspace if workspace is not None else _default_ws.get(nbytes, dev)
    N, F = points.shape[0], cfg.F
    if out is None:
        i32 = dict(dtype=torch.int32, device=dev)
        out = DynVoxels(feats=torch.empty((N, F), dtype=torch.float32, device=dev),
                        voxel_coors=torch.empty((N, 4), **i32), num_points=torch.empty((N,), **i32),
                        voxel_num=torch.empty((B + 1,), **i32), point2voxel=torch.empty((N,), **i32),
                        coors=torch.empty((N, 4), **i32) if (want_coors and coors_in is None) else None)
    st = L.rpc_dynamic_scatter_mean_fwd(ctypes.byref(cfg), _ptr(points), n_arr, B, _ptr(raw),
                                        ctypes.byref(bounds) if bounds is not None else None,
                                        _ptr(coors_in), _ptr(out.coors), _ptr(out.point2voxel), _ptr(out.feats),
                                        _ptr(out.voxel_coors), _ptr(out.num_points), _ptr(out.voxel_num),
                                        ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(st, "rpc_dynamic_scatter_mean_fwd")
    return out


def vfe_mean_fwd(voxels: torch.Tensor, num_points: torch.Tensor, num_features: int,
                 m_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    dev = _need_cuda(voxels, num_points)
    voxels = _f32c(voxels)
    num_points = num_points.to(torch.int32).contiguous()
    M, P, C = voxels.shape
    out = torch.empty((M, num_features), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_vfe_mean_fwd(_ptr(voxels), _ptr(num_points), M, P, C, num_features, _ptr(m_dev),
                                     _ptr(out), _stream(dev))
    _lib.check(st, "rpc_vfe_mean_fwd")
    return out


def vfe_mean_bwd(grad_out: torch.Tensor, num_points: torch.Tensor, P: int, C: int,
                 m_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    dev = _need_cuda(grad_out, num_points)
    grad_out = _f32c(grad_out)
    num_points = num_points.to(torch.int32).contiguous()
    M, F = grad_out.shape
    g = torch.empty((M, P, C), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_vfe_mean_bwd(_ptr(grad_out), _ptr(num_points), M, P, C, F, _ptr(m_dev), _ptr(g),
                                     _stream(dev))
    _lib.check(st, "rpc_vfe_mean_bwd")
    return g


PHASES = ("clear", "insert", "flag", "assign", "rank", "gather")


def phase_times_us(cfg: VoxelCfg, n_per_sample: Sequence[int], workspace: Optional[torch.Tensor] = None,
                   dev: Optional[torch.device] = None):
    """{phase: us} of the most recent profiled ``hard_voxelize`` call (``_lib.profile_enable(True)``),
    from the %globaltimer stamps CTA 0 of the persistent kernel leaves in the workspace (host sync)."""
    n_arr = _lib.counts(n_per_sample)
    off = _lib.lib().rpc_workspace_stamps_offset(ctypes.byref(cfg), len(n_per_sample), n_arr)
    if off < 0:
        raise _lib.RpcError(-2, "rpc_workspace_stamps_offset")
    ws = workspace if workspace is not None else _hard_ws.current(dev)
    st = ws[off:off + 56].view(torch.int64).cpu().tolist()
    return {PHASES[k]: (st[k + 1] - st[k]) / 1e3 for k in range(6)}


DYN_PHASES = ("clear", "mark", "coarse_sums", "coarse_scan", "fine_bits", "fine_sums", "rank", "sum", "mean")


def dyn_phase_times_us(cfg: VoxelCfg, n_per_sample: Sequence[int], workspace: Optional[torch.Tensor] = None,
                       dev: Optional[torch.device] = None):
    """{phase: us} of the most recent profiled ``dynamic_scatter_mean`` call (host sync)"""
    n_arr = _lib.counts(n_per_sample)
    off = _lib.lib().rpc_dynamic_stamps_offset(ctypes.byref(cfg), len(n_per_sample), n_arr)
    if off < 0:
        raise _lib.RpcError(-2, "rpc_dynamic_stamps_offset")
    ws = workspace if workspace is not None else _default_ws.current(dev)
    st = ws[off:off + 80].view(torch.int64).cpu().tolist()
    return {DYN_PHASES[k]: (st[k + 1] - st[k]) / 1e3 for k in range(9)}


# ------------------------------------------------------------------------------------------------
# reference-order composition (perturb the kept points inside the padded voxel tensor)
# ------------------------------------------------------------------------------------------------
MASK_MODES = {"rowsum": 0, "count": 1}


@dataclass
class ValidRows:
    """``rpc_valid_rows`` outputs, sized for all M*P rows; the first ``count`` entries are defined."""
    row_idx: torch.Tensor          # [M*P] i32
    row2k: torch.Tensor            # [M*P] i32
    rows: torch.Tensor             # [M*P, C] f32
    count: torch.Tensor            # [1] i32 on the device

    def k(self) -> int:
        return int(self.count.item())   # host sync (the reference's boolean index syncs here too)


def valid_rows(voxels: torch.Tensor, num_points: torch.Tensor, mask_mode: str = "rowsum",
               m_dev: Optional[torch.Tensor] = None) -> ValidRows:
    dev = _need_cuda(voxels, num_points)
    voxels = _f32c(voxels)
    num_points = num_points.to(torch.int32).contiguous()
    M, P, C = voxels.shape
    L = _lib.lib()
    R = M * P
    i32 = dict(dtype=torch.int32, device=dev)
    out = ValidRows(torch.empty((R,), **i32), torch.empty((R,), **i32),
                    torch.empty((R, C), dtype=torch.float32, device=dev), torch.zeros((1,), **i32))
    nbytes = L.rpc_valid_rows_workspace_bytes(M, P)
    ws = _default_ws.get(nbytes, dev)
    st = L.rpc_valid_rows(_ptr(voxels), _ptr(num_points), M, P, C, MASK_MODES[mask_mode], _ptr(m_dev),
                          _ptr(out.row_idx), _ptr(out.row2k), _ptr(out.rows), _ptr(out.count),
                          ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(st, "rpc_valid_rows")
    return out


def scatter_rows(voxels: torch.Tensor, row2k: torch.Tensor, rows: torch.Tensor) -> torch.Tensor:
    dev = _need_cuda(voxels, row2k, rows)
    voxels, rows = _f32c(voxels), _f32c(rows)
    M, P, C = voxels.shape
    if rows.numel() == 0:                       # no valid row: the copy alone
        return voxels.clone()
    out = torch.empty_like(voxels)
    st = _lib.lib().rpc_scatter_rows(_ptr(voxels), _ptr(row2k), _ptr(rows), M, P, C, _ptr(out), _stream(dev))
    _lib.check(st, "rpc_scatter_rows")
    return out


def gather_rows(src: torch.Tensor, row_idx: torch.Tensor, K: int, C: int,
                k_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    dev = _need_cuda(src, row_idx)
    src = _f32c(src)
    out = torch.empty((K, C), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_gather_rows(_ptr(src), _ptr(row_idx), K, _ptr(k_dev), C, _ptr(out), _stream(dev))
    _lib.check(st, "rpc_gather_rows")
    return out


def perturb_apply_fwd(x: torch.Tensor, raw: torch.Tensor, bounds: PerturbCfg, want_p: bool = True):
    dev = _need_cuda(x, raw)
    x, raw = _f32c(x), _f32c(raw)
    K, C = x.shape
    y = torch.empty_like(x)
    p = torch.empty_like(x) if want_p else None
    st = _lib.lib().rpc_perturb_apply_fwd(_ptr(x), _ptr(raw), ctypes.byref(bounds), K, None, C, _ptr(y), _ptr(p),
                                          _stream(dev))
    _lib.check(st, "rpc_perturb_apply_fwd")
    return y, p


def perturb_apply_bwd(grad_p: torch.Tensor, raw: torch.Tensor, bounds: PerturbCfg) -> torch.Tensor:
    dev = _need_cuda(grad_p, raw)
    grad_p, raw = _f32c(grad_p), _f32c(raw)
    K, C = raw.shape
    g = torch.empty_like(raw)
    st = _lib.lib().rpc_perturb_apply_bwd(_ptr(grad_p), _ptr(raw), ctypes.byref(bounds), K, None, C, _ptr(g),
                                          _stream(dev))
    _lib.check(st, "rpc_perturb_apply_bwd")
    return g


# ------------------------------------------------------------------ attention gate (first slice of SURVEY 8f row 2)
def attention_gate_fwd(xn: torch.Tensor, t: torch.Tensor, w1: torch.Tensor, b1: torch.Tensor, w2: torch.Tensor,
                       b2: torch.Tensor, want_gate: bool = False):
    """``t * sigmoid(Linear(relu(Linear(xn))))`` -- ``raw_perturbations * self.attention(xn)`` of
    /root/reference/models/adversarial/voxel_perturber.py:203-205 in one pass.  Returns ``out`` (and the gate [K, 1])."""
    dev = _need_cuda(xn, t, w1, b1, w2, b2)
    xn, t = _f32c(xn), _f32c(t)
    w1, b1, w2, b2 = _f32c(w1), _f32c(b1), _f32c(w2), _f32c(b2)
    K, C = xn.shape
    mid = max(C // 2, 1)
    if t.shape != xn.shape or tuple(w1.shape) != (mid, C) or b1.numel() != mid or w2.numel() != mid or b2.numel() != 1:
        raise ValueError("attention_gate: parameter shapes do not match Linear(C, C//2) -> Linear(C//2, 1)")
    out = torch.empty_like(t)
    gate = torch.empty((K, 1), dtype=torch.float32, device=dev) if want_gate else None
    st = _lib.lib().rpc_attention_gate_fwd(_ptr(xn), _ptr(t), _ptr(w1), _ptr(b1), _ptr(w2), _ptr(b2), K, C, _ptr(out),
                                           _ptr(gate), _stream(dev))
    _lib.check(st, "rpc_attention_gate_fwd")
    return (out, gate) if want_gate else out


def attention_gate_bwd(g_out: torch.Tensor, xn: torch.Tensor, t: torch.Tensor, w1: torch.Tensor, b1: torch.Tensor,
                       w2: torch.Tensor, b2: torch.Tensor, want_grad_xn: bool = False):
    """-> (grad_t, grad_xn or None, grad_w1, grad_b1, grad_w2, grad_b2); the gate is recomputed, parameter gradients
    are reduced in fp64."""
    dev = _need_cuda(g_out, xn, t, w1, b1, w2, b2)
    g_out, xn, t = _f32c(g_out), _f32c(xn), _f32c(t)
    w1, b1, w2, b2 = _f32c(w1), _f32c(b1), _f32c(w2), _f32c(b2)
    K, C = xn.shape
    grad_t = torch.empty_like(t)
    grad_xn = torch.empty_like(xn) if want_grad_xn else None
    # the four parameter gradients are views of one buffer (the caller clamps them in one pass)
    flat = torch.empty((sum(q.numel() for q in (w1, b1, w2, b2)),), dtype=torch.float32, device=dev)
    views, o = [], 0
    for q in (w1, b1, w2, b2):
        views.append(flat[o:o + q.numel()].view(q.shape))
        o += q.numel()
    gw1, gb1, gw2, gb2 = views
    scratch = _scratch64(_lib.lib().rpc_attention_scratch_bytes() // 8, dev)
    st = _lib.lib().rpc_attention_gate_bwd(_ptr(g_out), _ptr(xn), _ptr(t), _ptr(w1), _ptr(b1), _ptr(w2), _ptr(b2), K, C,
                                           _ptr(scratch), _ptr(grad_t), _ptr(grad_xn), _ptr(gw1), _ptr(gb1), _ptr(gw2),
                                           _ptr(gb2), _stream(dev))
    _lib.check(st, "rpc_attention_gate_bwd")
    attention_gate_bwd.flat = flat                                # (the buffer behind the four views, for the one-pass clamp)
    return grad_t, grad_xn, gw1, gb1, gw2, gb2


def linear_wgrad(g: torch.Tensor, a: torch.Tensor, want_bias: bool = True, bn=None):
    """(grad_weight [M, N], grad_bias [M] or None) of ``out = a @ W.T + b`` from ``g = dL/d(out)`` [K, M] and the layer
    input ``a`` [K, N]; M, N <= 128 (the Linear layers of /root/reference/models/adversarial/voxel_perturber.py:82-101).
    ``bn=(mean, invstd, gamma, beta)``: ``a`` is the previous layer's pre-activation and the input is recomputed as
    relu(batch_norm(a)) while it is staged (``rpc_linear_wgrad_bn``).  128 x 64 / 64 x 128 layers run on the tensor cores."""
    dev = _need_cuda(g, a)
    g, a = _f32c(g), _f32c(a)
    K, M = g.shape
    N = a.shape[1]
    if a.shape[0] != K:
        raise ValueError("linear_wgrad: g and a must have the same number of rows")
    gw = torch.empty((M, N), dtype=torch.float32, device=dev)
    gb = torch.empty((M,), dtype=torch.float32, device=dev) if want_bias else None
    b4 = [None] * 4 if bn is None else [_f32c(t) for t in bn]
    st = _lib.lib().rpc_linear_wgrad_bn(_ptr(g), _ptr(a), _ptr(b4[0]), _ptr(b4[1]), _ptr(b4[2]), _ptr(b4[3]), K, M, N, _ptr(gw),
                                        _ptr(gb), _stream(dev))
    _lib.check(st, "rpc_linear_wgrad_bn")
    return gw, gb


def linear_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, transposed: bool = False,
               bn: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]] = None, tanh: bool = False,
               moments: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``f(x) @ B.T + bias`` for x [K, NI], NI, NO <= 128, on the tensor cores as fp32-accurate 3 x TF32
    (``rpc_linear_fwd``; the Linear layers of /root/reference/models/adversarial/voxel_perturber.py:82-101 and the input
    gradients autograd derives from them).  ``transposed=False``: B = w [NO, NI] (forward, nn.Linear weight);
    ``transposed=True``: B = w.T with w [NI, NO] (input gradient: x = dL/dy, w = the same weight tensor).
    ``bn=(mean, invstd, gamma, beta)`` applies the previous layer's BatchNorm + ReLU to x while it is loaded;
    ``tanh`` applies the trunk's final Tanh; ``moments`` (float64 [2 * NO], zeroed by the call) receives the column sums
    and sums of squares of the result -- what the next BatchNorm needs (``bn_finalize``)."""
    dev = _need_cuda(x, w, bias, moments, out)
    x, w = _f32c(x), _f32c(w)
    K, NI = x.shape
    NO = w.shape[1] if transposed else w.shape[0]
    if (w.shape[0] if transposed else w.shape[1]) != NI:
        raise ValueError("linear_fwd: weight does not match the input width")
    y = out if out is not None else torch.empty((K, NO), dtype=torch.float32, device=dev)
    b4 = [None] * 4 if bn is None else [_f32c(t) for t in bn]
    st = _lib.lib().rpc_linear_fwd(_ptr(x), K, None, NI, NO, _ptr(w), int(transposed), _ptr(None if bias is None else _f32c(bias)),
                                   _ptr(b4[0]), _ptr(b4[1]), _ptr(b4[2]), _ptr(b4[3]), int(tanh), _ptr(moments), _ptr(y),
                                   _stream(dev))
    _lib.check(st, "rpc_linear_fwd")
    return y


def bn_finalize(moments: torch.Tensor, K: int, eps: float, momentum: float, running_mean: Optional[torch.Tensor] = None,
                running_var: Optional[torch.Tensor] = None):
    """(save_mean, save_invstd) of a training-mode BatchNorm1d from the column moments ``linear_fwd`` accumulated;
    running statistics updated in place like nn.BatchNorm1d."""
    dev = _need_cuda(moments, running_mean, running_var)
    H = moments.numel() // 2
    mean = torch.empty((H,), dtype=torch.float32, device=dev)
    invstd = torch.empty((H,), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_bn_finalize(_ptr(moments), K, None, H, float(eps), float(momentum), _ptr(running_mean), _ptr(running_var),
                                    _ptr(mean), _ptr(invstd), _stream(dev))
    _lib.check(st, "rpc_bn_finalize")
    return mean, invstd


def bn_relu_apply(y: torch.Tensor, mean: torch.Tensor, invstd: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``relu(batch_norm(y))`` with the saved statistics (the backward's recomputation of a layer input)."""
    dev = _need_cuda(y, mean, invstd, gamma, beta, out)
    y = _f32c(y)
    K, H = y.shape
    z = out if out is not None else torch.empty_like(y)
    st = _lib.lib().rpc_bn_relu_apply(_ptr(y), K, None, H, _ptr(mean), _ptr(invstd), _ptr(_f32c(gamma)), _ptr(_f32c(beta)),
                                      _ptr(z), _stream(dev))
    _lib.check(st, "rpc_bn_relu_apply")
    return z


def _ptr_array(tensors):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def trunk_desc(C: int, widths: Sequence[int], weights, biases, gammas, betas, running_means, running_vars, eps, momentum):
    """``rpc_trunk_desc`` of a trunk with ``len(widths)`` hidden layers; the tensors must stay alive while it is used."""
    d = _lib.TrunkDesc()
    d.C, d.nl = int(C), len(widths)
    for l, h in enumerate(widths):
        d.width[l] = int(h)
        d.gamma[l], d.beta[l] = gammas[l].data_ptr(), betas[l].data_ptr()
        d.running_mean[l] = None if running_means[l] is None else running_means[l].data_ptr()
        d.running_var[l] = None if running_vars[l] is None else running_vars[l].data_ptr()
        d.eps[l], d.momentum[l] = float(eps[l]), float(momentum[l])
    for l in range(len(widths) + 1):
        d.w[l], d.b[l] = weights[l].data_ptr(), biases[l].data_ptr()
    return d


def trunk_fwd(desc, xn: torch.Tensor, widths: Sequence[int]):
    """Training-mode forward of the whole trunk (``rpc_trunk_fwd``): -> (raw [K, C], ybuf, ys, stats).  ``ys[l]`` [K, widths[l]]
    are the pre-activations (views of the flat ``ybuf``), ``stats`` the per-layer (mean, invstd) the backward needs."""
    dev = _need_cuda(xn)
    K, C = xn.shape
    tot = sum(widths)
    ybuf = torch.empty((K * tot,), dtype=torch.float32, device=dev)
    ys, o = [], 0
    for h in widths:
        ys.append(ybuf[o:o + K * h].view(K, h))
        o += K * h
    stats = torch.empty((2 * tot,), dtype=torch.float32, device=dev)
    scratch = _scratch64(2 * tot, dev)
    raw = torch.empty((K, C), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_trunk_fwd(ctypes.byref(desc), _ptr(xn), K, _ptr_array(ys), _ptr(stats), _ptr(scratch), _ptr(raw), _stream(dev))
    _lib.check(st, "rpc_trunk_fwd")
    return raw, ybuf, ys, stats


def trunk_fwd_eval(desc, xn: torch.Tensor, widths: Sequence[int], stats: torch.Tensor) -> torch.Tensor:
    """Eval-mode forward of the whole trunk (``rpc_trunk_fwd_eval``): ``stats`` = per layer (running_mean, 1 / sqrt(running_var
    + eps)) concatenated; -> raw [K, C].  No autograd."""
    dev = _need_cuda(xn, stats)
    K, C = xn.shape
    work = torch.empty((2 * K * max(max(widths), C),), dtype=torch.float32, device=dev)
    raw = torch.empty((K, C), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_trunk_fwd_eval(ctypes.byref(desc), _ptr(xn), K, _ptr(stats), _ptr(work), _ptr(raw), _stream(dev))
    _lib.check(st, "rpc_trunk_fwd_eval")
    return raw


def trunk_bwd(desc, xn: torch.Tensor, widths: Sequence[int], ys, stats: torch.Tensor, raw: torch.Tensor, g_raw: torch.Tensor,
              param_shapes, want_grad_xn: bool = False):
    """Backward of ``trunk_fwd`` (``rpc_trunk_bwd``): -> (grad_w list, grad_b list, grad_gamma list, grad_beta list, grad_xn,
    flat) where ``flat`` is the one buffer all parameter gradients are views of."""
    dev = _need_cuda(xn, g_raw)
    K, C = xn.shape
    g_raw = _f32c(g_raw)
    maxw = max(max(widths), C)
    work = torch.empty((2 * K * maxw,), dtype=torch.float32, device=dev)
    scratch = _scratch64(2 * 128, dev)
    nl = len(widths)
    n_w = sum(a * b for a, b in param_shapes)
    flat = torch.empty((n_w + sum(a for a, _ in param_shapes) + 2 * sum(widths),), dtype=torch.float32, device=dev)
    gw, gb, gg, gbe, o = [], [], [], [], 0
    for a, b in param_shapes:
        gw.append(flat[o:o + a * b].view(a, b)); o += a * b
    for a, _ in param_shapes:
        gb.append(flat[o:o + a]); o += a
    for h in widths:
        gg.append(flat[o:o + h]); o += h
    for h in widths:
        gbe.append(flat[o:o + h]); o += h
    gxn = torch.empty_like(xn) if want_grad_xn else None
    st = _lib.lib().rpc_trunk_bwd(ctypes.byref(desc), _ptr(xn), K, _ptr_array(ys), _ptr(stats), _ptr(raw), _ptr(g_raw), _ptr(work),
                                  _ptr(scratch), _ptr_array(gw), _ptr_array(gb), _ptr_array(gg), _ptr_array(gbe), _ptr(gxn),
                                  _stream(dev))
    _lib.check(st, "rpc_trunk_bwd")
    assert nl == len(gg)
    return gw, gb, gg, gbe, gxn, flat


def bn_relu_fwd(y: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float, momentum: float,
                running_mean: Optional[torch.Tensor] = None, running_var: Optional[torch.Tensor] = None):
    """Training-mode ``relu(batch_norm(y))`` for y [K, H], H <= 128 (nn.BatchNorm1d -> nn.ReLU of
    /root/reference/models/adversarial/voxel_perturber.py:84-98); running statistics updated in place.
    Returns (z, save_mean, save_invstd)."""
    dev = _need_cuda(y, gamma, beta, running_mean, running_var)
    y = _f32c(y)
    K, H = y.shape
    z = torch.empty_like(y)
    mean = torch.empty((H,), dtype=torch.float32, device=dev)
    invstd = torch.empty((H,), dtype=torch.float32, device=dev)
    scratch = _scratch64(_lib.lib().rpc_bn_scratch_bytes() // 8, dev)
    st = _lib.lib().rpc_bn_relu_fwd(_ptr(y), K, H, _ptr(_f32c(gamma)), _ptr(_f32c(beta)), float(eps), float(momentum),
                                    _ptr(running_mean), _ptr(running_var), _ptr(scratch), _ptr(mean), _ptr(invstd),
                                    _ptr(z), _stream(dev))
    _lib.check(st, "rpc_bn_relu_fwd")
    return z, mean, invstd


def linear_dgrad_bnred(g: torch.Tensor, w: torch.Tensor, v: torch.Tensor, bn):
    """Input gradient ``g @ w`` (w [NI, NO], the nn.Linear weight) whose epilogue also reduces the backward of the
    BatchNorm + ReLU below (``v`` [K, NO] = that layer's pre-activation, ``bn`` = its (mean, invstd, gamma, beta)):
    -> (grad_in [K, NO], sums float64 [2 * NO]) for ``bn_relu_bwd(..., sums=sums)`` (``rpc_linear_dgrad_bnred``)."""
    dev = _need_cuda(g, w, v)
    g, w, v = _f32c(g), _f32c(w), _f32c(v)
    K, NI = g.shape
    NO = w.shape[1]
    out = torch.empty((K, NO), dtype=torch.float32, device=dev)
    sums = torch.empty((2 * NO,), dtype=torch.float64, device=dev)
    b4 = [_f32c(t) for t in bn]
    st = _lib.lib().rpc_linear_dgrad_bnred(_ptr(g), K, NI, NO, _ptr(w), _ptr(v), _ptr(b4[0]), _ptr(b4[1]), _ptr(b4[2]), _ptr(b4[3]),
                                           _ptr(sums), _ptr(out), _stream(dev))
    _lib.check(st, "rpc_linear_dgrad_bnred")
    return out, sums


def bn_relu_bwd(g: torch.Tensor, y: torch.Tensor, mean: torch.Tensor, invstd: torch.Tensor, gamma: torch.Tensor,
                beta: torch.Tensor, sums: Optional[torch.Tensor] = None):
    """-> (grad_y, grad_gamma, grad_beta) of ``bn_relu_fwd``; the ReLU mask is recomputed from y.  ``sums``: the reductions
    already made by the producer of g (``linear_dgrad_bnred``): the reduction pass is skipped."""
    dev = _need_cuda(g, y, mean, invstd, gamma, beta)
    g, y = _f32c(g), _f32c(y)
    K, H = y.shape
    gy = torch.empty_like(y)
    gg = torch.empty((H,), dtype=torch.float32, device=dev)
    gb = torch.empty((H,), dtype=torch.float32, device=dev)
    scratch = _scratch64(_lib.lib().rpc_bn_scratch_bytes() // 8, dev)
    if sums is not None:
        st = _lib.lib().rpc_bn_relu_bwd_sums(_ptr(g), _ptr(y), _ptr(mean), _ptr(invstd), _ptr(_f32c(gamma)), _ptr(_f32c(beta)), K, H,
                                             _ptr(sums), _ptr(gy), _ptr(gg), _ptr(gb), _stream(dev))
        _lib.check(st, "rpc_bn_relu_bwd_sums")
        return gy, gg, gb
    st = _lib.lib().rpc_bn_relu_bwd(_ptr(g), _ptr(y), _ptr(mean), _ptr(invstd), _ptr(_f32c(gamma)), _ptr(_f32c(beta)), K, H,
                                    _ptr(scratch), _ptr(gy), _ptr(gg), _ptr(gb), _stream(dev))
    _lib.check(st, "rpc_bn_relu_bwd")
    return gy, gg, gb


# ------------------------------------------------------------------ perturber statistics
def normalize_fwd(x: torch.Tensor) -> torch.Tensor:
    """``clamp(x / (std(x, 0) + 1e-6), -10, 10)`` with the NaN / Inf scale guard, no host sync
    (/root/reference/models/adversarial/voxel_perturber.py:157-168)."""
    dev = _need_cuda(x)
    x = _f32c(x)
    K, C = x.shape
    xn = torch.empty_like(x)
    mom = _scratch64(_lib.lib().rpc_moments_bytes() // 8, dev)
    st = _lib.lib().rpc_normalize_fwd(_ptr(x), K, None, C, _ptr(mom), _ptr(xn), _stream(dev))
    _lib.check(st, "rpc_normalize_fwd")
    return xn


def perturb_losses_fwd(p: torch.Tensor):
    """(losses[4] = l2_norm, intensity_loss, bias_loss, imbalance_loss; consts[32]) of voxel_perturber.py:268-299."""
    dev = _need_cuda(p)
    p = _f32c(p)
    K, C = p.shape
    mom = _scratch64(_lib.lib().rpc_moments_bytes() // 8, dev)
    losses = torch.empty((4,), dtype=torch.float32, device=dev)
    consts = torch.empty((32,), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_perturb_losses_fwd(_ptr(p), K, None, C, _ptr(mom), _ptr(losses), _ptr(consts), _stream(dev))
    _lib.check(st, "rpc_perturb_losses_fwd")
    return losses, consts


def perturb_losses_bwd(p: torch.Tensor, consts: torch.Tensor, grad_losses: torch.Tensor) -> torch.Tensor:
    dev = _need_cuda(p, consts, grad_losses)
    p, grad_losses = _f32c(p), _f32c(grad_losses)
    K, C = p.shape
    g = torch.empty_like(p)
    st = _lib.lib().rpc_perturb_losses_bwd(_ptr(p), _ptr(consts), _ptr(grad_losses), K, None, C, _ptr(g), _stream(dev))
    _lib.check(st, "rpc_perturb_losses_bwd")
    return g


def voxel_centers(coors: torch.Tensor, cfg: VoxelCfg, m_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``voxel_centers`` of mmdet3d's ``Det3DDataPreprocessor.voxelize``: [M, 3] (x, y, z) centres of the voxels whose
    coors rows end in (z, y, x)."""
    dev = _need_cuda(coors)
    coors = coors.contiguous()
    M, ncol = coors.shape
    out = torch.empty((M, 3), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_voxel_centers(ctypes.byref(cfg), _ptr(coors), ncol, M, _ptr(m_dev), _ptr(out), _stream(dev))
    _lib.check(st, "rpc_voxel_centers")
    return out


def momentum_apply_fwd(x: torch.Tensor, y: torch.Tensor, last: Optional[torch.Tensor], scaling: float, alpha: float,
                       y_is_delta: bool = False):
    """(out = x + sp, sp = (y - x) * scaling [+ alpha * last], l2 = ||sp||_2 as a 0-d device tensor);
    /root/reference/models/detectors/strong_adversarial_voxelnet.py:156-180."""
    dev = _need_cuda(x, y, last)
    x, y = _f32c(x), _f32c(y)
    last = _f32c(last) if last is not None else None
    out, sp = torch.empty_like(x), torch.empty_like(x)
    sumsq = torch.empty((1,), dtype=torch.float64, device=dev)
    l2 = torch.empty((), dtype=torch.float32, device=dev)
    st = _lib.lib().rpc_momentum_apply_fwd(_ptr(x), _ptr(y), _ptr(last), int(y_is_delta), scaling, alpha, x.numel(), _ptr(out), _ptr(sp),
                                           _ptr(sumsq), _ptr(l2), _stream(dev))
    _lib.check(st, "rpc_momentum_apply_fwd")
    return out, sp, l2


def momentum_apply_bwd(g_out: Optional[torch.Tensor], g_l2: Optional[torch.Tensor], sp: torch.Tensor, l2: torch.Tensor,
                       scaling: float, want_x: bool = True, want_y: bool = True, y_is_delta: bool = False):
    dev = _need_cuda(sp, l2)
    g_out = _f32c(g_out) if g_out is not None else None
    g_l2 = _f32c(g_l2) if g_l2 is not None else None
    gx = torch.empty_like(sp) if want_x else None
    gy = torch.empty_like(sp) if want_y else None
    if want_x or want_y:
        st = _lib.lib().rpc_momentum_apply_bwd(_ptr(g_out), _ptr(g_l2), _ptr(sp), _ptr(l2), int(y_is_delta), scaling, sp.numel(),
                                               _ptr(gx), _ptr(gy), _stream(dev))
        _lib.check(st, "rpc_momentum_apply_bwd")
    return gx, gy


def points_range_filter(points: torch.Tensor, point_cloud_range: Sequence[float]):
    """mmdet3d ``PointsRangeFilter``: rows with ``lo < (x, y, z) < hi`` in their original order.  Returns
    ``(out [N, C], index [N] int32, count [1] int32)``: the first ``count`` rows of ``out`` / ``index`` are the kept
    points and their source rows (no host sync; slice with ``int(count)`` when a tight tensor is needed)."""
    dev = _need_cuda(points)
    points = _f32c(points)
    N, C = points.shape
    L = _lib.lib()
    out = torch.empty_like(points)
    index = torch.empty((N,), dtype=torch.int32, device=dev)
    count = torch.zeros((1,), dtype=torch.int32, device=dev)
    rng = (ctypes.c_float * 6)(*[float(v) for v in point_cloud_range])
    nbytes = L.rpc_valid_rows_workspace_bytes(N, 1)
    ws = _default_ws.get(nbytes, dev)
    st = L.rpc_points_range_filter(_ptr(points), N, C, rng, _ptr(out), _ptr(index), _ptr(count), ws.data_ptr(),
                                   ws.numel(), _stream(dev))
    _lib.check(st, "rpc_points_range_filter")
    return out, index, count
