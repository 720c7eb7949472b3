"""Seeded synthetic point clouds of the BASELINE.json shapes: there is no network for
KITTI / nuScenes, so benches and tests use these.  numpy only; fp32 output.

``lidar``     spinning-LiDAR-like sweep: beams x azimuth columns, ground plane + box obstacles,
              range noise, 0.5 % null returns (0,0,0,0) -- realistic occupancy (several points per
              voxel near the sensor, many out-of-range points for KITTI's forward-only range).
``uniform``   xyz uniform over the range expanded by 10 % (almost every point its own voxel).
``clustered`` Gaussian blobs (many points per voxel: exercises the max_num_points overflow).
"""
from __future__ import annotations

from typing import Dict, Sequence

import os

import numpy as np

KITTI = dict(name="kitti", voxel_size=[0.05, 0.05, 0.1], point_cloud_range=[0, -40, -3, 70.4, 40, 1],
             max_num_points=5, max_voxels=(16000, 40000), C=4, N=120000, beams=64,
             elev=(-24.8, 2.0), sensor_h=1.73, sweeps=1)
NUSCENES = dict(name="nuscenes", voxel_size=[0.075, 0.075, 0.2], point_cloud_range=[-54, -54, -5, 54, 54, 3],
                max_num_points=10, max_voxels=(120000, 160000), C=5, N=300000, beams=32,
                elev=(-30.0, 10.0), sensor_h=1.84, sweeps=10)
PRESETS: Dict[str, dict] = {"kitti": KITTI, "nuscenes": NUSCENES}


def _lidar(rng: np.random.Generator, N: int, C: int, beams: int, elev, sensor_h: float, sweeps: int) -> np.ndarray:
    per = -(-N // sweeps)
    out = []
    for s in range(sweeps):
        cols = -(-per // beams)
        el = np.deg2rad(np.linspace(elev[0], elev[1], beams))[:, None]
        az = (np.arange(cols)[None, :] + rng.uniform(0, 1)) * (2 * np.pi / cols)
        el = np.broadcast_to(el, (beams, cols)).reshape(-1)
        az = np.broadcast_to(az, (beams, cols)).reshape(-1)
        with np.errstate(divide="ignore"):
            r_ground = np.where(el < 0, sensor_h / np.tan(-el), np.inf)
        r = np.minimum(r_ground, 120.0)
        obj = rng.uniform(size=r.shape) < 0.15
        r = np.where(obj, np.minimum(r, rng.uniform(5, 60, size=r.shape)), r)
        r = r + rng.normal(0, 0.02, size=r.shape)
        x = r * np.cos(el) * np.cos(az)
        y = r * np.cos(el) * np.sin(az)
        z = r * np.sin(el)
        pts = np.zeros((r.shape[0], C), np.float32)
        shift = rng.uniform(-0.5, 0.5, size=2) if sweeps > 1 else np.zeros(2)
        pts[:, 0] = x + shift[0]
        pts[:, 1] = y + shift[1]
        pts[:, 2] = z
        pts[:, 3] = rng.uniform(0, 1, size=r.shape)
        if C > 4:
            pts[:, 4] = 0.05 * s
        # column-major firing order (all beams of one azimuth, then the next), like a real spin
        pts = pts.reshape(beams, cols, C).transpose(1, 0, 2).reshape(-1, C)[:per]
        out.append(pts)
    pts = np.concatenate(out)[:N]
    null = rng.uniform(size=pts.shape[0]) < float(os.environ.get("RPC_SYNTH_NULL", "0.005"))
    pts[null] = 0.0
    return np.ascontiguousarray(pts, np.float32)


def make_cloud(preset: str = "kitti", N: int | None = None, dist: str = "lidar", seed: int = 1234,
               C: int | None = None) -> np.ndarray:
    cfg = PRESETS[preset]
    N = cfg["N"] if N is None else int(N)
    C = cfg["C"] if C is None else int(C)
    rng = np.random.default_rng(seed)
    lo = np.array(cfg["point_cloud_range"][:3], np.float64)
    hi = np.array(cfg["point_cloud_range"][3:], np.float64)
    if N == 0:
        return np.zeros((0, C), np.float32)
    if dist == "lidar":
        return _lidar(rng, N, C, cfg["beams"], cfg["elev"], cfg["sensor_h"], cfg["sweeps"])
    pts = np.zeros((N, C), np.float32)
    if dist == "uniform":
        ext = (hi - lo) * 0.05
        pts[:, :3] = rng.uniform(lo - ext, hi + ext, size=(N, 3))
    elif dist == "clustered":
        cen = rng.uniform(lo, hi, size=(1000, 3))
        pts[:, :3] = cen[rng.integers(0, 1000, N)] + rng.normal(0, 0.3, size=(N, 3))
    else:
        raise ValueError(dist)
    pts[:, 3:] = rng.uniform(0, 1, size=(N, C - 3))
    return pts


def make_batch(preset: str, B: int, N: int | None = None, dist: str = "lidar", seed: int = 1234):
    """list of B clouds, seeds ``seed + sample_idx``"""
    return [make_cloud(preset, N, dist, seed + i) for i in range(B)]


def make_raw(points: np.ndarray, seed: int = 99, amp: float = 1.5) -> np.ndarray:
    """stand-in for the perturber MLP output (tanh * attention, in [-1,1]); ``amp`` > 1 makes part of
    the rows hit the clamp so that the masked backward is exercised."""
    rng = np.random.default_rng(seed)
    return rng.uniform(-amp, amp, size=points.shape).astype(np.float32)
