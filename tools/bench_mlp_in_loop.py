"""SURVEY 8d, config 2 "one run with the real MLP in the loop reported separately": the REFERENCE-ORDER training step
of the adversarial input path with the real VoxelPerturber (MLP + attention, training-mode BatchNorm) instead of the
U(-1, 1) stand-in --

    voxelize (hard, whole batch)                        Det3DDataPreprocessor.voxelize, adversarial_voxelnet.py:66-71
    valid rows -> VoxelPerturber -> scatter back        extract_feat, adversarial_voxelnet.py:81-117
    HardSimpleVFE                                       :135-137
    loss = <feats, w> + the four regularisers; backward to the perturber's parameters (hooks clamp to +-0.1)

timed with CUDA events per step (median of 30 after 5 warm-ups), three ways: every fused kernel of this library on
(`fused_stats`, `fused_attention`, `fused_wgrad`), the last one off, all off (the ATen formulas the reference runs around its MLP), and the split of
the fused run into voxelize / perturber / VFE+backward.  `trunk on tcgen05` runs the whole trunk through _TrunkFn (rpc_linear_fwd: 3 x TF32 tcgen05 products, BatchNorm fused into
the neighbouring layers); in the others its products are cuBLAS.
    python tools/bench_mlp_in_loop.py [kitti|nuscenes ...]"""
import os
import statistics
import sys

import numpy as np
import torch

import robustpointclouds_b200 as rpc
from robustpointclouds_b200 import synth
from robustpointclouds_b200.adversarial import extract_voxel_features
from robustpointclouds_b200.voxelization import voxelize

dev = torch.device("cuda:0")


def events(fn, n=int(os.environ.get("RPC_BENCH_N", 30)), warm=int(os.environ.get("RPC_BENCH_WARM", 5))):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    return statistics.median(ts)


for preset in sys.argv[1:] or ["kitti", "nuscenes"]:
    d = synth.PRESETS[preset]
    B = 6 if preset == "kitti" else 4
    C = d["C"]
    clouds = [torch.from_numpy(c).to(dev) for c in synth.make_batch(preset, B)]
    N = sum(int(c.shape[0]) for c in clouds)
    layer = rpc.Voxelization(d["voxel_size"], d["point_cloud_range"], d["max_num_points"], d["max_voxels"]).train()
    vfe = rpc.HardSimpleVFE(num_features=C)
    vd0 = voxelize(clouds, layer)
    M = int(vd0["voxels"].shape[0])
    w = torch.randn((M, C), device=dev, generator=torch.Generator(dev).manual_seed(1))
    res = {}
    variants = (("trunk on tcgen05", True, True, True, True), ("fused", True, True, True, False),
                ("fused, ATen BatchNorm + ReLU", True, True, False, False),
                ("fused, nn.Linear backward, ATen BatchNorm + ReLU", True, False, False, False), ("aten", False, False, False, False))
    if os.environ.get("RPC_BENCH_VARIANTS"):                        # e.g. "fused" for an ncu launch list of one variant
        variants = tuple(v for v in variants if v[0] in os.environ["RPC_BENCH_VARIANTS"].split(";"))
    for name, fused, wgrad, bn, trunk in variants:
        torch.manual_seed(0)
        adv = rpc.VoxelPerturber(hidden_channels=[64, 128, 64], voxel_size=d["voxel_size"], nan_guards="device",
                                 fused_stats=fused, fused_attention=fused, fused_wgrad=wgrad, fused_bn=bn,
                                 fused_trunk=trunk).build(C, dev).train()
        adv._return_loss_dict = True
        state = {}

        def t_vox():
            state["vd"] = voxelize(clouds, layer)

        def t_fwd():
            feats, ld = extract_voxel_features(state["vd"], vfe, adv, training=True, epoch=3)
            state["loss"] = (feats * w).sum() + sum(ld.values())

        def t_bwd():
            for q in adv.parameters():
                q.grad = None
            state["loss"].backward()

        def step():
            t_vox(); t_fwd(); t_bwd()

        step()
        K = int(rpc.ops.valid_rows(state["vd"]["voxels"], state["vd"]["num_points"], "rowsum").k())
        tot = events(step)
        t_vox()
        tv = events(t_vox)
        tf = events(lambda: (t_fwd(),))                              # forward only (graph is dropped each time)
        res[name] = tot
        adv.flush_metrics()
        print(f"{preset} B={B} N={N} M={M} K={K} [{name}]: step {tot * 1e3:.0f} us = {N / tot / 1e3:.1f} M points/s "
              f"(voxelize {tv * 1e3:.0f} us, perturb + VFE forward {tf * 1e3:.0f} us, backward {(tot - tv - tf) * 1e3:.0f} us)")
    if "aten" in res and "trunk on tcgen05" in res:
        print(f"{preset}: trunk on tcgen05 vs cuBLAS + per-module kernels: {res['fused'] / res['trunk on tcgen05']:.2f}x, vs ATen: {res['aten'] / res['trunk on tcgen05']:.2f}x")
    if "aten" in res and "fused" in res:
        print(f"{preset}: fused kernels around the MLP vs ATen formulas: {res['aten'] / res['fused']:.2f}x")
