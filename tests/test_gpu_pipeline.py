"""StepPipeline (pinned host buffers in and out, three streams, double buffering) must return exactly what the
direct C-ABI calls return, step after step, in both the hard and the dynamic configuration."""
import numpy as np
import pytest
import torch

from oracle import ref_path as orc
from robustpointclouds_b200 import PerturbCfg, StepPipeline, VoxelCfg, ops, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dynamic", [False, True])
@pytest.mark.parametrize("use_graph", [False, True])
def test_pipeline_matches_direct_calls(dynamic, use_graph):
    dev = torch.device("cuda:0")
    k = synth.KITTI
    C, P, V = k["C"], k["max_num_points"], 3000
    n = [7001, 5000]
    cfg = VoxelCfg.make(k["voxel_size"], k["point_cloud_range"], P, V, C, C)
    eb, fb = orc.perturber_bounds(C, True)
    bounds = PerturbCfg.make(eb, fb)
    pipe = StepPipeline(cfg, bounds, n, dev, depth=2, dynamic=dynamic, use_graph=use_graph)
    cap = sum(n) if dynamic else sum(min(x, V) for x in n)
    gf = torch.randn((cap, C), device=dev, generator=torch.Generator(dev).manual_seed(3))
    batches = []
    for s in range(5):
        clouds = [synth.make_cloud("kitti", m, "lidar", 50 + 10 * s + i) for i, m in enumerate(n)]
        raws = [synth.make_raw(c, 70 + 10 * s + i) for i, c in enumerate(clouds)]
        if s % 2:                                            # the halves of one pinned buffer: moved by one copy
            hp, hr = pipe.alloc_host_inputs()
            hp.copy_(torch.from_numpy(np.concatenate(clouds)))
            hr.copy_(torch.from_numpy(np.concatenate(raws)))
            batches.append((hp, hr))
        else:                                                # two unrelated pinned tensors: two copies
            batches.append((torch.from_numpy(np.concatenate(clouds)).pin_memory(),
                            torch.from_numpy(np.concatenate(raws)).pin_memory()))
    tickets, got = [], []
    for s, (hp, hr) in enumerate(batches):
        tickets.append(pipe.submit(hp, hr, gf))
        if s >= 1:                                           # results are read one step behind the submissions
            got.append({k2: v.clone() for k2, v in pipe.result(tickets[s - 1]).items()})
    got.append({k2: v.clone() for k2, v in pipe.result(tickets[-1]).items()})
    with pytest.raises(ValueError):
        pipe.result(tickets[0])
    for (hp, hr), res in zip(batches, got):
        pts, raw = hp.to(dev), hr.to(dev)
        if dynamic:
            ref = ops.dynamic_scatter_mean(pts, n, cfg, raw=raw, bounds=bounds)
            names = dict(coors=ref.coors, voxel_coors=ref.voxel_coors, num_points=ref.num_points, feats=ref.feats)
        else:
            ref = ops.hard_voxelize(pts, n, cfg, raw=raw, bounds=bounds, want_feats=True, want_maps=True)
            names = dict(voxels=ref.voxels, coors=ref.coors, num_points=ref.num_points, feats=ref.feats)
        graw = ops.step_bwd(gf, ref.num_points, ref.point2voxel, cfg, raw=raw, bounds=bounds)
        M = int(ref.voxel_num[-1])
        assert res["voxel_num"].tolist() == ref.voxel_num.cpu().tolist()
        for name, t in names.items():
            a, b = res[name], t.cpu()
            if name != "coors" or not dynamic:
                a, b = a[:M], b[:M]
            if name == "feats" and dynamic:                  # red.add order differs from run to run
                torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-4)
            else:
                assert torch.equal(a, b), name
        if dynamic:
            torch.testing.assert_close(res["grad_raw"], graw.cpu(), rtol=1e-5, atol=1e-6)
        else:
            assert torch.equal(res["grad_raw"], graw.cpu())


def test_pipeline_lean_host_fields():
    """host_fields=LEAN_FIELDS: the padded voxels stay on the device, everything else arrives unchanged"""
    dev = torch.device("cuda:0")
    k = synth.KITTI
    C, P, V = k["C"], k["max_num_points"], 3000
    n = [6001, 4000]
    cfg = VoxelCfg.make(k["voxel_size"], k["point_cloud_range"], P, V, C, C)
    eb, fb = orc.perturber_bounds(C, True)
    bounds = PerturbCfg.make(eb, fb)
    full = StepPipeline(cfg, bounds, n, dev)
    lean = StepPipeline(cfg, bounds, n, dev, host_fields=StepPipeline.LEAN_FIELDS)
    assert lean.d2h_bytes < full.d2h_bytes - sum(min(x, V) for x in n) * P * C * 4 + 1024
    gf = torch.randn((sum(min(x, V) for x in n), C), device=dev, generator=torch.Generator(dev).manual_seed(5))
    hp, hr = full.alloc_host_inputs()
    hp.copy_(torch.from_numpy(np.concatenate([synth.make_cloud("kitti", m, "lidar", 90 + i) for i, m in enumerate(n)])))
    hr.copy_(torch.from_numpy(synth.make_raw(hp.numpy(), 91)))
    a = full.result(full.submit(hp, hr, gf))
    b = lean.result(lean.submit(hp, hr, gf))
    assert "voxels" in a and "voxels" not in b
    for name in b:
        assert torch.equal(a[name], b[name]), name
    M = int(a["voxel_num"][-1])
    assert torch.equal(lean.outs[0].voxels[:M].cpu(), a["voxels"][:M])       # still produced, on the device
    with pytest.raises(ValueError):
        StepPipeline(cfg, bounds, n, dev, host_fields=("voxels", "nonsense"))


def test_pipeline_raw_on_device_gradient_on_device():
    """raw_on_device=True, grad_to_host=False: only points cross PCIe on the way in, the gradient stays where the
    perturber's backward reads it; every result equals the host-fed pipeline's, over several steps and both slots"""
    dev = torch.device("cuda:0")
    k = synth.KITTI
    C, P, V = k["C"], k["max_num_points"], 3000
    n = [6001, 4000]
    cfg = VoxelCfg.make(k["voxel_size"], k["point_cloud_range"], P, V, C, C)
    eb, fb = orc.perturber_bounds(C, True)
    bounds = PerturbCfg.make(eb, fb)
    full = StepPipeline(cfg, bounds, n, dev, host_fields=StepPipeline.LEAN_FIELDS)
    real = StepPipeline(cfg, bounds, n, dev, host_fields=StepPipeline.LEAN_FIELDS, raw_on_device=True, grad_to_host=False)
    assert real.h2d_bytes * 2 == full.h2d_bytes and real.d2h_bytes < full.d2h_bytes - sum(n) * C * 4 + 1024
    gf = torch.randn((sum(min(x, V) for x in n), C), device=dev, generator=torch.Generator(dev).manual_seed(5))
    hp, hr = full.alloc_host_inputs()
    for step in range(3):
        hp.copy_(torch.from_numpy(np.concatenate([synth.make_cloud("kitti", m, "lidar", 70 + 10 * step + i) for i, m in enumerate(n)])))
        hr.copy_(torch.from_numpy(synth.make_raw(hp.numpy(), 71 + step)))
        real.set_raw(hr.to(dev))
        a = full.result(full.submit(hp, hr, gf))
        t = real.submit(hp, None, gf)
        b = real.result(t)
        assert "grad_raw" in a and "grad_raw" not in b
        for name in b:
            assert torch.equal(a[name], b[name]), name
        assert torch.equal(real.graw[t % real.depth].cpu(), a["grad_raw"])
    with pytest.raises(ValueError):
        real.submit(hp, hr, gf)
    with pytest.raises(ValueError):
        full.submit(hp, None, gf)
