"""GPU parity of the post-VFE perturbation variant: the fused momentum kernels through the C-ABI
against the oracle's restatement of strong_adversarial_voxelnet.py:114-192 and against the ATen expressions the
reference runs.  Elementwise outputs are bit-exact; the norm and the gradients within 1e-6 relative."""
import numpy as np
import pytest
import torch

from oracle import ref_path as orc
from robustpointclouds_b200 import EnhancedPerturbation, VoxelPerturber, ops

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


@pytest.mark.parametrize("M,C", [(1, 4), (1000, 4), (96000, 4), (160001, 5)])
@pytest.mark.parametrize("momentum", [False, True])
@pytest.mark.parametrize("delta", [False, True])
def test_momentum_kernels_vs_oracle(M, C, momentum, delta):
    rng = np.random.default_rng(M + C)
    x = rng.normal(size=(M, C)).astype(np.float32) * 10
    y = (rng.normal(size=(M, C)) * 0.1).astype(np.float32) if delta else (x + rng.normal(size=(M, C)) * 0.1).astype(np.float32)
    last = (rng.normal(size=(M, C)) * 0.1).astype(np.float32) if momentum else None
    s, a = 1.7, 0.9
    ro, rsp, rl2 = orc.enhanced_perturbation(x, y, last, s, a, delta)
    t = lambda v: None if v is None else torch.from_numpy(v).to(DEV)          # noqa: E731
    out, sp, l2 = ops.momentum_apply_fwd(t(x), t(y), t(last), s, a, delta)
    assert np.array_equal(out.cpu().numpy(), ro) and np.array_equal(sp.cpu().numpy(), rsp)
    assert float(l2) == pytest.approx(rl2, rel=1e-6)
    go = rng.normal(size=(M, C)).astype(np.float32)
    gx, gy = ops.momentum_apply_bwd(t(go), torch.tensor(0.7, device=DEV), sp, l2, s, True, True, delta)
    rgx, rgy = orc.enhanced_perturbation_bwd(go, 0.7, rsp, s, delta)
    np.testing.assert_allclose(gx.cpu().numpy(), rgx, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gy.cpu().numpy(), rgy, rtol=1e-5, atol=1e-6)


def test_enhanced_perturbation_matches_aten_sequence():
    """three calls with momentum, autograd to the adversary's parameters, against the reference's ATen expressions"""
    torch.manual_seed(0)
    adv = VoxelPerturber(hidden_channels=[8, 16, 32]).to(DEV)
    adv.train()
    ep = EnhancedPerturbation(adv)
    ep._epoch = 4
    last_ref = None
    hist = []
    for it in range(3):
        x = torch.randn(5000, 4, device=DEV) * 5
        ep._iteration = it + 1
        state = torch.get_rng_state()
        out, l2 = ep.apply_enhanced_perturbations(x)
        for q in adv.parameters():
            q.grad = None
        (out.square().sum() * 1e-3 + 2 * l2).backward()
        got = [q.grad.clone() for q in adv.parameters()]
        # the reference's lines with ATen ops
        torch.set_rng_state(state)
        s = orc.adversarial_strength(4, it + 1, hist)
        assert ep._current_scaling == pytest.approx(s)
        yy, _ = adv(x)
        del adv.l2_norms[-1]                                       # the second forward tracked metrics again
        spr = (yy - x) * s
        if last_ref is not None:
            spr = spr + 0.9 * last_ref
        last_ref = spr.detach()
        ref_out = x + spr
        ref_l2 = torch.norm(spr, p=2)
        hist.append(ref_l2.item())
        for q in adv.parameters():
            q.grad = None
        (ref_out.square().sum() * 1e-3 + 2 * ref_l2).backward()
        assert torch.equal(out, ref_out)
        assert float(l2.detach()) == pytest.approx(float(ref_l2.detach()), rel=1e-6)
        for a, b in zip(got, [q.grad for q in adv.parameters()]):
            np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=1e-3, atol=1e-6)
    assert ep._attack_history == pytest.approx(hist, rel=1e-6)
    # size change resets the momentum (:170-173); eval mode / no adversary pass the features through (:142-143)
    out, _ = ep.apply_enhanced_perturbations(torch.randn(77, 4, device=DEV))
    assert out.shape == (77, 4) and ep._last_perturbations.shape == (77, 4)
    x = torch.randn(10, 4, device=DEV)
    assert ep.apply_enhanced_perturbations(x, training=False) == (x, 0.0)


def test_strength_schedule_matches_oracle():
    ep = EnhancedPerturbation(None, max_scaling=5.0)
    for epoch, it, hist in [(0, 0, []), (3, 2500, [0.05] * 60), (10, 9000, [0.2] * 51), (80, 30000, [1.0] * 100),
                            (5, 100, [0.05] * 50)]:
        ep._epoch, ep._iteration, ep._attack_history = epoch, it, list(hist)
        assert ep.update_adversarial_strength() == orc.adversarial_strength(epoch, it, hist)
    assert EnhancedPerturbation(None, dynamic_scaling=False).update_adversarial_strength() == 1.0


@pytest.mark.parametrize("preset,N", [("kitti", 120000), ("nuscenes", 300001), ("kitti", 1), ("kitti", 1025)])
def test_points_range_filter_and_shuffle(preset, N):
    """pipeline transforms in front of the voxelizer (kitti-3d-car.py:62-64): bit-exact, order kept"""
    from robustpointclouds_b200 import PointShuffle, PointsRangeFilter, synth
    pcr = synth.PRESETS[preset]["point_cloud_range"]
    pts = synth.make_cloud(preset, N, "lidar", 3)
    pts[::97, 0] = pcr[0]                                            # exactly on the lower face: dropped (open box)
    pts[5::101, 1] = pcr[4]                                          # exactly on the upper face: dropped
    if N > 10:
        pts[7, 2] = np.nan
    ref = orc.points_range_filter(pts, pcr)
    t = torch.from_numpy(pts).to(DEV)
    d = PointsRangeFilter(pcr)({"points": t})
    assert np.array_equal(d["points"].cpu().numpy(), ref, equal_nan=True)
    out, index, count = ops.points_range_filter(t, pcr)
    k = int(count)
    assert k == len(ref) and np.array_equal(pts[index[:k].cpu().numpy()], ref, equal_nan=True)
    g = torch.Generator(DEV).manual_seed(5)
    sh = PointShuffle(g).shuffle(t)
    perm = torch.randperm(N, device=DEV, generator=torch.Generator(DEV).manual_seed(5))
    assert torch.equal(torch.nan_to_num(sh, nan=-1.0), torch.nan_to_num(t[perm], nan=-1.0))
    assert ops.points_range_filter(torch.zeros((0, 4), device=DEV), pcr)[2].item() == 0
