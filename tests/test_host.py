"""CPU tests of the host side: the C-ABI library loads and exports every symbol the header declares, the
sizing / error entry points behave (no compute calls without a GPU), the registry / perturber host logic
matches the reference's rules, and the multi-rank plumbing works over gloo with world_size 2."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import robustpointclouds_b200 as rpc
from oracle import ref_path as orc
from robustpointclouds_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "rpc_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(rpc_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name
    assert _lib.lib().rpc_abi_version() == 1
    assert _lib.strerror(0) == "ok" and "NULL" in _lib.strerror(-1)


def test_sizing_and_errors_without_gpu():
    L = _lib.lib()
    cfg = rpc.VoxelCfg.make([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 16000, 4)
    g = (ctypes.c_int32 * 3)()
    assert L.rpc_grid_size(ctypes.byref(cfg), g) == 0 and list(g) == [1408, 1600, 40]
    n = _lib.counts([120000] * 6)
    assert L.rpc_voxel_capacity(ctypes.byref(cfg), 6, n) == 96000
    assert L.rpc_workspace_bytes(ctypes.byref(cfg), 6, n) > 6 * 120000 * 12
    assert L.rpc_voxel_capacity(ctypes.byref(cfg), 3, _lib.counts([10, 0, 20000])) == 10 + 0 + 16000
    # invalid configurations are refused before anything is launched
    bad = rpc.VoxelCfg.make([0.05, 0.0, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 16000, 4)
    assert L.rpc_grid_size(ctypes.byref(bad), g) == -2
    assert L.rpc_workspace_bytes(ctypes.byref(bad), 6, n) == 0
    assert L.rpc_voxel_capacity(ctypes.byref(cfg), 33, n) == -1            # B > RPC_MAX_BATCH
    assert L.rpc_workspace_bytes(ctypes.byref(cfg), 1, _lib.counts([-5])) == 0
    bad_c = rpc.VoxelCfg.make([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 16000, 9)
    assert L.rpc_voxelize_hard(ctypes.byref(bad_c), None, n, 6, None, None, None, None, None, None, None, None,
                               None, None, 0, None) == -2
    assert L.rpc_voxelize_hard(ctypes.byref(cfg), None, n, 6, None, None, None, None, None, None, None, None,
                               None, None, 0, None) == -1                   # NULL pointers
    assert L.rpc_valid_rows_workspace_bytes(96000, 5) >= 4 * (96000 * 5 // 1024 + 1)


def test_ops_refuse_cpu_tensors():
    cfg = rpc.VoxelCfg.make([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 100, 4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        rpc.ops.hard_voxelize(torch.zeros(10, 4), [10], cfg)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        rpc.ops.vfe_mean_fwd(torch.zeros(3, 5, 4), torch.ones(3, dtype=torch.int32), 4)


@pytest.mark.parametrize("C,training", [(4, True), (4, False), (5, True), (5, False), (6, True)])
def test_perturber_bounds_match_reference_rules(C, training):
    eb, fb = rpc.perturber_bounds(C, training, 0.2)
    oeb, ofb = orc.perturber_bounds(C, training, 0.2)
    assert np.array_equal(eb.numpy(), oeb) and np.array_equal(fb.numpy(), ofb)
    if C == 5:
        assert eb.tolist() == pytest.approx([0.2, 0.2, 0.2, 0.2, 0.0]) and fb[4] == 0


def test_perturber_bounds_below_four_features_crashes_like_reference():
    with pytest.raises(UnboundLocalError):
        rpc.perturber_bounds(3, True)


def test_registry_and_lazy_build():
    adv = rpc.build_adversary(dict(type="VoxelPerturber", sensor_error_bound=0.2, voxel_size=[0.05, 0.05, 0.1],
                                   use_spatial_attention=True, hidden_channels=[64, 128, 64]))
    assert isinstance(adv, rpc.VoxelPerturber) and rpc.ADVERSARIES.get("VoxelPerturber") is rpc.VoxelPerturber
    assert adv.model is None and adv.attention is None and len(list(adv.parameters())) == 0   # lazy, like the reference
    adv.build(4)
    assert sum(p.numel() for p in adv.parameters()) == 34641
    adv5 = rpc.VoxelPerturber(voxel_size=[0.075, 0.075, 0.2], hidden_channels=[64, 128, 64]).build(5)
    assert sum(p.numel() for p in adv5.parameters()) == 34772 and adv5.in_features == 5
    with pytest.raises(KeyError):
        rpc.build_adversary(dict(type="NoSuchAdversary"))
    bn = [m for m in adv.modules() if isinstance(m, torch.nn.BatchNorm1d)]
    assert len(bn) == 5 and all(m.eps == 1e-3 and m.momentum == 0.1 for m in bn)


def test_state_dict_keys_match_reference_fixture(golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "perturber_kitti_train.npz")))
    state = {k[len("state_"):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("state_")}
    adv = rpc.VoxelPerturber(hidden_channels=[int(h) for h in g["hidden"]]).build(int(g["C"]))
    missing, unexpected = adv.load_state_dict(state, strict=True)
    assert not missing and not unexpected


def test_param_grad_hook_clamps_like_reference():
    torch.manual_seed(0)
    adv = rpc.VoxelPerturber(hidden_channels=[16, 32, 64]).build(4)
    adv.train()
    out = adv.model(torch.randn(64, 4) * 5)          # the MLP alone runs anywhere; the fused tail is CUDA-only
    (out.sum() * 1e4).backward()
    gmax = max(float(p.grad.abs().max()) for p in adv.parameters() if p.grad is not None)
    assert gmax == pytest.approx(0.1)


def test_metrics_csv_roundtrip(tmp_path):
    adv = rpc.VoxelPerturber()
    adv.l2_norms, adv.l2_percentages, adv.constraint_violations = [0.1, 0.2], [1.0, 2.0], [0.0, 0.5]
    adv.perturbation_stats = [dict(l2_norm=0.1), dict(l2_norm=0.2)]
    f = tmp_path / "l2.csv"
    adv.save_l2_norms(str(f))
    rows = f.read_text().strip().splitlines()
    assert rows[0] == "L2 Norm,L2 Percentage,Constraint Violations" and len(rows) == 3
    assert (tmp_path / "l2_detailed.csv").exists() and adv.l2_norms == []


def test_shard_batch():
    assert rpc.shard_batch(6, 0, 1) == [0, 1, 2, 3, 4, 5]
    assert [rpc.shard_batch(8, r, 8) for r in range(8)] == [[r] for r in range(8)]
    assert rpc.shard_batch(6, 1, 4) == [1, 5] and rpc.shard_batch(2, 3, 4) == []
    got = sorted(s for r in range(4) for s in rpc.shard_batch(6, r, 4))
    assert got == list(range(6))


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)                                   # same weights on every rank
    adv = rpc.VoxelPerturber(hidden_channels=[8, 16, 32]).build(4)
    torch.manual_seed(100 + rank)                          # rank-local shard of the batch
    x = torch.randn(32, 4)
    out = adv.model(x) * adv.attention(x)
    (out.square().sum() * 50).backward()                   # the +-0.1 hook fires here, before the reduce
    local = torch.cat([p.grad.reshape(-1) for p in adv.parameters()]).clone()
    sync = rpc.FlatGradSync(adv.parameters())
    sync.all_reduce()
    after = torch.cat([p.grad.reshape(-1) for p in adv.parameters()])
    gathered = [torch.zeros_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    q.put((rank, sync.numel(), float((after - torch.stack(gathered).mean(0)).abs().max()), float(local.abs().max())))
    dist.destroy_process_group()


def test_flat_grad_allreduce_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=60) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n_params = sum(p.numel() for p in rpc.VoxelPerturber(hidden_channels=[8, 16, 32]).build(4).parameters())
    for rank, numel, err, gmax in res:
        assert numel == n_params and err < 1e-7 and gmax <= 0.1 + 1e-7


def test_dynamic_and_allreduce_entries_without_gpu():
    """argument validation of the round-1 additions happens on the host, before any launch"""
    L = _lib.lib()
    cfg = rpc.VoxelCfg.make([0.075, 0.075, 0.2], [-54, -54, -5, 54, 54, 3], 1, 1, 5)
    n = _lib.counts([300000] * 4)
    ws = L.rpc_dynamic_workspace_bytes(ctypes.byref(cfg), 4, n)
    # per point: cell + slot words, a 32-byte fine group and an accumulator row at most
    assert 1200000 * (4 + 4 + 32) < ws < 1200000 * 200
    assert L.rpc_dynamic_stamps_offset(ctypes.byref(cfg), 4, n) > 0
    assert L.rpc_dynamic_workspace_bytes(ctypes.byref(cfg), 40, n) == 0           # B > RPC_MAX_BATCH
    args = [None] * 9
    assert L.rpc_dynamic_scatter_mean_fwd(ctypes.byref(cfg), None, n, 4, None, None, *args[:7], None, 0, None) == -1
    bad = rpc.VoxelCfg.make([0.075, -1.0, 0.2], [-54, -54, -5, 54, 54, 3], 1, 1, 5)
    assert L.rpc_dynamic_scatter_mean_fwd(ctypes.byref(bad), None, n, 4, None, None, *args[:7], None, 0, None) == -2
    assert L.rpc_allreduce_flag_bytes() >= 64 * 16 * 4
    vp = ctypes.c_void_p
    one = (vp * 1)(None)
    assert L.rpc_allreduce_oneshot(one, one, one, 0, 17, 8, None, None, None, None, None) == -2   # world > 16
    assert L.rpc_allreduce_oneshot(one, one, one, 1, 1, 8, None, None, None, None, None) == -2    # rank >= world
    assert L.rpc_allreduce_oneshot(one, one, one, 0, 1, 8, None, None, None, None, None) == -1    # NULL pointers
    assert L.rpc_allreduce_oneshot(one, one, one, 0, 1, 0, None, None, None, None, None) == 0     # nothing to do


def test_perturber_statistics_entries_without_gpu():
    L = _lib.lib()
    assert L.rpc_moments_bytes() == 8 * 18                   # S_c, Q_c for 8 channels + the two row sums, fp64
    assert L.rpc_normalize_fwd(None, 10, None, 0, None, None, None) == -2         # C out of range
    assert L.rpc_normalize_fwd(None, 10, None, 9, None, None, None) == -2
    assert L.rpc_normalize_fwd(None, 10, None, 4, None, None, None) == -1         # NULL pointers
    assert L.rpc_normalize_fwd(None, 0, None, 4, None, None, None) == 0           # nothing to do
    assert L.rpc_perturb_losses_fwd(None, 10, None, 1, None, None, None, None) == -2
    assert L.rpc_perturb_losses_fwd(None, 10, None, 5, None, None, None, None) == -1
    assert L.rpc_perturb_losses_bwd(None, None, None, 10, None, 5, None, None) == -1
    assert L.rpc_perturb_losses_bwd(None, None, None, 0, None, 5, None, None) == 0
    cfg = rpc.VoxelCfg.make([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 16000, 4)
    assert L.rpc_voxel_centers(ctypes.byref(cfg), None, 5, 10, None, None, None) == -2       # coors rows of 3 or 4
    assert L.rpc_voxel_centers(ctypes.byref(cfg), None, 4, 10, None, None, None) == -1
    assert L.rpc_voxel_centers(ctypes.byref(cfg), None, 4, 0, None, None, None) == 0
    assert L.rpc_momentum_apply_fwd(None, None, None, 0, 1.0, 0.9, 10, None, None, None, None, None) == -1
    assert L.rpc_momentum_apply_bwd(None, None, None, None, 0, 1.0, -1, None, None, None) == -2
    assert L.rpc_momentum_apply_bwd(None, None, None, None, 0, 1.0, 0, None, None, None) == 0


def test_strength_schedule_host_logic():
    """strong_adversarial_voxelnet.py:114-138: epoch ramp, weak-attack boost, curriculum factor, cap"""
    ep = rpc.EnhancedPerturbation(None)
    assert ep.update_adversarial_strength() == 1.0
    ep._epoch, ep._iteration = 10, 5000
    assert ep.update_adversarial_strength() == pytest.approx(2.0 * 1.5)
    ep._attack_history = [0.05] * 51
    assert ep.update_adversarial_strength() == 5.0                       # 2.0 * 2.0 * 1.5 capped at max_scaling
    assert ep.update_adversarial_strength() == orc.adversarial_strength(10, 5000, [0.05] * 51)


def test_allreduce_ll_entries_without_gpu():
    L = _lib.lib()
    assert L.rpc_allreduce_ll_bytes(8, 34641) == 2 * 8 * 34642 * 8      # two parities x world x padded n x 8-byte words
    assert L.rpc_allreduce_ll_bytes(17, 8) == 0 and L.rpc_allreduce_ll_bytes(0, 8) == 0
    vp = ctypes.c_void_p
    one = (vp * 1)(None)
    assert L.rpc_allreduce_ll(one, 0, 17, 8, None, None, None, None, None) == -2      # world > 16
    assert L.rpc_allreduce_ll(one, 2, 2, 8, None, None, None, None, None) == -2       # rank >= world
    assert L.rpc_allreduce_ll(one, 0, 1, 8, None, None, None, None, None) == -1       # NULL pointers
    assert L.rpc_allreduce_ll(one, 0, 1, 0, None, None, None, None, None) == 0        # nothing to do


def test_range_filter_entry_without_gpu():
    L = _lib.lib()
    rng = (ctypes.c_float * 6)(0, -40, -3, 70.4, 40, 1)
    assert L.rpc_points_range_filter(None, 10, 2, rng, None, None, None, None, 0, None) == -2      # needs x, y, z
    assert L.rpc_points_range_filter(None, 10, 4, rng, None, None, None, None, 0, None) == -1      # NULL pointers
    pts = np.array([[1, 0, 0, 1], [0, 0, 0, 1], [70.4, 0, 0, 1], [5, -40, 0, 1], [np.nan, 0, 0, 1], [3, 3, 0.5, 1]], np.float32)
    assert orc.points_range_filter(pts, [0, -40, -3, 70.4, 40, 1]).tolist() == [[1, 0, 0, 1], [3, 3, 0.5, 1]]


def test_attention_gate_entries_without_gpu():
    L = _lib.lib()
    assert L.rpc_attention_scratch_bytes() == 8 * (4 * (8 + 2) + 1)               # W1, b1, w2, b2 at C = 8, fp64
    assert L.rpc_attention_gate_fwd(None, None, None, None, None, None, 10, 0, None, None, None) == -2     # C out of range
    assert L.rpc_attention_gate_fwd(None, None, None, None, None, None, 10, 9, None, None, None) == -2
    assert L.rpc_attention_gate_fwd(None, None, None, None, None, None, 10, 4, None, None, None) == -1     # NULL pointers
    assert L.rpc_attention_gate_fwd(None, None, None, None, None, None, 0, 4, None, None, None) == 0       # nothing to do
    args = [None] * 7
    assert L.rpc_attention_gate_bwd(*args, 10, 0, *([None] * 8)) == -2
    assert L.rpc_attention_gate_bwd(*args, 10, 4, *([None] * 8)) == -1


def test_linear_wgrad_entry_without_gpu():
    L = _lib.lib()
    assert L.rpc_linear_wgrad(None, None, 10, 0, 4, None, None, None) == -2        # M out of range
    assert L.rpc_linear_wgrad(None, None, 10, 64, 129, None, None, None) == -2     # N out of range
    assert L.rpc_linear_wgrad(None, None, 10, 64, 4, None, None, None) == -1       # NULL pointers


def test_bn_relu_entries_without_gpu():
    L = _lib.lib()
    assert L.rpc_bn_scratch_bytes() == 8 * 2 * 128
    assert L.rpc_bn_relu_fwd(None, 10, 0, None, None, 1e-3, 0.1, None, None, None, None, None, None, None) == -2
    assert L.rpc_bn_relu_fwd(None, 10, 129, None, None, 1e-3, 0.1, None, None, None, None, None, None, None) == -2
    assert L.rpc_bn_relu_fwd(None, 10, 64, None, None, 1e-3, 0.1, None, None, None, None, None, None, None) == -1
    assert L.rpc_bn_relu_bwd(None, None, None, None, None, None, 10, 64, None, None, None, None, None) == -1
