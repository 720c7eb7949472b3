"""Generate golden vectors from the REFERENCE's own code (run in the build container only).

    python tests/golden/make_golden.py

Imports /root/reference/models/adversarial/voxel_perturber.py (the real ``VoxelPerturber``) through
a minimal stand-in for ``mmengine.registry.Registry`` (mmengine is not installed; the class body
only needs ``register_module`` / ``build``), runs it on seeded inputs and stores inputs, weights and
outputs as small ``.npz`` fixtures next to this script.  /root/reference does not exist on the GPU
box, so tests read only the committed fixtures.

The voxelization fixtures (``voxel_micro.npz``) are HAND-DERIVED from the algorithm statement in
SURVEY 8a-U because nothing upstream pins them here (mmcv absent) -- see ``hand_cases`` below;
each expected array was written out by hand, not produced by our oracle.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def _load_reference_perturber():
    class Registry:  # 15-line shim of mmengine.registry.Registry
        def __init__(self, name, parent=None, scope=None):
            self.name, self.module_dict = name, {}

        def register_module(self, name=None, force=False, module=None):
            def deco(cls):
                self.module_dict[name or cls.__name__] = cls
                return cls
            return deco(module) if module is not None else deco

        def build(self, cfg):
            cfg = dict(cfg)
            return self.module_dict[cfg.pop("type")](**cfg)

    mmengine = types.ModuleType("mmengine")
    registry = types.ModuleType("mmengine.registry")
    registry.Registry = Registry
    mmengine.registry = registry
    sys.modules.setdefault("mmengine", mmengine)
    sys.modules.setdefault("mmengine.registry", registry)

    def load(name, path):
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod

    models = types.ModuleType("models")
    models.__path__ = []
    sys.modules["models"] = models
    builder = load("models.builder", f"{REF}/models/builder.py")
    vp = load("models.adversarial.voxel_perturber", f"{REF}/models/adversarial/voxel_perturber.py")
    return builder, vp


def synth_points(K, C, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.empty(K, C)
    x[:, 0] = torch.rand(K, generator=g) * 70.4
    x[:, 1] = torch.rand(K, generator=g) * 80 - 40
    x[:, 2] = torch.rand(K, generator=g) * 4 - 3
    x[:, 3] = torch.rand(K, generator=g)
    if C > 4:
        x[:, 4] = torch.randint(0, 10, (K,), generator=g).float() * 0.05
    return x


def perturber_case(builder, name, C, training, hidden, K, seed):
    torch.manual_seed(seed)
    adv = builder.build_adversary(dict(type="VoxelPerturber", sensor_error_bound=0.2,
                                       voxel_size=[0.05, 0.05, 0.1] if C == 4 else [0.075, 0.075, 0.2],
                                       use_spatial_attention=True, hidden_channels=hidden))
    adv.train()
    warm = synth_points(K, C, seed + 1)
    adv(warm)  # lazy build + one BN running-stat update
    adv.train(training)
    state = {k: v.detach().clone().numpy() for k, v in adv.state_dict().items()}

    x = synth_points(K, C, seed + 2)
    # fp64 shadow of the same module on the same input: arbitrates the tolerance of the parameter-gradient pin
    # (rtol 1e-2 against the fp32 run was two orders looser than the stated bar)
    import copy
    adv64 = copy.deepcopy(adv).double()
    adv64.train(training)
    captured = {}
    orig = adv._apply_physical_constraints

    def spy(p, vf):
        out = orig(p, vf)
        captured["p"] = out
        return out

    adv._apply_physical_constraints = spy
    h1 = adv.model.register_forward_hook(lambda m, i, o: captured.__setitem__("mlp_out", o.detach().clone()))
    h2 = adv.attention.register_forward_hook(lambda m, i, o: captured.__setitem__("attn", o.detach().clone()))
    y, ld = adv(x)
    h1.remove(); h2.remove()
    p = captured["p"]
    w = synth_points(K, C, seed + 3) * 0.01 - 0.2  # upstream gradient on the perturbed points
    p.retain_grad()
    total = (y * w).sum() + ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"] + 4 * ld["imbalance_loss"]
    for q in adv.parameters():
        q.grad = None
    total.backward()
    out = dict(x=x.numpy(), y=y.detach().numpy(), p=p.detach().numpy(), w=w.numpy(),
               mlp_out=captured["mlp_out"].numpy(), attn=captured["attn"].numpy(),
               grad_p=p.grad.numpy(), training=np.array(training), C=np.array(C), hidden=np.array(hidden),
               **{f"loss_{k}": v.detach().numpy() for k, v in ld.items()})
    for k, v in state.items():
        out[f"state_{k}"] = v
    for k, q in adv.named_parameters():
        out[f"pgrad_{k}"] = q.grad.numpy()  # after the +-0.1 clamp hook (voxel_perturber.py:465-475)
    # the fp64 run: same weights, same input, same loss; gradients after the same +-0.1 clamp hooks
    cap64 = {}
    orig64 = adv64._apply_physical_constraints

    def spy64(p_, vf):
        o = orig64(p_, vf)
        cap64["p"] = o
        return o

    adv64._apply_physical_constraints = spy64
    # smallest |ReLU input| of the fp64 run: a pre-activation within fp32 rounding of zero lets two correct fp32
    # implementations (CPU / GPU operation order) take different sides of the kink, which moves the parameter gradients
    # by O(1/K).  The first kitti_eval fixture (seed 200) had one at 6e-9 and failed on the B200 at 1e-4.
    margins = []
    hooks = [m.register_forward_hook(lambda m_, i_, o_: margins.append(float(o_.detach().abs().min())))
             for m in adv64.modules() if isinstance(m, torch.nn.BatchNorm1d)]
    hooks += [adv64.attention[0].register_forward_hook(lambda m_, i_, o_: margins.append(float(o_.detach().abs().min())))]
    y64, ld64 = adv64(x.double())
    for h_ in hooks:
        h_.remove()
    out["relu_margin"] = np.array(min(margins))
    print(name, "relu margin %.3g" % min(margins))
    assert min(margins) > 1e-5, "pick another seed: a ReLU input within fp32 rounding of zero"
    p64 = cap64["p"]
    p64.retain_grad()
    total64 = ((y64 * w.double()).sum() + ld64["l2_norm"] + 2 * ld64["intensity_loss"] + 3 * ld64["bias_loss"]
               + 4 * ld64["imbalance_loss"])
    for q in adv64.parameters():
        q.grad = None
    total64.backward()
    out["y64"] = y64.detach().numpy()
    out["grad_p64"] = p64.grad.numpy()
    for k, v in ld64.items():
        out[f"loss64_{k}"] = v.detach().numpy()
    for k, q in adv64.named_parameters():
        out[f"pgrad64_{k}"] = q.grad.numpy()
    np.savez_compressed(os.path.join(HERE, f"perturber_{name}.npz"), **out)
    print(name, "K", K, "C", C, "l2", float(ld["l2_norm"]), "max|pgrad|",
          max(float(np.abs(q.grad.numpy()).max()) for q in adv.parameters()))


def extract_feat_case(builder, name, C, seed):
    """adversarial_voxelnet.py:81-117 + HardSimpleVFE on a tiny padded voxel tensor (the lines are
    restated here because the detector class body needs mmdet3d; the adversary is the real one)."""
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed)
    M, P = 37, 5
    num_points = torch.randint(1, P + 1, (M,), generator=g).int()
    voxels = torch.zeros(M, P, C)
    for m in range(M):
        voxels[m, : num_points[m]] = synth_points(int(num_points[m]), C, seed + 10 + m)
    voxels[3, 0] = 0.0                                     # a real (0,0,0,0) null return
    if C == 4:
        voxels[5, 1] = torch.tensor([1.5, -2.5, 0.75, 0.25])   # real point whose features sum to 0
    adv = builder.build_adversary(dict(type="VoxelPerturber", hidden_channels=[16, 32, 64]))
    adv.train()
    adv(synth_points(64, C, seed + 1))
    state = {k: v.detach().clone().numpy() for k, v in adv.state_dict().items()}
    # ---- restated lines 85-117 ----
    num_voxels, max_points, num_features = voxels.shape
    voxels_2d = voxels.view(-1, num_features)
    valid_mask = voxels_2d.sum(dim=1) != 0
    valid_voxels = voxels_2d[valid_mask]
    perturbed_valid, loss_dict = adv(valid_voxels)
    voxels_2d_perturbed = voxels_2d + torch.zeros_like(voxels_2d)
    voxels_2d_perturbed[valid_mask] = perturbed_valid
    perturbed_voxels = voxels_2d_perturbed.view(num_voxels, max_points, num_features)
    # ---- HardSimpleVFE ----
    feats = perturbed_voxels[:, :, :C].sum(dim=1) / num_points.type_as(perturbed_voxels).view(-1, 1)
    gfe = torch.randn(M, C, generator=g)
    perturbed_valid.retain_grad()
    (feats * gfe).sum().backward()
    out = dict(voxels=voxels.numpy(), num_points=num_points.numpy(), valid_mask=valid_mask.numpy(),
               perturbed_voxels=perturbed_voxels.detach().numpy(), feats=feats.detach().numpy(),
               grad_feats=gfe.numpy(), grad_perturbed_valid=perturbed_valid.grad.numpy(),
               **{f"loss_{k}": v.detach().numpy() for k, v in loss_dict.items()})
    for k, v in state.items():
        out[f"state_{k}"] = v
    np.savez_compressed(os.path.join(HERE, f"extract_feat_{name}.npz"), **out)
    print(name, "valid rows", int(valid_mask.sum()), "by count", int(num_points.sum()))


def hand_cases():
    """Hand-derived voxelization vectors (SURVEY 8a-U walked by hand; see comments)."""
    vs = np.array([1.0, 1.0, 1.0], np.float32)
    pcr = np.array([0, 0, 0, 4, 3, 2], np.float32)          # grid gx=4, gy=3, gz=2
    pts = np.array([
        [0.5, 0.5, 0.5, 10],    # 0 -> cell (z0,y0,x0)  voxel 0 slot 0
        [3.5, 2.5, 1.5, 11],    # 1 -> (1,2,3)          voxel 1 slot 0
        [0.6, 0.4, 0.1, 12],    # 2 -> (0,0,0)          voxel 0 slot 1
        [4.0, 1.0, 1.0, 13],    # 3 x == hi -> c=4 >= 4 dropped
        [-0.1, 1.0, 1.0, 14],   # 4 x < lo  -> dropped
        [0.0, 0.0, 0.0, 15],    # 5 exactly lo -> (0,0,0) voxel 0 slot 2 -> P=2 overflow, not stored
        [2.2, 1.1, 0.9, 16],    # 6 -> (0,1,2)          voxel 2 slot 0
        [1.0, 2.999, 1.0, 17],  # 7 -> (1,2,1)          4th distinct cell: refused when V=3
        [3.9, 2.1, 1.9, 18],    # 8 -> (1,2,3)          voxel 1 slot 1 (existing voxel after cap)
        [np.nan, 1.0, 1.0, 19], # 9 NaN -> dropped
        [1.2, 2.5, 1.5, 20],    # 10 -> (1,2,1)         refused again (cell never admitted)
    ], np.float32)
    P, V = 2, 3
    coors = np.array([[0, 0, 0], [1, 2, 3], [0, 1, 2]], np.int32)
    num = np.array([2, 2, 1], np.int32)
    vox = np.zeros((3, 2, 4), np.float32)
    vox[0, 0], vox[0, 1] = pts[0], pts[2]
    vox[1, 0], vox[1, 1] = pts[1], pts[8]
    vox[2, 0] = pts[6]
    dyn = np.array([[0, 0, 0], [1, 2, 3], [0, 0, 0], [-1, -1, -1], [-1, -1, -1], [0, 0, 0], [0, 1, 2],
                    [1, 2, 1], [1, 2, 3], [-1, -1, -1], [1, 2, 1]], np.int32)
    p2v = np.array([0, 1, 0, -1, -1, -1, 2, -1, 1, -1, -1], np.int32)
    p2s = np.array([0, 0, 1, -1, -1, -1, 0, -1, 1, -1, -1], np.int32)
    # same cloud, V large, P large: 4 voxels in first-occurrence order
    coors_big = np.array([[0, 0, 0], [1, 2, 3], [0, 1, 2], [1, 2, 1]], np.int32)
    num_big = np.array([3, 2, 1, 2], np.int32)
    # DynamicScatter mean: voxels sorted lexicographically by (z,y,x)
    ds_coors = np.array([[0, 0, 0], [0, 1, 2], [1, 2, 1], [1, 2, 3]], np.int32)
    ds_count = np.array([3, 1, 2, 2], np.int32)
    ds_map = np.array([0, 3, 0, -1, -1, 0, 1, 2, 3, -1, 2], np.int32)
    ds_feats = np.stack([
        (pts[0] + pts[2] + pts[5]) / np.float32(3), pts[6], (pts[7] + pts[10]) / np.float32(2),
        (pts[1] + pts[8]) / np.float32(2)]).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "voxel_micro.npz"), vs=vs, pcr=pcr, pts=pts, P=P, V=V,
                        coors=coors, num=num, vox=vox, dyn=dyn, p2v=p2v, p2s=p2s, coors_big=coors_big,
                        num_big=num_big, ds_coors=ds_coors, ds_count=ds_count, ds_map=ds_map,
                        ds_feats=ds_feats, grid=np.array([4, 3, 2], np.int32))
    print("voxel_micro written")


if __name__ == "__main__":
    builder, _ = _load_reference_perturber()
    perturber_case(builder, "kitti_train", 4, True, [64, 128, 64], 311, 100)
    perturber_case(builder, "kitti_eval", 4, False, [16, 32, 64], 127, 203)  # seed chosen for the ReLU margin
    perturber_case(builder, "nus_train", 5, True, [16, 32, 64], 293, 300)
    extract_feat_case(builder, "kitti", 4, 400)
    extract_feat_case(builder, "nus", 5, 500)
    hand_cases()
