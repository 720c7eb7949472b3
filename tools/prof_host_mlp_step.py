"""cProfile of the host side of the reference-order KITTI step with the real perturber (where the wall time goes when
the GPU is idle half of the step)"""
import cProfile, pstats, sys, torch
import robustpointclouds_b200 as rpc
from robustpointclouds_b200 import synth
from robustpointclouds_b200.adversarial import extract_voxel_features
from robustpointclouds_b200.voxelization import voxelize
dev = torch.device("cuda:0"); preset = "kitti"; d = synth.PRESETS[preset]; B = 6; C = d["C"]
clouds = [torch.from_numpy(c).to(dev) for c in synth.make_batch(preset, B)]
layer = rpc.Voxelization(d["voxel_size"], d["point_cloud_range"], d["max_num_points"], d["max_voxels"]).train()
vfe = rpc.HardSimpleVFE(num_features=C)
adv = rpc.VoxelPerturber(hidden_channels=[64, 128, 64], voxel_size=d["voxel_size"], nan_guards="device").build(C, dev).train()
adv._return_loss_dict = True
vd = voxelize(clouds, layer); M = int(vd["voxels"].shape[0]); w = torch.randn((M, C), device=dev)
def step():
    vd = voxelize(clouds, layer)
    feats, ld = extract_voxel_features(vd, vfe, adv, training=True, epoch=3)
    loss = (feats * w).sum() + sum(ld.values())
    for q in adv.parameters(): q.grad = None
    loss.backward()
for _ in range(5): step()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(30): step()
torch.cuda.synchronize(); pr.disable()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(28)
