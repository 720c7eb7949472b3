"""diagnostic: full objective through our VoxelPerturber vs an fp64 ATen chain of deep-copied modules, per module"""
import copy, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import robustpointclouds_b200 as rpc
from robustpointclouds_b200.voxel_perturber import perturber_bounds
name = sys.argv[1] if len(sys.argv) > 1 else "kitti_eval"
g = dict(np.load(f"tests/golden/perturber_{name}.npz", allow_pickle=True))
state = {k[len("state_"):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("state_")}
DEV = torch.device("cuda:0"); C = int(g["C"]); hidden = [int(h) for h in g["hidden"]]; training = bool(g["training"])
adv = rpc.VoxelPerturber(hidden_channels=hidden, voxel_size=[0.05, 0.05, 0.1] if C == 4 else [0.075, 0.075, 0.2])
adv.build(C); adv.load_state_dict(state); adv = adv.to(DEV); adv.train(training)
x = torch.from_numpy(g["x"]).to(DEV); w = torch.from_numpy(g["w"]).to(DEV)
adv._prepare(x)
adv64 = copy.deepcopy(adv).double()
for q in adv64.parameters(): q._backward_hooks = None
for q in adv.parameters(): q._backward_hooks = None
acts, grads = {}, {}
def fh(tag):
    def f(m, i, o):
        acts[tag] = o.detach().double().cpu().clone()
        if o.requires_grad: o.register_hook(lambda gr: grads.__setitem__(tag, gr.detach().double().cpu().clone()))
    return f
hs = [layer.register_forward_hook(fh(("m", i))) for i, layer in enumerate(adv.model)]
y, ld = adv(x)
((y * w).sum() + ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"] + 4 * ld["imbalance_loss"]).backward()
p32 = {k: q.grad.detach().double().cpu() for k, q in adv.named_parameters()}
# fp64 chain
acts64, grads64 = {}, {}
xd = x.double()
xn = torch.clamp(xd / (xd.std(0, keepdim=True) + 1e-6), -10, 10)
h = xn
for i, layer in enumerate(adv64.model):
    h = layer(h)
    acts64[("m", i)] = h.detach().cpu().clone()
    h.register_hook(lambda gr, i=i: grads64.__setitem__(("m", i), gr.detach().cpu().clone()))
raw = h * adv64.attention(xn)
eb, fb = perturber_bounds(C, training); eb = eb.to(DEV).double(); fb = fb.to(DEV).double()
p = torch.nan_to_num(torch.clamp(raw * eb, -fb, fb)); y64 = xd + p
l2 = torch.norm(p, p=2, dim=1).mean(); inten = p[:, 3].abs().mean(); bias = p.mean(0).abs().mean(); imb = p.std(0).std()
((y64 * w.double()).sum() + l2 + 2 * inten + 3 * bias + 4 * imb).backward()
p64 = {k: q.grad.detach().cpu() for k, q in adv64.named_parameters()}
rel = lambda a, b: float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))
print("y", rel(y.detach().double().cpu(), y64.detach().cpu()))
for i in range(len(adv.model)):
    t = ("m", i)
    print(i, type(adv.model[i]).__name__, "act %.1e" % rel(acts[t], acts64[t]), "grad_out %.1e" % (rel(grads[t], grads64[t]) if t in grads else -1))
for k in p32:
    print(k, "pgrad rel-to-scale %.1e" % rel(p32[k], p64[k]), "scale %.2g" % float(p64[k].abs().max()),
          "vs fixture pgrad64 %.1e" % rel(p64[k], torch.from_numpy(np.asarray(g["pgrad64_" + k], np.float64))))
