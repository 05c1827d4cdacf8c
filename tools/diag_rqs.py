"""GPU-box diagnostic (not a test): spline layer product vs CPU oracle, worst elements listed."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")):
    sys.path.insert(0, p)
import torch
import oracle
import vcnf_b200 as nf
from cases import LAYER_CASES, randomize
from harness import eval_layer


def worst(name, a, b, k=8, thr=1e-3):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    d = (a - b).abs()
    d = torch.where(torch.isnan(d), torch.full_like(d, float("inf")), d)
    print("%s: max %.3g  n(>%.0e)=%d  nan_a=%d nan_b=%d of %d" % (name, float(d.max()), thr, int((d > thr).sum()),
          int(torch.isnan(a).sum()), int(torch.isnan(b).sum()), d.numel()))
    flat = d.flatten().topk(min(k, d.numel()))
    for v, i in zip(flat.values, flat.indices):
        if v <= thr:
            break
        idx = list(torch.unravel_index(i, d.shape))
        idx = [int(j) for j in idx]
        print("   ", idx, "got %.6g ref %.6g" % (float(a[tuple(idx)]), float(b[tuple(idx)])))


def pair(build, scale, seed=0):
    torch.manual_seed(seed)
    ref = build(oracle)
    randomize(ref, scale=scale)
    torch.manual_seed(seed)
    prod = build(nf)
    prod.load_state_dict(ref.state_dict())
    return ref, prod.cuda()


def round_trip(B, scale=0.05):
    print("== round trip B=%d" % B)
    ref, prod = pair(lambda m: m.flows.CoupledRationalQuadraticSpline(29, 2, 128), scale)
    x = 1.5 * torch.randn(B, 29, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        y, ld = prod.inverse(x.cuda())
        yo, ldo = ref.inverse(x)
        worst("density y", y, yo)
        worst("density ld", ld, ldo)
        x1, ld1 = prod.forward(y)
        x1o, ld1o = ref.forward(y.cpu())
        worst("sampling x (same y)", x1, x1o)
        worst("sampling ld", ld1, ld1o)
        worst("round trip oracle x1o vs x", x1o, x)
        worst("round trip product x1 vs x", x1, x)


def golden_grad():
    print("== golden crqs_linear forward")
    g = torch.load(os.path.join(ROOT, "tests", "golden", "layers.pt"))["crqs_linear"]
    torch.manual_seed(0)
    layer = LAYER_CASES["crqs_linear"]["build"](nf)
    layer.load_state_dict(g["state"])
    layer = layer.cuda()
    got = eval_layer(layer, g["x"].cuda(), "forward")
    ref = g["forward"]
    worst("z", got["z"], ref["z"], thr=1e-4)
    worst("z vs f64", got["z"], g["forward_f64"]["z"], thr=1e-4)
    worst("ref z vs f64", ref["z"], g["forward_f64"]["z"], thr=1e-4)
    worst("gx", got["gx"], ref["gx"], thr=1e-1)
    # the same through the oracle in fp64
    torch.manual_seed(0)
    o = LAYER_CASES["crqs_linear"]["build"](oracle)
    o.load_state_dict(g["state"])
    r64 = eval_layer(o.double(), g["x"].double(), "forward")
    worst("gx vs oracle f64", got["gx"], r64["gx"], thr=1e-1)
    worst("ref gx vs oracle f64", ref["gx"], r64["gx"], thr=1e-1)


if __name__ == "__main__":
    golden_grad()
    for B in (64, 4096, 65536):
        round_trip(B)
