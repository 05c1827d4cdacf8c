import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests", "golden"), os.path.join(ROOT, "tests")]
import torch
import vcnf_b200 as nf
from cases import LAYER_CASES_R2
from harness import eval_layer
name, direction = "prq_coupling_img_uncond", "inverse"
g = torch.load(os.path.join(ROOT, "tests/golden/layers_r2.pt"))[name]
torch.manual_seed(0)
layer = LAYER_CASES_R2[name]["build"](nf); layer.load_state_dict(g["state"]); layer = layer.cuda()
got = eval_layer(layer, g["x"].cuda(), direction)
ref, r64 = g[direction], g[direction + "_f64"]
print("z err", float((got["z"] - ref["z"]).abs().max()), "ref32-64", float((ref["z"].double() - r64["z"]).abs().max()))
print("ld err", float((got["log_det"] - ref["log_det"]).abs().max()), "ref32-64", float((ref["log_det"].double() - r64["log_det"]).abs().max()))
for k in ["x"] + list(ref["grads"]):
    a = got["gx"] if k == "x" else got["grads"][k]
    b = ref["gx"] if k == "x" else ref["grads"][k]
    c = r64["gx"] if k == "x" else r64["grads"][k]
    sc = float(b.abs().max())
    print("%-50s scale %.3g  got-ref32 %.3g  got-ref64 %.3g  ref32-ref64 %.3g" % (k, sc, float((a - b).abs().max()) / sc, float((a.double() - c).abs().max()) / sc, float((b.double() - c).abs().max()) / sc))
# per-sample contribution: which sample drives the error? evaluate per-row gradient of x
e = (got["gx"] - ref["gx"]).abs().amax(dim=(1, 2, 3))
print("gx err per sample", [round(float(v), 5) for v in e])
print("z range", float(ref["z"].abs().max()), "x", float(g["x"].abs().max()))
