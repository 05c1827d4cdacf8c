"""Generate tests/golden/*.pt from the UNMODIFIED reference (run in the build container only).

    python tests/golden/make_golden.py

For every case in cases.py: build the reference layer / model under torch.manual_seed(0),
perturb its parameters (cases.randomize), save the state_dict taken BEFORE the first call
(so data-dependent ActNorm init is part of what is checked) and the outputs / gradients of
each call.  fp32 cases also store an fp64 evaluation (`*_f64`) for the tail clause of the
tolerance.
"""
import copy
import os
import sys
import warnings

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_shim import load_reference          # noqa: E402
from cases import LAYER_CASES, LAYER_CASES_R2, LAYER_CASES_R3, MODEL_CASES, randomize   # noqa: E402
from harness import eval_layer, eval_model   # noqa: E402

warnings.filterwarnings("ignore")


def layer_records(nf, cases, full64=False):
    layers = {}
    for name, case in cases.items():
        torch.manual_seed(0)
        dtype = case.get("dtype", torch.float32)
        layer = case["build"](nf)
        if case.get("randomize", True):
            randomize(layer, scale=case.get("scale", 0.3))
        layer = layer.to(dtype)
        x = case["x"]().to(dtype)
        rec = {"state": copy.deepcopy(layer.state_dict()), "x": x}
        for direction in ("inverse", "forward"):
            fresh = copy.deepcopy(layer)
            rec[direction] = eval_layer(fresh, x, direction, case.get("nograd", False))
            if dtype == torch.float32:
                f64 = copy.deepcopy(layer).double()
                r = eval_layer(f64, x.double(), direction, case.get("nograd", False))
                rec[direction + "_f64"] = r if full64 else {"z": r["z"], "log_det": r["log_det"]}
        layers[name] = rec
        print("layer", name, tuple(x.shape), "ok")
    return layers


def main():
    nf = load_reference()
    if "--r3" in sys.argv:                 # bounded-spline cases only: the earlier fixtures stay byte-identical
        torch.save(layer_records(nf, LAYER_CASES_R3, full64=True), os.path.join(HERE, "layers_r3.pt"))
        print("layers_r3.pt", os.path.getsize(os.path.join(HERE, "layers_r3.pt")) // 1024, "KiB")
        return
    if "--r2" in sys.argv:                 # round-2 rows only: the round-1 fixtures stay byte-identical
        torch.save(layer_records(nf, LAYER_CASES_R2, full64=True), os.path.join(HERE, "layers_r2.pt"))
        print("layers_r2.pt", os.path.getsize(os.path.join(HERE, "layers_r2.pt")) // 1024, "KiB")
        return
    torch.save(layer_records(nf, LAYER_CASES), os.path.join(HERE, "layers.pt"))

    models = {}
    for name, case in MODEL_CASES.items():
        torch.manual_seed(0)
        dtype = case.get("dtype", torch.float32)
        model = case["build"](nf)
        randomize(model, scale=case.get("scale", 0.3))
        model = model.to(dtype)
        x = case["x"]()
        y = case["y"]() if "y" in case else None
        rec = {"state": copy.deepcopy(model.state_dict()), "x": x, "y": y}
        for call in case["calls"]:
            rec[call] = eval_model(copy.deepcopy(model), x, call, y)
        models[name] = rec
        print("model", name, {c: float(rec[c]["value"].float().mean()) for c in case["calls"]})
    torch.save(models, os.path.join(HERE, "models.pt"))
    for f in ("layers.pt", "models.pt"):
        print(f, os.path.getsize(os.path.join(HERE, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
