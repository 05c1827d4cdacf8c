"""Evaluate one golden case on any implementation with the reference's public layout."""
import torch

from cases import LAYER_CASES, MODEL_CASES, randomize, _g  # noqa: F401


def _probe(shape, seed, dtype):
    return torch.randn(shape, generator=_g(seed), dtype=torch.float32).to(dtype)


def eval_layer(layer, x, direction, nograd=False):
    """Returns dict(z, log_det, gx, grads{name}) for loss = <z,w1> + <log_det,w2>."""
    if nograd:      # reference path that autograd rejects (in-place on a chunk view, coupling.py:140)
        with torch.no_grad():
            z, ld = getattr(layer, direction)(x.clone())
        ld = ld if torch.is_tensor(ld) else torch.tensor(float(ld))
        return {"z": z.cpu().clone(), "log_det": ld.cpu().clone(), "gx": None, "grads": {}}
    x = x.clone().requires_grad_(True)
    z, ld = getattr(layer, direction)(x)
    out = {"z": z.detach().cpu().clone()}
    loss = (z * _probe(z.shape, 101, z.dtype).to(z.device)).sum()
    if torch.is_tensor(ld):
        out["log_det"] = ld.detach().cpu().clone()
        if ld.requires_grad:
            loss = loss + (ld * _probe(ld.shape, 102, ld.dtype).to(ld.device)).sum()
    else:
        out["log_det"] = torch.tensor(float(ld))
    names = [n for n, p in layer.named_parameters() if p.requires_grad]
    params = [p for _, p in layer.named_parameters() if p.requires_grad]
    grads = torch.autograd.grad(loss, [x] + params, allow_unused=True)
    out["gx"] = grads[0].detach().cpu().clone()
    out["grads"] = {n: (g.detach().cpu().clone() if g is not None else None) for n, g in zip(names, grads[1:])}
    return out


def eval_model(model, x, call, y=None):
    """forward_kld: loss + all param grads; log_prob: values + grads of its mean."""
    torch.manual_seed(1234)            # vdeq noise
    x = x.clone()
    args = (x,) if y is None else (x, y)
    val = getattr(model, call)(*args)
    loss = val if val.dim() == 0 else val.mean()
    names = [n for n, p in model.named_parameters() if p.requires_grad]
    params = [p for _, p in model.named_parameters() if p.requires_grad]
    grads = torch.autograd.grad(loss, params, allow_unused=True)
    return {"value": val.detach().cpu().clone(),
            "grads": {n: (g.detach().cpu().clone() if g is not None else None) for n, g in zip(names, grads)}}
