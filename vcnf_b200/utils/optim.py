"""Optimisation helpers (reference: normflow/utils/optim.py:10-22)."""


def set_requires_grad(module, flag):
    for param in module.parameters():
        param.requires_grad = flag


def clear_grad(model):
    for param in model.parameters():
        param.grad = None
