"""Binary feature masks (reference: normflow/utils/masks.py:5-53)."""
import torch


def create_alternating_binary_mask(features, even=True):
    mask = torch.zeros(features).byte()
    start = 0 if even else 1
    mask[start::2] += 1
    return mask


def create_mid_split_binary_mask(features):
    mask = torch.zeros(features).byte()
    midpoint = features // 2 if features % 2 == 0 else features // 2 + 1
    mask[:midpoint] += 1
    return mask


def create_random_binary_mask(features, seed=None):
    mask = torch.zeros(features).byte()
    weights = torch.ones(features).float()
    num_samples = features // 2 if features % 2 == 0 else features // 2 + 1
    if seed is None:
        indices = torch.multinomial(input=weights, num_samples=num_samples, replacement=False)
    else:
        gen = torch.Generator()
        gen.manual_seed(seed)
        indices = torch.multinomial(input=weights, num_samples=num_samples, replacement=False, generator=gen)
    mask[indices] += 1
    return mask
