"""Deterministic, construction-order-independent parameter initialisation shared by the
reference run (gen_golden.py), the oracle and the CUDA modules.  TEST INFRASTRUCTURE ONLY."""
import math

import torch


def reseed_parameters(model, seed=1234):
    """Overwrite every parameter from one seeded generator in sorted-name order.
    conv/linear weights ~ N(0, sqrt(2/fan_out)); 1-D 'weight' (RangeBN gamma) ~ U(0.5,1.5);
    every bias ~ N(0, 0.05).  Unlike the model's own init this leaves no all-zero
    gamma/bias, so every quantizer sees a non-degenerate range."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in sorted(model.named_parameters()):
            if p.dim() > 1:
                fan = p.size(0) * (p[0, 0].numel() if p.dim() > 2 else 1)
                v = torch.randn(p.shape, generator=g) * math.sqrt(2.0 / fan)
            elif name.endswith("weight"):
                v = torch.rand(p.shape, generator=g) + 0.5
            else:
                v = torch.randn(p.shape, generator=g) * 0.05
            p.copy_(v.to(p.device))
    return model


def synthetic_batch(shape, num_classes, seed=123):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(shape, generator=g), torch.randint(0, num_classes, (shape[0],), generator=g)
