"""Golden vectors for the calibration / deployment helpers, generated from the UNMODIFIED reference
(models/modules/quantize.py:301-312, 386-396, 547-592): set_measure_mode -> calibrate -> quantized eval, set_quant_mode,
overwrite_params, freeze_quant_params, QuantMeasure with momentum=0.  TEST INFRASTRUCTURE ONLY.

    python -m oracle.gen_golden_n2      ->  tests/golden/n2_helpers.npz
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import ref_import  # noqa: E402


def main():
    _, rq = ref_import.load()
    from n2_net import N2Net
    torch.manual_seed(2024)
    net = N2Net(rq)
    with torch.no_grad():
        for p in net.parameters():
            p.copy_(torch.randn_like(p) * 0.1)
        net.bn1.weight.uniform_(0.5, 1.5)
    out = {}
    for k, v in net.state_dict().items():
        out["init/" + k] = v.detach().numpy().copy()
    gen = torch.Generator().manual_seed(7)
    xs = [torch.randn(8, 16, 8, 8, generator=gen) * (1 + i) + 0.3 * i for i in range(3)]
    xe = torch.randn(8, 16, 8, 8, generator=gen)
    out["xs"] = torch.stack(xs).numpy()
    out["xe"] = xe.numpy()
    # 1. calibration: measure mode, train-mode forwards (quantizers pass through, ranges accumulate)
    net.train()
    rq.set_measure_mode(net, True)
    with torch.no_grad():
        for i, x in enumerate(xs):
            out["measure_out%d" % i] = net(x).numpy().copy()
    for k, v in net.state_dict().items():
        out["calibrated/" + k] = v.detach().numpy().copy()
    # 2. quantized evaluation with the calibrated ranges
    rq.set_measure_mode(net, False)
    net.eval()
    with torch.no_grad():
        out["eval_out"] = net(xe).numpy().copy()
        # 3. set_quant_mode(False): the float network
        rq.set_quant_mode(net, False)
        out["float_out"] = net(xe).numpy().copy()
        rq.set_quant_mode(net, True)
        out["eval_out_again"] = net(xe).numpy().copy()
    # 4. freeze_quant_params: QuantMeasure modules leave training mode -> running ranges frozen in a train-mode forward
    net.train()
    rq.freeze_quant_params(net)
    with torch.no_grad():
        out["frozen_train_out"] = net(xs[0]).numpy().copy()
    for k, v in net.state_dict().items():
        out["frozen/" + k] = v.detach().numpy().copy()
    # 5. overwrite_params: quantized weights baked into the parameters (with the stored ranges)
    rq.overwrite_params(net)
    for k, v in net.state_dict().items():
        out["overwritten/" + k] = v.detach().numpy().copy()
    # 6. QuantMeasure(momentum=0) is the cumulative average (quantize.py:217: `momentum or ... n/(n+1)`)
    qm = rq.QuantMeasure(8, momentum=0).train()
    with torch.no_grad():
        for x in xs:
            out["mom0_out"] = qm(x).numpy().copy()
    for k, v in qm.state_dict().items():
        out["mom0/" + k] = v.numpy().copy()
    path = os.path.join(ROOT, "tests", "golden", "n2_helpers.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, len(out), "arrays")


if __name__ == "__main__":
    main()
