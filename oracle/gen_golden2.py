"""Golden vectors for the quantize2.py flavour (row A9), produced by the UNMODIFIED reference file
/root/reference/models/modules/quantize2.py (imported by path; build container only):

    python -m oracle.gen_golden2

Writes tests/golden/q2_*.npz with inputs, outputs and the uniform(-0.5, 0.5) noise the reference drew (call order)."""
import importlib.util
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_import  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def load_ref2():
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location("ref_quantize2", os.path.join(ref_import.REF_ROOT, "models", "modules",
                                                                                "quantize2.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def save(name, d):
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name), **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v))
                                                    for k, v in d.items()})
    print("wrote", name, len(d), "arrays")


def gen_fq(r2):
    g = torch.Generator().manual_seed(20)
    d, i = {}, 0
    for shape in [(4, 3, 5, 5), (8, 16), (2, 1000)]:
        x = torch.randn(shape, generator=g) * 1.7 + 0.4
        for bits in (4, 8):
            for signed in (False, True):
                for deq in (True, False):
                    d["x%d" % i] = x
                    d["meta%d" % i] = np.array([bits, int(signed), int(deq)])
                    d["y%d" % i] = r2.quantize(x, num_bits=bits, signed=signed, dequantize=deq)       # per-sample mean ranges
                    d["yglob%d" % i] = r2.quantize(x, num_bits=bits, flatten_dims=(0, -1), signed=signed, dequantize=deq)
                    qp = r2.calculate_qparams(x, num_bits=bits, flatten_dims=(1, -1), reduce_dim=None)
                    d["yrow%d" % i] = r2.quantize(x, qparams=qp, signed=signed, dequantize=deq)       # per-row (weights)
                    d["zp%d" % i], d["rng%d" % i] = qp.zero_point, qp.range
                    i += 1
    d["count"] = i
    # stochastic branch with recorded noise
    log = []
    with ref_import.record_uniform(log):
        x = torch.randn(3, 7, 9, generator=g)
        d["sx"] = x
        d["sy"] = r2.quantize(x, num_bits=8, stochastic=True)
        d["sy_signed"] = r2.quantize(x, num_bits=6, signed=True, stochastic=True)
    d["snoise0"], d["snoise1"] = log
    save("q2_fq.npz", d)


def gen_grad(r2):
    g = torch.Generator().manual_seed(21)
    d = {}
    for i, (shape, fd) in enumerate([((4, 6, 5, 5), (1, -1)), ((16, 10), (0, -1)), ((3, 1024), (1, -1))]):
        x = torch.randn(shape, generator=g, requires_grad=True)
        go = torch.randn(shape, generator=g) * 1e-3 + 2e-4
        log = []
        with ref_import.record_uniform(log):
            y = r2.quantize_grad(x * 1.0, num_bits=8, flatten_dims=fd)
            y.backward(go)
        d["go%d" % i], d["dx%d" % i], d["noise%d" % i], d["fd%d" % i] = go, x.grad, log[0], np.array(fd)
    d["count"] = 3
    save("q2_grad.npz", d)


def gen_measure(r2):
    g = torch.Generator().manual_seed(22)
    d = {}
    for tag, kw in [("ema", dict(momentum=0.1)), ("meas", dict(measure=True))]:
        m = r2.QuantMeasure(8, shape_measure=(1, 1, 1, 1), flatten_dims=(1, -1), **kw).train()
        for s in range(3):
            x = torch.randn(5, 3, 4, 4, generator=g) * (1 + s) + 0.2 * s
            d["%s_x%d" % (tag, s)] = x
            d["%s_y%d" % (tag, s)] = m(x)
            d["%s_zp%d" % (tag, s)], d["%s_rng%d" % (tag, s)] = m.running_zero_point.clone(), m.running_range.clone()
            d["%s_n%d" % (tag, s)] = m.num_measured.clone()
        m.measure = False
        m.eval()
        x = torch.randn(5, 3, 4, 4, generator=g)
        d["%s_xe" % tag], d["%s_ye" % tag] = x, m(x)
    save("q2_measure.npz", d)


def gen_modules(r2):
    g = torch.Generator().manual_seed(23)
    d = {}
    cases = {"conv": (lambda: r2.QConv2d(8, 12, 3, padding=1), (4, 8, 6, 6)),
             "conv_s2_nobias": (lambda: r2.QConv2d(16, 8, 3, stride=2, padding=1, bias=False), (2, 16, 9, 9)),
             "conv_nobiprec": (lambda: r2.QConv2d(8, 8, 1, biprecision=False), (3, 8, 5, 5)),
             "linear": (lambda: r2.QLinear(32, 10), (6, 32)),
             "linear_nograd": (lambda: r2.QLinear(16, 4, num_bits_grad=None), (5, 16))}
    for tag, (mk, shape) in cases.items():
        torch.manual_seed(hash(tag) % 1000)
        m = mk().train()
        x = torch.randn(shape, generator=g, requires_grad=True)
        log = []
        with ref_import.record_uniform(log):
            out = m(x)
            go = torch.randn(out.shape, generator=g) * 1e-2
            out.backward(go)
        d[tag + "_w"] = m.weight
        if m.bias is not None:
            d[tag + "_b"], d[tag + "_db"] = m.bias, m.bias.grad
        d[tag + "_x"], d[tag + "_out"], d[tag + "_go"], d[tag + "_dx"], d[tag + "_dw"] = x, out, go, x.grad, m.weight.grad
        d[tag + "_nnoise"] = len(log)
        for k, n in enumerate(log):
            d["%s_noise%d" % (tag, k)] = n
        d[tag + "_zp"], d[tag + "_rng"] = m.quantize_input.running_zero_point, m.quantize_input.running_range
    save("q2_modules.npz", d)


TC_CASES = {"tc_conv": ("conv", dict(cin=16, cout=32, k=3, stride=1, pad=1, bias=True, biprec=True, bits_grad=8), (4, 16, 8, 8)),
            "tc_conv_s2_nobias": ("conv", dict(cin=32, cout=16, k=3, stride=2, pad=1, bias=False, biprec=True, bits_grad=8), (2, 32, 9, 9)),
            "tc_conv_1x1_nobiprec": ("conv", dict(cin=16, cout=16, k=1, stride=1, pad=0, bias=True, biprec=False, bits_grad=8), (3, 16, 5, 5)),
            "tc_conv_64": ("conv", dict(cin=64, cout=64, k=3, stride=1, pad=1, bias=False, biprec=True, bits_grad=8), (2, 64, 14, 14)),
            "tc_linear": ("lin", dict(cin=64, cout=10, bias=True, biprec=True, bits_grad=8), (6, 64)),
            "tc_linear_nograd": ("lin", dict(cin=32, cout=24, bias=True, biprec=True, bits_grad=None), (5, 32))}


def build_tc(ns, tag):
    kind, c, _ = TC_CASES[tag]
    if kind == "conv":
        return ns.QConv2d(c["cin"], c["cout"], c["k"], stride=c["stride"], padding=c["pad"], bias=c["bias"], num_bits_grad=c["bits_grad"],
                          biprecision=c["biprec"])
    return ns.QLinear(c["cin"], c["cout"], bias=c["bias"], num_bits_grad=c["bits_grad"], biprecision=c["biprec"])


def gen_modules_tc(r2):
    """Shapes the tensor-core kernels take (C_in % 16 == 0): quantize2.QConv2d / QLinear fwd + bwd with recorded noise."""
    g = torch.Generator().manual_seed(29)
    d = {}
    for i, (tag, (_, _, shape)) in enumerate(TC_CASES.items()):
        torch.manual_seed(100 + i)
        m = build_tc(r2, tag).train()
        x = torch.randn(shape, generator=g, requires_grad=True)
        log = []
        with ref_import.record_uniform(log):
            out = m(x)
            go = torch.randn(out.shape, generator=g) * 1e-2
            out.backward(go)
        d[tag + "_w"] = m.weight
        if m.bias is not None:
            d[tag + "_b"], d[tag + "_db"] = m.bias, m.bias.grad
        d[tag + "_x"], d[tag + "_out"], d[tag + "_go"], d[tag + "_dx"], d[tag + "_dw"] = x, out, go, x.grad, m.weight.grad
        d[tag + "_nnoise"] = len(log)
        for k, n in enumerate(log):
            d["%s_noise%d" % (tag, k)] = n
        d[tag + "_zp"], d[tag + "_rng"] = m.quantize_input.running_zero_point, m.quantize_input.running_range
    save("q2_modules_tc.npz", d)


def gen_rangebn(r2):
    g = torch.Generator().manual_seed(24)
    d = {}
    for tag, shape, cls in [("4d", (8, 6, 4, 4), r2.RangeBN), ("2d", (32, 8), r2.RangeBN1d)]:
        m = cls(shape[1], num_chunks=4 if tag == "4d" else 8).train()
        with torch.no_grad():
            m.bias.copy_(torch.randn(shape[1], generator=g) * 0.1)
        x = torch.randn(shape, generator=g, requires_grad=True)
        log = []
        with ref_import.record_uniform(log):
            out = m(x * 1.0)                       # the input quantizer works in place: feed a non-leaf
            go = torch.randn(out.shape, generator=g) * 1e-2
            out.backward(go)
        d[tag + "_x"], d[tag + "_out"], d[tag + "_go"], d[tag + "_dx"] = x, out, go, x.grad
        d[tag + "_gamma"], d[tag + "_beta"], d[tag + "_dgamma"], d[tag + "_dbeta"] = m.weight, m.bias, m.weight.grad, m.bias.grad
        d[tag + "_noise"] = log[0]
        d[tag + "_rm"], d[tag + "_rv"] = m.running_mean, m.running_var
        d[tag + "_zp"], d[tag + "_rng"] = m.quantize_input.running_zero_point, m.quantize_input.running_range
        m.eval()
        xe = torch.randn(shape, generator=g)
        d[tag + "_xe"], d[tag + "_oute"] = xe, m(xe.clone())
    save("q2_rangebn.npz", d)


if __name__ == "__main__":
    r2 = load_ref2()
    torch.set_num_threads(1)
    gen_fq(r2)
    gen_grad(r2)
    gen_measure(r2)
    gen_modules(r2)
    gen_modules_tc(r2)
    gen_rangebn(r2)
