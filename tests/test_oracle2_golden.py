"""Pin oracle/qref2.py (quantize2.py flavour, row A9) bit-exactly against tests/golden/q2_*.npz, which
oracle/gen_golden2.py produced by running the UNMODIFIED reference quantize2.py.  CPU only."""
import numpy as np
import torch

from conftest import golden
from oracle import qref2


def T(a):
    return torch.from_numpy(np.asarray(a))


def same(a, b, what=""):
    a = a.detach().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert np.array_equal(a.view(np.int32), b.view(np.int32)), "%s: max abs diff %g" % (what, np.abs(a - b).max())


def test_fq2_cases():
    g = golden("q2_fq.npz")
    for i in range(int(g["count"])):
        x = T(g["x%d" % i])
        bits, signed, deq = [int(v) for v in g["meta%d" % i]]
        same(qref2.quantize(x, bits, signed=bool(signed), dequantize=bool(deq)), g["y%d" % i], "mean %d" % i)
        same(qref2.quantize(x, bits, flatten_dims=(0, -1), signed=bool(signed), dequantize=bool(deq)), g["yglob%d" % i], "glob")
        zp, rng = qref2.qparams(x, (1, -1), None)
        same(zp, g["zp%d" % i])
        same(rng, g["rng%d" % i])
        same(qref2.fq2(x, zp, rng, bits, bool(signed), bool(deq)), g["yrow%d" % i], "row %d" % i)
    x = T(g["sx"])
    same(qref2.quantize(x, 8, noise=T(g["snoise0"])), g["sy"], "stochastic")
    same(qref2.quantize(x, 6, signed=True, noise=T(g["snoise1"])), g["sy_signed"], "stochastic signed")


def test_grad2():
    g = golden("q2_grad.npz")
    for i in range(int(g["count"])):
        fd = tuple(int(v) for v in g["fd%d" % i])
        same(qref2.quantize_grad_backward(T(g["go%d" % i]), 8, fd, 0, False, T(g["noise%d" % i])), g["dx%d" % i], "grad %d" % i)


def test_measure2():
    g = golden("q2_measure.npz")
    for tag, measure in (("ema", False), ("meas", True)):
        st = dict(running_zero_point=torch.zeros(1, 1, 1, 1), running_range=torch.zeros(1, 1, 1, 1), num_measured=torch.zeros(1))
        for s in range(3):
            x = T(g["%s_x%d" % (tag, s)])
            zp, rng = qref2.measure_update(st, x, 8, (1, -1), 0.1, measure)
            y = x if measure else qref2.fq2(x, zp, rng, 8)
            same(y, g["%s_y%d" % (tag, s)])
            same(st["running_zero_point"], g["%s_zp%d" % (tag, s)])
            same(st["running_range"], g["%s_rng%d" % (tag, s)])
            same(st["num_measured"], g["%s_n%d" % (tag, s)])
        same(qref2.fq2(T(g[tag + "_xe"]), st["running_zero_point"], st["running_range"], 8), g[tag + "_ye"])


CASES = {"conv": (((1, 1), (1, 1), (1, 1), 1), True, 8), "conv_s2_nobias": (((2, 2), (1, 1), (1, 1), 1), True, 8),
         "conv_nobiprec": (((1, 1), (0, 0), (1, 1), 1), False, 8), "linear": (None, True, 8), "linear_nograd": (None, True, None)}


def test_modules2():
    g = golden("q2_modules.npz")
    for tag, (conv, biprec, bits_g) in CASES.items():
        x = T(g[tag + "_x"]).clone().requires_grad_()
        w = T(g[tag + "_w"]).clone().requires_grad_()
        b = T(g[tag + "_b"]).clone().requires_grad_() if tag + "_b" in g.files else None
        st = dict(running_zero_point=torch.zeros(1, 1, 1, 1) if conv else torch.zeros(1),
                  running_range=torch.zeros(1, 1, 1, 1) if conv else torch.zeros(1), num_measured=torch.zeros(1))
        noise = [T(g["%s_noise%d" % (tag, k)]) for k in range(int(g[tag + "_nnoise"]))]
        out = qref2.qcontract(x, w, b, st, conv, 8, 8, bits_g, biprec, True, noise)
        out.backward(T(g[tag + "_go"]))
        assert torch.allclose(out, T(g[tag + "_out"]), rtol=1e-5, atol=1e-6), tag
        assert torch.allclose(x.grad, T(g[tag + "_dx"]), rtol=1e-4, atol=1e-7), tag
        assert torch.allclose(w.grad, T(g[tag + "_dw"]), rtol=1e-4, atol=1e-6), tag
        if b is not None:
            assert torch.allclose(b.grad, T(g[tag + "_db"]), rtol=1e-5, atol=1e-7), tag
        same(st["running_zero_point"], g[tag + "_zp"])
        same(st["running_range"], g[tag + "_rng"])
        assert not noise


def test_rangebn2():
    g = golden("q2_rangebn.npz")
    for tag, chunks in (("4d", 4), ("2d", 8)):
        x = T(g[tag + "_x"]).clone().requires_grad_()
        gamma = T(g[tag + "_gamma"]).clone().requires_grad_()
        beta = T(g[tag + "_beta"]).clone().requires_grad_()
        shp = (1, 1, 1, 1) if tag == "4d" else (1, 1)
        qst = dict(running_zero_point=torch.zeros(shp), running_range=torch.zeros(shp), num_measured=torch.zeros(1))
        st = dict(running_mean=torch.zeros(x.shape[1]), running_var=torch.zeros(x.shape[1]))
        zp, rng = qref2.measure_update(qst, x, 8, (1, -1), 0.1)
        xq = qref2.ste(x, qref2.fq2(x, zp, rng, 8))
        x4 = xq if xq.dim() == 4 else xq.unsqueeze(-1).unsqueeze(-1)
        noise = [T(g[tag + "_noise"]).reshape(x4.shape)]
        out = qref2.GradQ.apply(qref2.rangebn2_forward(x4, st, gamma, beta, True, 0.1, chunks), 8, (1, -1), noise)
        out = out.reshape(g[tag + "_out"].shape)
        out.backward(T(g[tag + "_go"]))
        same(out, g[tag + "_out"], tag + " out")
        for a, k in ((x.grad, "_dx"), (gamma.grad, "_dgamma"), (beta.grad, "_dbeta")):
            assert torch.allclose(a, T(g[tag + k]), rtol=1e-5, atol=1e-8), (tag, k)
        same(st["running_mean"], g[tag + "_rm"])
        same(st["running_var"], g[tag + "_rv"])
        xe = T(g[tag + "_xe"])
        xqe = qref2.fq2(xe, qst["running_zero_point"], qst["running_range"], 8)
        xqe = xqe if xqe.dim() == 4 else xqe.unsqueeze(-1).unsqueeze(-1)
        same(qref2.rangebn2_forward(xqe, st, gamma, beta, False).reshape(g[tag + "_oute"].shape), g[tag + "_oute"], tag + " eval")
