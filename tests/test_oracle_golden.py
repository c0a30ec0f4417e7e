"""Pin the CPU oracle (oracle/qref*.py) against outputs of the unmodified reference
(tests/golden/*.npz, produced by oracle/gen_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import qref, qref_modules, seeding


def T(a):
    return torch.from_numpy(np.asarray(a))


def same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape
    assert np.array_equal(a.view(np.int32), b.view(np.int32)), "max abs diff %g" % np.abs(a - b).max()


def test_kat():
    g = golden("kat.npz")
    x, w = T(g["x"]), T(g["w"])
    same(qref.fq_affine(x, 8, -1.0, 1.0), g["y8"])
    same(qref.fq_affine(x, 4, -1.0, 1.0), g["y4"])
    same(qref.fq_affine(x, 8), g["y8_data"])
    same(qref.fq_affine(w, 8, w.min(-1)[0].view(2, 1), w.max(-1)[0].view(2, 1)), g["w_rows"])
    same(qref.fq_affine(torch.full((1, 4), 0.5), 8), g["const"])
    same(T(g["round_in"]).round(), g["round_out"])
    # SURVEY.md section 4, KAT 1 (literal)
    assert g["y8"].tolist() == [[-1.0, -0.498039186000824, -0.0039215087890625, 0.301960825920105, 1.0, 1.0]]


def test_fq_cases():
    g = golden("fq_cases.npz")
    for i in range(int(g["count"])):
        x = T(g["x%d" % i])
        bits, lo, hi = g["meta%d" % i]
        bits = int(bits)
        y, codes = qref.fq_affine(x, bits, float(lo), float(hi), want_codes=True)
        same(y, g["given%d" % i])
        assert codes.min() >= 0 and codes.max() <= 2 ** bits - 1 and torch.equal(codes, codes.round())
        same(qref.fq_affine(x, bits), g["data%d" % i])
        rshape = (x.shape[0],) + (1,) * (x.dim() - 1)
        rows = x.flatten(1)
        same(qref.fq_affine(x, bits, rows.min(-1)[0].view(rshape), rows.max(-1)[0].view(rshape)), g["rows%d" % i])
        same(qref.fq_affine(x, bits, x.min(), x.max()), g["scalar_t%d" % i])


def test_fq_grad():
    g = golden("fq_grad.npz")
    for i in range(int(g["count"])):
        gr, noise = T(g["g%d" % i]), T(g["noise%d" % i])
        same(qref.fq_grad(gr, int(g["bits%d" % i]), noise), g["out%d" % i])
    # SURVEY KAT 6: scale/zero-point of the literal example
    gr = T(g["g8"])
    s = qref.affine_scale(float(gr.min()), float(gr.max()), 8)
    assert s == 2.7450981239477793e-05 and qref.zero_point(float(gr.min()), s, 8) == 109


@pytest.mark.parametrize("method", ["avg", "aciq"])
@pytest.mark.parametrize("mom", [None, 0.9])
def test_measure(method, mom):
    g = golden("measure.npz")
    tag = "%s_%s" % (method, "cum" if mom is None else "ema")
    m = qref_modules.QuantMeasure(8, momentum=mom, method=method).train()
    for step in range(3):
        y = m(T(g["%s_x%d" % (tag, step)]))
        same(y, g["%s_y%d" % (tag, step)])
        buf = torch.cat([m.running_min, m.running_max, m.num_measurements, m.running_var, m.running_mean])
        same(buf, g["%s_buf%d" % (tag, step)])
    m.eval()
    same(m(T(g["%s_xe" % tag])), g["%s_ye" % tag])


class Replay:
    def __init__(self, tensors):
        self.q = list(tensors)

    def __call__(self, shape):
        return self.q.pop(0)


@pytest.mark.parametrize("tag", ["4d", "big", "2d"])
def test_rangebn(tag):
    g = golden("rangebn.npz")
    C = g[tag + "_gamma"].shape[0]
    bn = qref_modules.RangeBN(C).train()
    with torch.no_grad():
        bn.weight.copy_(T(g[tag + "_gamma"]))
        bn.bias.copy_(T(g[tag + "_beta"]))
    x = T(g[tag + "_x"]).clone().requires_grad_()
    prev = qref.set_noise_provider(Replay([T(g[tag + "_noise"])]))
    try:
        out = bn(x)
        out.backward(T(g[tag + "_go"]))
    finally:
        qref.set_noise_provider(prev)
    same(out.detach(), g[tag + "_out"])
    same(x.grad, g[tag + "_dx"])
    same(bn.weight.grad, g[tag + "_dgamma"])
    same(bn.bias.grad, g[tag + "_dbeta"])
    same(bn.running_mean, g[tag + "_rm"])
    same(bn.running_var, g[tag + "_rv"])
    bn.eval()
    with torch.no_grad():
        same(bn(x.detach()), g[tag + "_out_eval"])


CONTRACT = {
    "conv_bp": dict(cin=3, cout=8, k=3, stride=1, pad=1, groups=1, bias=False, biprec=True),
    "conv_s2": dict(cin=4, cout=6, k=3, stride=2, pad=1, groups=1, bias=True, biprec=True),
    "conv_dw": dict(cin=6, cout=6, k=3, stride=1, pad=1, groups=6, bias=True, biprec=True),
    "conv_1x1": dict(cin=8, cout=5, k=1, stride=2, pad=0, groups=1, bias=False, biprec=True),
    "conv_nb": dict(cin=3, cout=4, k=3, stride=1, pad=1, groups=1, bias=True, biprec=False),
    "lin_bp": dict(cin=24, cout=10, bias=True, biprec=True),
    "lin_nb": dict(cin=24, cout=10, bias=True, biprec=False),
}


def build_contract(ns, tag, g):
    c = CONTRACT[tag]
    if tag.startswith("conv"):
        m = ns.QConv2d(c["cin"], c["cout"], c["k"], stride=c["stride"], padding=c["pad"], groups=c["groups"],
                       bias=c["bias"], num_bits=8, num_bits_weight=8, num_bits_grad=8, biprecision=c["biprec"])
    else:
        m = ns.QLinear(c["cin"], c["cout"], bias=c["bias"], num_bits=8, num_bits_weight=8, num_bits_grad=8,
                       biprecision=c["biprec"])
    with torch.no_grad():
        m.weight.copy_(T(g[tag + "_w"]))
        if m.bias is not None:
            m.bias.copy_(T(g[tag + "_b"]))
    return m.train()


@pytest.mark.parametrize("tag", sorted(CONTRACT))
def test_contract(tag):
    g = golden("contract.npz")
    m = build_contract(qref_modules, tag, g)
    x = T(g[tag + "_x"]).clone().requires_grad_()
    prev = qref.set_noise_provider(Replay([T(g[tag + "_noise"])]))
    try:
        out = m(x)
        out.backward(T(g[tag + "_go"]))
    finally:
        qref.set_noise_provider(prev)
    # same ATen conv/addmm kernels on both sides -> exact on this container; tolerance guards
    # against a different oneDNN/MKL blocking on another host
    np.testing.assert_allclose(out.detach().numpy(), g[tag + "_out"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(x.grad.numpy(), g[tag + "_dx"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(m.weight.grad.numpy(), g[tag + "_dw"], rtol=1e-5, atol=1e-7)
    if m.bias is not None:
        np.testing.assert_allclose(m.bias.grad.numpy(), g[tag + "_db"], rtol=1e-5, atol=1e-7)
    same(m.weight_min.detach().flatten(), g[tag + "_wmin"])
    same(m.weight_max.detach().flatten(), g[tag + "_wmax"])


def test_resnet20_three_steps():
    """Config 1 of BASELINE.json at batch 16: three SGD steps of ResNet-20/CIFAR through the
    oracle modules reproduce the reference's losses (same seeds => same noise draws)."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import models
    g = golden("resnet20.npz")
    # The fixture was generated single-threaded; with 8 threads oneDNN's different summation order
    # flips a few codes and this chaotic (lr=0.1, batch 16, random data) run drifts 0.5% / 5% at
    # steps 2 / 3 - that drift is the tolerance band used for the CUDA loss-curve tests.
    nthreads = torch.get_num_threads()
    torch.set_num_threads(1)
    try:
        _resnet20_three_steps(models, g)
    finally:
        torch.set_num_threads(nthreads)


def _resnet20_three_steps(models, g):
    torch.manual_seed(123)
    model = models.resnet_quantized(qref_modules, depth=20, dataset="cifar10")
    assert sorted(model.state_dict()) == [str(k) for k in g["keys"]]
    seeding.reseed_parameters(model, 1234)
    x, y = seeding.synthetic_batch((16, 3, 32, 32), 10, seed=123)
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    crit = torch.nn.CrossEntropyLoss()
    model.train()
    torch.manual_seed(77)
    losses = []
    for step in range(3):
        out = model(x)
        if step == 0:
            np.testing.assert_allclose(out.detach().numpy(), g["logits0"], rtol=1e-4, atol=1e-5)
        loss = crit(out, y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    np.testing.assert_allclose(losses, g["losses"], rtol=1e-5)
    sd = model.state_dict()
    sums = np.array([float(sd[k].double().sum()) for k in sorted(sd)])
    asums = np.array([float(sd[k].double().abs().sum()) for k in sorted(sd)])
    np.testing.assert_allclose(asums, g["asums"], rtol=5e-3, atol=1e-4)
    assert np.abs(sums - g["sums"]).max() <= 5e-3 * max(1.0, np.abs(g["asums"]).max())
