"""Generate tests/golden/*.npz by executing the UNMODIFIED reference (build container only).

    python -m oracle.gen_golden            # from the repo root; needs /root/reference

The reference has no tests or golden vectors of its own; these files are
the pin for ``oracle/qref.py`` and, through it, for the CUDA path.  Inputs are stored next to
the outputs so the tests never need the reference tree.  Noise tensors drawn by the reference's
stochastic rounding (quantize.py:93) are recorded in call order and stored too.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_import, seeding  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def npy(t):
    return t.detach().cpu().numpy()


def save(name, **arrs):
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name), **{k: (npy(v) if torch.is_tensor(v) else np.asarray(v))
                                                    for k, v in arrs.items()})
    print("wrote", name, {k: np.asarray(v).shape if not torch.is_tensor(v) else tuple(v.shape)
                           for k, v in arrs.items()})


def gen_kat(rq):
    x = torch.tensor([[-1, -.5, 0, .3, 1, 1.7]])
    w = torch.tensor([[-1, .25, .5], [0, .1, .2]])
    d = dict(x=x, w=w)
    d["y8"] = rq.quantize(x, 8, min_value=-1.0, max_value=1.0)
    d["y4"] = rq.quantize(x, 4, min_value=-1.0, max_value=1.0)
    d["y8_data"] = rq.quantize(x, 8)
    d["w_rows"] = rq.quantize(w, 8, min_value=w.min(-1)[0].view(2, 1), max_value=w.max(-1)[0].view(2, 1))
    d["const"] = rq.quantize(torch.full((1, 4), 0.5), 8)
    d["round_in"] = torch.tensor([0.5, 1.5, 2.5, 3.5, -0.5, 254.5])
    d["round_out"] = d["round_in"].clone().round_()
    save("kat.npz", **d)


def gen_fq(rq):
    g = torch.Generator().manual_seed(0)
    d = {}
    idx = 0
    for shape in [(4, 37), (3, 5, 7, 2), (8, 1031)]:
        for bits in (2, 4, 8, 16):
            x = torch.randn(shape, generator=g) * 3
            lo, hi = float(x.min()) * 0.6, float(x.max()) * 0.7       # clamping exercised
            rows = x.flatten(1)
            rshape = (shape[0],) + (1,) * (len(shape) - 1)
            rlo, rhi = rows.min(-1)[0].view(rshape), rows.max(-1)[0].view(rshape)
            d["x%d" % idx] = x
            d["meta%d" % idx] = np.array([bits, lo, hi], dtype=np.float64)
            d["given%d" % idx] = rq.quantize(x, bits, min_value=lo, max_value=hi)
            d["data%d" % idx] = rq.quantize(x, bits)
            d["rows%d" % idx] = rq.quantize(x, bits, min_value=rlo, max_value=rhi)
            d["scalar_t%d" % idx] = rq.quantize(x, bits, min_value=x.min(), max_value=x.max())
            idx += 1
    d["count"] = idx
    save("fq_cases.npz", **d)


def gen_fq_grad(rq):
    g = torch.Generator().manual_seed(1)
    d = {}
    idx = 0
    for shape, shift in [((4, 33), 0.0), ((2, 3, 5, 5), 0.0), ((6, 17), 5.0), ((6, 17), -5.0), ((1, 5), 0.0)]:
        for bits in (8, 4):
            gr = torch.randn(shape, generator=g) * 1e-3 + shift * 1e-3
            if shape == (1, 5):
                gr = torch.tensor([[-0.003, 0, 0.001, 0.004, -0.0005]])
            torch.manual_seed(7 + idx)
            log = []
            leaf = torch.zeros(shape, requires_grad=True)
            with ref_import.record_uniform(log):
                rq.quantize_grad(leaf, num_bits=bits).backward(gr)
            assert len(log) == 1
            d["g%d" % idx], d["noise%d" % idx], d["out%d" % idx] = gr, log[0], leaf.grad
            d["bits%d" % idx] = bits
            idx += 1
    d["count"] = idx
    save("fq_grad.npz", **d)


def gen_measure(rq):
    g = torch.Generator().manual_seed(2)
    d = {}
    for method in ("avg", "aciq"):
        for mom in (None, 0.9):
            tag = "%s_%s" % (method, "cum" if mom is None else "ema")
            m = rq.QuantMeasure(8, momentum=mom, method=method)
            m.train()
            for step in range(3):
                x = torch.randn((4, 3, 5, 5), generator=g) * (1 + step) + 0.3 * step
                d["%s_x%d" % (tag, step)] = x
                d["%s_y%d" % (tag, step)] = m(x)
                d["%s_buf%d" % (tag, step)] = torch.cat([m.running_min, m.running_max, m.num_measurements,
                                                         m.running_var, m.running_mean])
            m.eval()
            x = torch.randn((4, 3, 5, 5), generator=g) * 2
            d["%s_xe" % tag], d["%s_ye" % tag] = x, m(x)
    save("measure.npz", **d)


def gen_rangebn(rq):
    g = torch.Generator().manual_seed(3)
    d = {}
    for tag, shape in (("4d", (16, 4, 4, 4)), ("big", (32, 6, 3, 5)), ("2d", (32, 8))):
        C = shape[1]
        bn = rq.RangeBN(C)
        with torch.no_grad():
            bn.weight.copy_(torch.rand(C, generator=g) + 0.5)
            bn.bias.copy_(torch.randn(C, generator=g) * 0.1)
        x = (torch.randn(shape, generator=g) * 2 + 0.5).requires_grad_()
        go = torch.randn(shape, generator=g) * 1e-2
        bn.train()
        torch.manual_seed(11)
        log = []
        with ref_import.record_uniform(log):
            out = bn(x)
            out.backward(go)
        assert len(log) == 1
        d.update({tag + "_x": x, tag + "_go": go, tag + "_gamma": bn.weight, tag + "_beta": bn.bias,
                  tag + "_out": out, tag + "_dx": x.grad, tag + "_dgamma": bn.weight.grad,
                  tag + "_dbeta": bn.bias.grad, tag + "_noise": log[0],
                  tag + "_rm": bn.running_mean.clone(), tag + "_rv": bn.running_var.clone(),
                  tag + "_qm": torch.cat([bn.quantize_input.running_min, bn.quantize_input.running_max])})
        bn.eval()
        with torch.no_grad():
            d[tag + "_out_eval"] = bn(x.detach())
    save("rangebn.npz", **d)


def gen_contract(rq):
    g = torch.Generator().manual_seed(4)
    d = {}
    cases = [
        ("conv_bp", "conv", dict(cin=3, cout=8, k=3, stride=1, pad=1, groups=1, bias=False, biprec=True), (4, 3, 8, 8)),
        ("conv_s2", "conv", dict(cin=4, cout=6, k=3, stride=2, pad=1, groups=1, bias=True, biprec=True), (4, 4, 9, 9)),
        ("conv_dw", "conv", dict(cin=6, cout=6, k=3, stride=1, pad=1, groups=6, bias=True, biprec=True), (4, 6, 7, 7)),
        ("conv_1x1", "conv", dict(cin=8, cout=5, k=1, stride=2, pad=0, groups=1, bias=False, biprec=True), (4, 8, 6, 6)),
        ("conv_nb", "conv", dict(cin=3, cout=4, k=3, stride=1, pad=1, groups=1, bias=True, biprec=False), (4, 3, 6, 6)),
        ("lin_bp", "lin", dict(cin=24, cout=10, bias=True, biprec=True), (8, 24)),
        ("lin_nb", "lin", dict(cin=24, cout=10, bias=True, biprec=False), (8, 24)),
    ]
    for tag, kind, c, shape in cases:
        if kind == "conv":
            m = rq.QConv2d(c["cin"], c["cout"], c["k"], stride=c["stride"], padding=c["pad"], groups=c["groups"],
                           bias=c["bias"], num_bits=8, num_bits_weight=8, num_bits_grad=8, biprecision=c["biprec"])
        else:
            m = rq.QLinear(c["cin"], c["cout"], bias=c["bias"], num_bits=8, num_bits_weight=8, num_bits_grad=8,
                           biprecision=c["biprec"])
        with torch.no_grad():
            m.weight.copy_(torch.randn(m.weight.shape, generator=g) * 0.2)
            if m.bias is not None:
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
        x = (torch.randn(shape, generator=g) + 0.2).requires_grad_()
        m.train()
        torch.manual_seed(21)
        log = []
        with ref_import.record_uniform(log):
            out = m(x)
            go = torch.randn(out.shape, generator=g) * 1e-2
            out.backward(go)
        assert len(log) == 1
        d.update({tag + "_x": x, tag + "_w": m.weight, tag + "_go": go, tag + "_out": out, tag + "_dx": x.grad,
                  tag + "_dw": m.weight.grad, tag + "_noise": log[0],
                  tag + "_wmin": m.weight_min.detach().flatten(), tag + "_wmax": m.weight_max.detach().flatten()})
        if m.bias is not None:
            d[tag + "_b"], d[tag + "_db"] = m.bias, m.bias.grad
    save("contract.npz", **d)


def gen_resnet20(ref_models):
    torch.manual_seed(123)
    model = ref_models.resnet_quantized(depth=20, dataset="cifar10")
    seeding.reseed_parameters(model, 1234)
    x, y = seeding.synthetic_batch((16, 3, 32, 32), 10, seed=123)
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    crit = torch.nn.CrossEntropyLoss()
    model.train()
    torch.manual_seed(77)
    losses, logits0 = [], None
    for step in range(3):
        out = model(x)
        if step == 0:
            logits0 = out.detach().clone()
        loss = crit(out, y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    sd = model.state_dict()
    keys = sorted(sd)
    sums = np.array([float(sd[k].double().sum()) for k in keys])
    asums = np.array([float(sd[k].double().abs().sum()) for k in keys])
    model.eval()
    with torch.no_grad():
        ev = model(x)
    save("resnet20.npz", losses=np.array(losses), logits0=logits0, keys=np.array(keys), sums=sums, asums=asums,
         eval_logits=ev)


def main():
    ref_models, rq = ref_import.load()
    torch.set_num_threads(1)          # fixed summation order for the fixtures
    gen_kat(rq)
    gen_fq(rq)
    gen_fq_grad(rq)
    gen_measure(rq)
    gen_rangebn(rq)
    gen_contract(rq)
    gen_resnet20(ref_models)


if __name__ == "__main__":
    main()
