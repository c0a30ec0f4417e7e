"""One eager ResNet-18 (or MobileNet: argv[2]) b256 8-bit train step between cudaProfilerStart/Stop, for ncu
(`ncu --profile-from-start off ...`).  Not a benchmark: numbers printed under a profiler are never reported."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import models, quantize as Q  # noqa: E402


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    name = sys.argv[2] if len(sys.argv) > 2 else "resnet18"          # or "mobilenet" (BASELINE configs[3])
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    model = (models.mobilenet_quantized(Q) if name == "mobilenet" else models.resnet_quantized(Q, depth=18, dataset="imagenet"))
    model = model.to(dev).train()
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
    from qtorch_b200 import glue
    crit = glue.CrossEntropyLoss()
    x = torch.randn(batch, 3, 224, 224, device=dev)
    y = torch.randint(0, 1000, (batch,), device=dev)

    def step():
        loss = crit(model(x), y) + model.regularization()
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    loss = step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("loss", float(loss))


if __name__ == "__main__":
    main()
