"""qtorch_b200: B200-native (sm_100a) implementation of the fake-quantization hot path of
amishacorns/quantized.pytorch (models/modules/quantize.py) behind the reference's Python API.

``quantize``    - the API mirror (QuantMeasure, QConv2d, QLinear, RangeBN, quantize, quantize_grad ...)
``models``      - resnet_quantized / mobilenet_quantized consumers
``_abi``        - ctypes binding of the C-ABI library ``libqb200.so`` (include/qb200.h)
There is no CPU fallback: every compute entry point raises if the CUDA library cannot be loaded.
"""
__version__ = "0.1.0"
