"""Quantised linear modules (reference: models/quant_linear.py:10-65).

`QuantLinear` = optional calibration / fake-quant of the input + F.linear (pure torch).  `iLinear` = int8 x int8
inner product with the bias folded into the accumulator domain and one output scale; here the contraction is the
tcgen05 int8 GEMM (P.linear -> rnnt_gemm_i8_i32 + rnnt_dequant_i32_f32).  The RNNT model at the reference HEAD does
not instantiate either (modeling_rnnt.py:54,60,65 are commented out); the classes are kept for surface parity.
"""
import torch
from torch import Tensor
from torch.nn import functional as F
from torch.nn.parameter import Parameter

from . import _C as P
from .quant_modules import QUANT_LINEAR_ACTIVA, QUANT_LINEAR_WEIGHT, TensorQuantizer, WeightQuantizer


class QuantLinear(torch.nn.Linear):
    def __init__(self, in_features: int, out_features: int, bias: bool = True, **kwargs) -> None:
        super().__init__(in_features, out_features, bias, **kwargs)

    def _init_quantizers(self, run_mode=None) -> None:
        QUANT_LINEAR_ACTIVA.mode = "calib" if run_mode == "calib" else run_mode
        QUANT_LINEAR_WEIGHT.mode = "quant" if run_mode == "calib" else run_mode
        self.input_quantizer = TensorQuantizer(QUANT_LINEAR_ACTIVA)
        self.weight_quantizer = WeightQuantizer(QUANT_LINEAR_WEIGHT)

    def _init_parameters(self, weight: Tensor, bias: Tensor) -> None:
        self.weight = weight
        self.bias = bias

    def _quant_parameters(self, run_mode) -> None:
        if run_mode == "fake_quant":
            self.weight_quantizer.amax = torch.max(self.weight.abs())
            self.weight = Parameter(self.weight_quantizer(self.weight), requires_grad=False)

    def forward(self, x: Tensor) -> Tensor:
        quantizer = getattr(self, "input_quantizer", None)
        if quantizer is not None and quantizer.mode is not None:
            x = quantizer(x)
        return F.linear(x, self.weight, self.bias)


class iLinear(QuantLinear):
    def __init__(self, in_features: int, out_features: int, bias: bool = True, **kwargs) -> None:
        super().__init__(in_features, out_features, bias, **kwargs)

    def _quant_parameters(self, run_mode) -> None:
        if run_mode == "quant":
            self.weight_quantizer.amax = torch.max(self.weight.abs())
            self.weight = Parameter(self.weight_quantizer(self.weight), requires_grad=False)
            b_scale = self.input_quantizer.scale * self.weight_quantizer.scale
            self.bias = Parameter(self.bias * b_scale, requires_grad=False)
            self.o_scale = 1 / b_scale

    def forward(self, x: Tensor) -> Tensor:
        # x int8 [.., K]; weight int8 [O, K] (K-major pack); y fp32 = (float(acc) + bias) * o_scale
        return P.linear(x, self.weight, self.bias, self.o_scale.item(), None)
