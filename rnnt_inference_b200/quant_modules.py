"""Quantizers and weight packing (reference: models/quant_modules.py).

Same classes, buffers and scale algebra as the reference (scale = 127/amax, round-half-even, clamp
[-128,127], `_amax` registered as a buffer so calibrated checkpoints load).  The one deliberate difference:
`transpose_tile_weight{,_bf16}` keep their names and call signature but do not produce the AMX VNNI tiles of
quant_modules.py:177-208 — an x86 detail.  The sm_100a kernels read weights K-major through TMA
(SWIZZLE_128B) straight into tcgen05 shared-memory descriptors, for which the natural packed form is the
plain row-major [out_features, in_features] matrix, K zero-padded to a multiple of 64 (int8) / 32 (bf16)
when `padding` is set.
"""
import torch
import torch.nn as nn
from torch import Tensor


def round_and_clamp(inputs: Tensor, _min: float, _max: float):
    return torch.clamp(inputs.round(), _min, _max)


def clamp_and_round(inputs: Tensor, _min: float, _max: float):
    return torch.round(torch.clamp(inputs, _min, _max))


class QuantDescriptor:
    def __init__(self, axis=None, amax=None, mode="quant", **kwargs):
        self._axis = axis
        self._amax = amax
        self._mode = mode
        self._update_amax = kwargs.pop("update_amax", False)
        self._narrow_bound = kwargs.pop("narrow_bound", False)

    @property
    def axis(self):
        return self._axis

    @property
    def amax(self):
        return self._amax

    @property
    def mode(self):
        return self._mode

    @mode.setter
    def mode(self, mode):
        self._mode = mode

    @property
    def update_amax(self):
        return self._update_amax

    @property
    def narrow_bound(self):
        return self._narrow_bound


QUANT_LINEAR_ACTIVA = QuantDescriptor(axis=None, amax=None, mode="fake_quant", update_amax=False)
QUANT_LINEAR_WEIGHT = QuantDescriptor(axis=None, amax=None, mode="fake_quant")
QUANT_LSTM_ACTIVA = QuantDescriptor(axis=None, amax=None, mode="fake_quant", update_amax=False)
QUANT_LSTM_WEIGHT = QuantDescriptor(axis=None, amax=None, mode="fake_quant")


class TensorQuantizer(nn.Module):
    """Per-tensor symmetric int8 quantizer; modes calib / quant / fake_quant (quant_modules.py:59-152)."""

    def __init__(self, quant_desc=QuantDescriptor(), **kwargs):
        super().__init__()
        self._mode = quant_desc.mode
        self._axis = quant_desc._axis
        self._max_bound = torch.tensor(127.0, dtype=torch.float32)
        self._min_bound = -self._max_bound if quant_desc.narrow_bound else -self._max_bound - 1
        self.amax = torch.tensor(quant_desc._amax) if quant_desc._amax is not None else torch.tensor(0.0)
        self._scale = torch.tensor(0.0)
        self._name = kwargs.pop("name", "TensorQuantizer")
        self._update_amax = quant_desc.update_amax
        self._track_amax = False
        self._amaxs = []

    @property
    def mode(self):
        return self._mode

    @property
    def amax(self):
        return self._amax if hasattr(self, "_amax") else None

    @amax.setter
    def amax(self, value):
        if not hasattr(self, "_amax"):
            self.register_buffer("_amax", value)
        else:
            self._amax = value

    @property
    def scale(self):
        return self._max_bound.to(self.amax.device) / self.amax

    @scale.setter
    def scale(self, value):
        self._scale = value

    def calib_amax(self, inputs):
        cur_amax = inputs.abs().max()
        self.amax = torch.max(self.amax.to(cur_amax.device), cur_amax)
        if self._track_amax:
            self._amaxs.append(cur_amax)

    def _quant_forward(self, inputs: Tensor) -> Tensor:
        outputs = round_and_clamp(inputs * self.scale, float(self._min_bound), float(self._max_bound))
        return outputs.type(torch.int8)

    def _fake_quant_forward(self, inputs):
        if self._update_amax:
            if self._axis is not None:
                self.amax = torch.max(inputs.abs(), self._axis).values.unsqueeze(self._axis)
            else:
                self.amax = torch.max(inputs.abs())
        outputs = round_and_clamp(inputs * self.scale, float(self._min_bound), float(self._max_bound))
        outputs /= self.scale
        return outputs

    def forward(self, inputs: Tensor) -> Tensor:
        outputs = inputs
        if self._mode == "calib":
            self.calib_amax(inputs)
        elif self._mode == "quant":
            outputs = self._quant_forward(inputs)
        elif self._mode == "fake_quant":
            outputs = self._fake_quant_forward(inputs)
        return outputs

    def __str__(self):
        return (f" name=${self._name}\n mode=${self._mode}\n axis=${self._axis}\n amax=${self._amax}\n"
                f" scale=${self._scale}\n max_bound=${self._max_bound}\n min_bound=${self._min_bound}")


def _pack_k_major(weight: Tensor, padding: bool, k_mult: int) -> Tensor:
    """weight: logical W^T of shape [K, OC] (the reference passes w.transpose(1, 0)).  Returns [OC, K(_padded)]."""
    k, oc = weight.shape
    packed = weight.transpose(1, 0)
    if padding and k % k_mult != 0:
        packed = torch.nn.functional.pad(packed, (0, k_mult - k % k_mult), "constant", 0)
    return packed.contiguous()


def transpose_tile_weight_bf16(weight, padding: bool = False):
    """Same call as quant_modules.py:177-193; B200 layout: K-major [OC(_pad to 32 when padding), K] bf16."""
    packed = _pack_k_major(weight, False, 32)
    if padding and packed.shape[0] % 32 != 0:
        packed = torch.nn.functional.pad(packed, (0, 0, 0, 32 - packed.shape[0] % 32), "constant", 0.0)
    return packed.contiguous()


def transpose_tile_weight(weight, padding: bool = False):
    """Same call as quant_modules.py:196-212; B200 layout: K-major [OC, K(_pad to 64 when padding)] int8."""
    return _pack_k_major(weight, padding, 64)


class WeightQuantizer(TensorQuantizer):
    def __init__(self, quant_desc=QuantDescriptor(), **kwargs):
        super().__init__(quant_desc)

    def _quant_forward(self, inputs: Tensor, first_layer: bool = False) -> Tensor:
        outputs = round_and_clamp(inputs * self.scale, float(self._min_bound), float(self._max_bound))
        outputs = outputs.type(torch.int8)
        if first_layer:
            return transpose_tile_weight(outputs.transpose(1, 0), True)
        return transpose_tile_weight(outputs.transpose(1, 0))

    def _fake_quant_forward(self, inputs):
        if self._scale is None:
            if self._axis is not None:
                self.amax = torch.max(inputs.abs(), self._axis).values.unsqueeze(self._axis)
            else:
                self.amax = torch.max(inputs.abs())
            self.scale = self._max_bound / self.amax
        outputs = round_and_clamp(inputs * self.scale, float(self._min_bound), float(self._max_bound))
        outputs /= self.scale
        return outputs

    def forward(self, inputs: Tensor) -> Tensor:
        outputs = inputs
        if self._mode == "quant":
            outputs = self._quant_forward(inputs)
        elif self._mode == "fake_quant":
            outputs = self._fake_quant_forward(inputs)
        return outputs
