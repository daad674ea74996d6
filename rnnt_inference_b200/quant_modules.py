"""Symmetric int8 quantisers and weight packing for the B200 path.

Interface contract with the reference (models/quant_modules.py), kept so that calibrated MLPerf checkpoints and the
loader code of quant_lstm.py / quant_linear.py work unchanged:
  * a quantiser is an nn.Module whose range lives in a buffer named `_amax` (that is the key calibrated checkpoints
    carry), exposes `.amax`, `.scale` (= 127 / amax), `.mode`, and is called on a tensor;
  * mode "quant" -> int8 = clamp(round_half_even(x * scale), -128, 127); "fake_quant" -> the same grid back in fp32;
    "calib" -> passes x through and widens amax to cover it; None / "f32" -> identity;
  * descriptors QUANT_{LSTM,LINEAR}_{ACTIVA,WEIGHT} are the mutable mode switches the LSTM / Linear builders flip
    before constructing their quantisers;
  * `transpose_tile_weight{,_bf16}(w_t, padding)` take the logical W^T [K, OC] like the reference's packers.
Everything else is this package's own: there is no VNNI / AMX tiling (quant_modules.py:177-208 is an x86 layout) —
the sm_100a kernels read weights K-major through TMA (SWIZZLE_128B) straight into tcgen05 shared-memory
descriptors, so the packed form is the plain row-major [OC, K] matrix, K zero-padded to 64 (int8) on request.
"""
import torch
from torch import Tensor, nn

INT8_MAX = 127.0
INT8_MIN = -128.0


def round_and_clamp(x: Tensor, lo: float, hi: float) -> Tensor:
    """Round half to even, then saturate (the order matters only at +-0.5 beyond the bounds, where both agree)."""
    return x.round().clamp_(lo, hi)


class QuantDescriptor:
    """Mode switch shared by the quantisers built from it: mode in {None, "f32", "calib", "quant", "fake_quant"}."""

    def __init__(self, axis=None, amax=None, mode="quant", update_amax=False, narrow_bound=False):
        self.axis = axis
        self.amax = amax
        self.mode = mode
        self.update_amax = update_amax
        self.narrow_bound = narrow_bound


QUANT_LINEAR_ACTIVA = QuantDescriptor(mode="fake_quant")
QUANT_LINEAR_WEIGHT = QuantDescriptor(mode="fake_quant")
QUANT_LSTM_ACTIVA = QuantDescriptor(mode="fake_quant")
QUANT_LSTM_WEIGHT = QuantDescriptor(mode="fake_quant")


class TensorQuantizer(nn.Module):
    """Per-tensor symmetric int8 quantiser of activations."""

    def __init__(self, quant_desc=None, name="TensorQuantizer"):
        super().__init__()
        desc = quant_desc if quant_desc is not None else QuantDescriptor()
        self._mode = desc.mode
        self._axis = desc.axis
        self._update_amax = desc.update_amax
        self._lo = -INT8_MAX if desc.narrow_bound else INT8_MIN
        self._name = name
        self.register_buffer("_amax", torch.tensor(float(desc.amax) if desc.amax is not None else 0.0))

    # -- range / scale
    @property
    def mode(self):
        return self._mode

    @property
    def amax(self) -> Tensor:
        return self._amax

    @amax.setter
    def amax(self, value):
        self._amax = torch.as_tensor(value, dtype=torch.float32).detach().to(self._amax.device)

    @property
    def scale(self) -> Tensor:
        # ONE correctly rounded fp32 division, tensor / tensor like the reference's `_max_bound / amax`.  (A Python float
        # divided by a tensor is evaluated by torch as reciprocal() * scalar — two roundings, one ulp off for some amax,
        # which moves weights that sit on a .5 rounding tie.)
        return torch.full_like(self._amax, INT8_MAX) / self._amax

    def _range_of(self, x: Tensor) -> Tensor:
        if self._axis is None:
            return x.abs().max()
        return x.abs().max(self._axis).values.unsqueeze(self._axis)

    def calib_amax(self, x: Tensor):
        self.amax = torch.maximum(self._amax, x.abs().max().to(self._amax.device))

    # -- the three mappings
    def _grid(self, x: Tensor) -> Tensor:
        return round_and_clamp(x * self.scale, self._lo, INT8_MAX)

    def _quant_forward(self, x: Tensor) -> Tensor:
        return self._grid(x).to(torch.int8)

    def _fake_quant_forward(self, x: Tensor) -> Tensor:
        if self._update_amax:
            self.amax = self._range_of(x)
        return self._grid(x) / self.scale

    def forward(self, x: Tensor) -> Tensor:
        if self._mode == "calib":
            self.calib_amax(x)
            return x
        if self._mode == "quant":
            return self._quant_forward(x)
        if self._mode == "fake_quant":
            return self._fake_quant_forward(x)
        return x

    def extra_repr(self):
        return f"name={self._name}, mode={self._mode}, axis={self._axis}, amax={float(self._amax.max()):.6g}"


class WeightQuantizer(TensorQuantizer):
    """Per-layer symmetric int8 quantiser of weights; never calibrates from data (its amax is set from the weights)."""

    def __init__(self, quant_desc=None, name="WeightQuantizer"):
        super().__init__(quant_desc, name)

    def _quant_forward(self, w: Tensor, first_layer: bool = False) -> Tensor:
        """[OC, K] fp32 -> packed int8 [OC, K] (K zero-padded to a multiple of 64 for the first layer, 240 -> 256)."""
        return transpose_tile_weight(self._grid(w).to(torch.int8).transpose(1, 0), padding=first_layer)

    def forward(self, w: Tensor) -> Tensor:
        if self._mode == "quant":
            return self._quant_forward(w)
        if self._mode == "fake_quant":
            if float(self._amax.max()) == 0.0:
                self.amax = self._range_of(w)
            return self._grid(w) / self.scale
        return w


def _k_major(w_t: Tensor, k_multiple: int, pad_k: bool) -> Tensor:
    """w_t: logical W^T [K, OC] -> contiguous [OC, K], K zero-padded up to `k_multiple` when asked."""
    w = w_t.transpose(1, 0)
    k = w.shape[1]
    if pad_k and k % k_multiple:
        w = torch.nn.functional.pad(w, (0, k_multiple - k % k_multiple))
    return w.contiguous()


def transpose_tile_weight(weight: Tensor, padding: bool = False) -> Tensor:
    """int8 packer: K-major [OC, K], K padded to a multiple of 64 when `padding` (one 128-byte TMA row = 2 x 64)."""
    return _k_major(weight, 64, padding)


def transpose_tile_weight_bf16(weight: Tensor, padding: bool = False) -> Tensor:
    """bf16 packer: K-major [OC, K]; `padding` rounds OC up to 32 rows of zeros (fc2: 29 -> 32 logits)."""
    w = _k_major(weight, 32, False)
    if padding and w.shape[0] % 32:
        w = torch.nn.functional.pad(w, (0, 0, 0, 32 - w.shape[0] % 32))
    return w.contiguous()
