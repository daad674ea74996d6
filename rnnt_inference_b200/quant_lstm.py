"""Quantised LSTM modules (reference: models/quant_lstm.py).

Same classes, constructor signatures, attribute names and scale algebra as the reference so that a calibrated
MLPerf checkpoint loads the same way:

  iLSTM            int8 stack; forward -> P.lstm_amx_int8 (quant_lstm.py:80-102)
  QuantLSTM        fp32 / calib / fake_quant stack, python loop over layers (quant_lstm.py:105-128)
  QuantLSTMLayer   one layer of the above, python loop over timesteps (quant_lstm.py:131-183)
  iLSTMLayer       weight quantisation + bias folding of one int8 layer (quant_lstm.py:186-230)

On the B200 path the int8 stack runs as: one tcgen05 projection GEMM per layer over all timesteps + one
persistent recurrent kernel per layer (see csrc/gemm_i8.cu, csrc/lstm_recurrent.cu).
"""
from typing import List, Optional, Tuple

import torch
from torch import Tensor, nn
from torch.nn import functional as F
from torch.nn.parameter import Parameter

from . import _C as P
from .quant_modules import (QUANT_LSTM_ACTIVA, QUANT_LSTM_WEIGHT, TensorQuantizer, WeightQuantizer)


class iLSTM(torch.nn.LSTM):
    def __init__(self, input_size, hidden_size, num_layers, skip_quant_y, **kwargs):
        super().__init__(input_size, hidden_size, num_layers, **kwargs)
        self.skip_quant_y = skip_quant_y
        self.weights = []                        # per layer [w_ih int8, w_hh int8, b_ih, b_hh_folded]
        self.rb_scale = torch.zeros(num_layers)  # host-side scalars, read with .item() like the reference
        self.in_scale = torch.zeros(num_layers)
        self.out_scale = torch.zeros(num_layers)
        self.run_mode = None

    # -- construction of the per-layer cells (quant_lstm.py:21-53)
    def _init_layers(self, run_mode=None):
        self.run_mode = run_mode
        activation_mode = "calib" if run_mode == "calib" else run_mode
        weight_mode = "quant" if run_mode == "calib" else run_mode
        QUANT_LSTM_ACTIVA.mode = activation_mode
        QUANT_LSTM_WEIGHT.mode = weight_mode
        in_features = self.input_size
        for layer in range(self.num_layers):
            cell_cls = iLSTMLayer if run_mode == "quant" else QuantLSTMLayer
            cell = cell_cls(in_features, self.hidden_size)
            if run_mode is not None and run_mode != "f32":
                cell._init_quantizers(WeightQuantizer(QUANT_LSTM_WEIGHT), TensorQuantizer(QUANT_LSTM_ACTIVA),
                                      TensorQuantizer(QUANT_LSTM_ACTIVA))
            cell._init_parameters(*self.all_weights[layer][:4])
            setattr(self, f"lstm{layer}", cell)
            in_features = self.hidden_size

    # -- quantise + drop the fp32 copies held by nn.LSTM (quant_lstm.py:55-64)
    def _process_parameters(self, run_mode):
        for layer in range(self.num_layers):
            if run_mode in ("quant", "fake_quant"):
                getattr(self, f"lstm{layer}")._quant_parameters(self.weights, self.rb_scale)
            for name in self._all_weights[layer][:4]:
                delattr(self, name)

    # -- out_scale of layer l = in_scale of layer l+1 (quant_lstm.py:66-78)
    def _propagate_quantizers(self):
        for layer in range(self.num_layers):
            cell = getattr(self, f"lstm{layer}")
            if layer + 1 < self.num_layers:
                cell.output_quantizer = getattr(self, f"lstm{layer + 1}").input_quantizer
            self.in_scale[layer] = cell.input_quantizer.scale.item()
            self.out_scale[layer] = cell.output_quantizer.scale.item()

    def _apply(self, fn, recurse=True):
        # self.weights is a plain list (as in the reference): move it with the module, keep the scale vectors on the host
        out = super()._apply(fn, recurse)
        self.weights = [[fn(t) for t in layer] for layer in self.weights]
        return out

    def forward(self, x: Tensor, hx: Optional[List[Tensor]], cx: Optional[List[Tensor]], tile_steps=None, true_rows=None
                ) -> Tuple[Tensor, List[Tensor], List[Tensor]]:
        if hx is None and cx is None:
            n = x.size(1)
            hx = [torch.zeros((n, self.hidden_size), dtype=torch.int8, device=x.device) for _ in range(self.num_layers)]
            cx = [torch.zeros((n, self.hidden_size), dtype=torch.float16, device=x.device) for _ in range(self.num_layers)]
        return P.lstm_amx_int8(x, hx, cx, self.weights, self.rb_scale, self.in_scale, self.out_scale, self.skip_quant_y,
                               tile_steps=tile_steps, true_rows=true_rows)


class QuantLSTM(iLSTM):
    """fp32 / calib / fake_quant stack: plain torch, device-agnostic (quant_lstm.py:105-128)."""

    def __init__(self, input_size, hidden_size, num_layers, skip_quant_y, **kwargs):
        super().__init__(input_size, hidden_size, num_layers, skip_quant_y, **kwargs)

    def forward(self, x: Tensor, hx: Optional[List[Tensor]], cx: Optional[List[Tensor]]
                ) -> Tuple[Tensor, List[Tensor], List[Tensor]]:
        if hx is None and cx is None:
            n = x.size(1)
            hx = [torch.zeros((n, self.hidden_size), dtype=torch.float32, device=x.device) for _ in range(self.num_layers)]
            cx = [torch.zeros((n, self.hidden_size), dtype=torch.float32, device=x.device) for _ in range(self.num_layers)]
        for layer in range(self.num_layers):
            x, hx[layer], cx[layer] = getattr(self, f"lstm{layer}")(x, hx[layer], cx[layer])
        return x, hx, cx


class QuantLSTMLayer(nn.LSTMCell):
    def __init__(self, input_size: int, hidden_size: int, **kwargs) -> None:
        super().__init__(input_size, hidden_size, **kwargs)

    def _init_quantizers(self, weight_quantizer: WeightQuantizer = None, input_quantizer: TensorQuantizer = None,
                         output_quantizer: TensorQuantizer = None) -> None:
        self.weight_quantizer = weight_quantizer
        self.input_quantizer = input_quantizer
        self.output_quantizer = output_quantizer

    def _init_parameters(self, w_ih: Tensor, w_hh: Tensor, b_ih: Tensor, b_hh: Tensor) -> None:
        self.weight_ih, self.weight_hh, self.bias_ih, self.bias_hh = w_ih, w_hh, b_ih, b_hh

    def _layer_weight_amax(self):
        # one amax per layer over cat[w_ih, w_hh] (quant_lstm.py:152-155)
        return torch.max(torch.cat([self.weight_ih, self.weight_hh], 1).abs())

    def _quant_parameters(self, weights, rb_scale_list) -> None:
        self.weight_quantizer.amax = self._layer_weight_amax()
        self.weight_ih = Parameter(self.weight_quantizer(self.weight_ih), requires_grad=False)
        self.weight_hh = Parameter(self.weight_quantizer(self.weight_hh), requires_grad=False)

    def forward(self, x: Tensor, hx: Tensor, cx: Tensor, quant_y: bool = False) -> Tuple[Tensor, Tensor, Tensor]:
        steps = []
        quantize_inputs = hasattr(self, "input_quantizer") and self.input_quantizer is not None \
            and self.input_quantizer.mode is not None
        for t in range(x.shape[0]):
            xt = x[t]
            if quantize_inputs:
                # calibrates / fake-quantises x_t and h together with ONE scale (quant_lstm.py:166-171)
                joined = self.input_quantizer(torch.cat([xt, hx], 1))
                xt, hx = joined.split([xt.size(1), hx.size(1)], 1)
            gates = F.linear(xt, self.weight_ih, self.bias_ih) + F.linear(hx, self.weight_hh, self.bias_hh)
            i_g, f_g, g_g, o_g = gates.chunk(4, 1)
            cx = torch.sigmoid(f_g) * cx + torch.sigmoid(i_g) * torch.tanh(g_g)
            hx = torch.sigmoid(o_g) * torch.tanh(cx)
            steps.append(hx)
        return torch.stack(steps, 0), hx, cx


class iLSTMLayer(QuantLSTMLayer):
    def __init__(self, input_size: int, hidden_size: int, first_layer=False, **kwargs) -> None:
        super().__init__(input_size, hidden_size, **kwargs)
        self.first_layer = first_layer

    def _quant_parameters(self, weights, rb_scale) -> None:
        """quant_lstm.py:193-215: int8 weights with one per-layer scale, bias folded into the accumulator domain
        bias_hh = (b_ih + b_hh) * in_scale * w_scale, rb_scale = 1 / (in_scale * w_scale)."""
        self.weight_quantizer.amax = self._layer_weight_amax()
        self.weight_ih = Parameter(self.weight_quantizer._quant_forward(self.weight_ih, self.first_layer), requires_grad=False)
        self.weight_hh = Parameter(self.weight_quantizer._quant_forward(self.weight_hh, self.first_layer), requires_grad=False)
        b_scale = self.input_quantizer.scale * self.weight_quantizer.scale
        self.bias_hh = Parameter((self.bias_hh + self.bias_ih) * b_scale, requires_grad=False)
        self.rb_scale = 1 / b_scale.item()
        self.in_quant_scale = self.input_quantizer.scale.item()
        self.out_quant_scale = self.output_quantizer.scale.item()
        rb_scale[len(weights)] = self.rb_scale
        weights.append([self.weight_ih, self.weight_hh, self.bias_ih, self.bias_hh])

    def forward(self, x: Tensor, hx: Tensor, cx: Tensor, skip_quant_y: bool) -> Tuple[Tensor, Tensor, Tensor]:
        if self.first_layer and x.shape[-1] % 64 != 0:
            x = F.pad(x, (0, 64 - x.shape[-1] % 64), "constant", 0)
        return P.lstm_layer_amx_int8(x, hx, cx, self.weight_ih, self.weight_hh, self.bias_ih, self.bias_hh, self.rb_scale,
                                     self.in_quant_scale, self.out_quant_scale, skip_quant_y)
