"""RNN-T model modules (reference: models/modeling_rnnt.py).

RNNT = Transcription (2xLSTM 240->1024, StackTime x2, 3xLSTM 2048->1024) + Prediction (Embedding 28x320,
2xLSTM 320) + Joint (1344->512, ReLU, 512->29) + GreedyDecoderUpdate.  Constructor and forward signatures follow
the reference so that decoder.py-style callers work unchanged; in run_mode "quant" (+ enable_bf16) every forward
dispatches to librnnt_b200.so on the module's CUDA device.
"""
import os
from typing import List, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor
from torch.nn import Linear
from torch.nn.parameter import Parameter

from . import _C as P
from .config import RNNTParam
from .quant_lstm import QuantLSTM, iLSTM
from .quant_modules import transpose_tile_weight_bf16
from .utils import migrate_state_dict


class RNNT(torch.nn.Module):
    def __init__(self, model_path=None, run_mode=None, enable_bf16=False, load_jit=False, device=None):
        """model_path: path of an MLPerf-format checkpoint (or the already loaded dict).  device: where the model
        runs; defaults to the current CUDA device for run_mode "quant" (and for the fp32 modes with enable_bf16, whose
        prediction / joint operators are GPU kernels) and to the CPU for the plain fp32 modes."""
        super().__init__()
        if run_mode == "quant" and not enable_bf16:
            raise RuntimeError("run_mode='quant' runs the prediction/joint networks in bf16: pass enable_bf16=True "
                               "(the configuration of run.sh / launch_sut.sh)")
        self.transcription = Transcription(run_mode)
        self.prediction = Prediction(enable_bf16)
        self.joint = Joint(enable_bf16)
        self.update = GreedyDecoderUpdate()
        self.run_mode = run_mode
        self.enable_bf16 = enable_bf16
        self._decode_weights = None
        if model_path is not None:
            calibrated = run_mode not in (None, "f32", "calib")
            self._load_model(model_path, run_mode, enable_bf16, load_jit, calibrated)
            if device is None:
                # the int8 encoder and the bf16 prediction / joint operators exist on the GPU only
                on_gpu = run_mode == "quant" or (enable_bf16 and torch.cuda.is_available())
                device = torch.device("cuda", torch.cuda.current_device()) if on_gpu else torch.device("cpu")
            self.to(device)

    def _load_model(self, model_path, run_mode=None, enable_bf16=False, load_jit=False, saved_quantizers=False):
        """modeling_rnnt.py:34-81.  With saved quantizers the per-layer cells (and their `_amax` buffers) must exist
        BEFORE load_state_dict so that the calibrated ranges load; otherwise they are built from the loaded fp32
        LSTM weights afterwards."""
        if load_jit:
            raise RuntimeError("TorchScript archives of the reference embed torch.ops.intel_mlperf CPU kernels and cannot "
                               "run here; load the state-dict checkpoint (rnnt.pt / rnnt_calib.pt) instead")
        model = torch.load(model_path, map_location="cpu") if isinstance(model_path, (str, os.PathLike)) else model_path
        state_dict = migrate_state_dict(model)
        trans = self.transcription
        if saved_quantizers:
            trans.pre_rnn._init_layers(run_mode)
            trans.post_rnn._init_layers(run_mode)
            self.load_state_dict(state_dict, strict=False)
        else:
            self.load_state_dict(state_dict, strict=False)
            trans.pre_rnn._init_layers(run_mode)
            trans.post_rnn._init_layers(run_mode)
        trans.pre_rnn.lstm0.first_layer = True
        if run_mode == "quant":
            # chain the activation scales before folding them into the weights: out(pre.lstm1) = in(post.lstm0)
            trans.pre_rnn.lstm1.output_quantizer = trans.post_rnn.lstm0.input_quantizer
            for rnn in (trans.pre_rnn, trans.post_rnn):
                for layer in range(rnn.num_layers - 1):
                    getattr(rnn, f"lstm{layer}").output_quantizer = getattr(rnn, f"lstm{layer + 1}").input_quantizer
        trans.pre_rnn._process_parameters(run_mode)
        trans.post_rnn._process_parameters(run_mode)
        if run_mode == "quant":
            trans.pre_rnn._propagate_quantizers()
            trans.post_rnn._propagate_quantizers()
            trans.pre_quantizer = trans.pre_rnn.lstm0.input_quantizer
            trans.post_quantizer = trans.post_rnn.lstm0.input_quantizer
        self.prediction.prepack_weights()
        if enable_bf16:
            self.joint.prepack_weights()

    def decode_weights(self):
        """Device pointers of the prediction/joint weights for the fused decode kernel (cached)."""
        if self._decode_weights is None:
            pr, jn = self.prediction, self.joint
            w = pr.pred_rnn.weights
            self._decode_weights = P.make_decode_weights(
                pr.embed.weight.data, [w[0][0], w[1][0]], [w[0][1], w[1][1]], [w[0][3], w[1][3]], jn.linear1_pred.weight.data,
                jn.linear1_bias.data, jn.linear2.weight.data, jn.linear2.bias.data)
        return self._decode_weights

    def _apply(self, fn, recurse=True):
        self._decode_weights = None
        return super()._apply(fn, recurse)


class Transcription(torch.nn.Module):
    def __init__(self, run_mode, **kwargs):
        super().__init__()
        rnn_cls = iLSTM if run_mode == "quant" else QuantLSTM
        hidden = RNNTParam.trans_hidden_size
        self.pre_rnn = rnn_cls(RNNTParam.trans_input_size, hidden, RNNTParam.pre_num_layers, skip_quant_y=False)
        self.post_rnn = rnn_cls(hidden * RNNTParam.stack_time_factor, hidden, RNNTParam.post_num_layers, skip_quant_y=True)
        self.stack_time = StackTime(RNNTParam.stack_time_factor, run_mode)
        self.run_mode = run_mode

    def forward(self, x: Tensor, x_lens: Tensor, pre_hx: List[Tensor], pre_cx: List[Tensor], post_hx: List[Tensor],
                post_cx: List[Tensor], lens_host=None
                ) -> Tuple[Tensor, List[Tensor], List[Tensor], List[Tensor], List[Tensor]]:
        """x {T, N, 240} fp32 (or int8 already quantised), x_lens {N} -> x {ceil(T/2), N, 1024} (bf16 in quant mode);
        the four state lists are updated in place and returned (modeling_rnnt.py:116-144)."""
        if lens_host is not None and self.run_mode == "quant":
            # ragged batch sorted by length: per-128-row-tile step counts for the recurrent kernels (pre: frames,
            # post: stacked frames); the skipped tail is exactly what stack_time / the decode loop never read
            from . import ops
            n = x.size(1)
            fac = RNNTParam.stack_time_factor
            frames = int(sum(lens_host))                                   # true (frame, utterance) pairs: telemetry only
            frames2 = int(sum((int(v) + fac - 1) // fac for v in lens_host))
            x, pre_hx, pre_cx = self.pre_rnn(x, pre_hx, pre_cx, tile_steps=ops.tile_steps_for(lens_host, n), true_rows=frames)
            x = self.stack_time(x, x_lens)
            x, post_hx, post_cx = self.post_rnn(x, post_hx, post_cx, tile_steps=ops.tile_steps_for(lens_host, n, fac),
                                                true_rows=frames2)
            return x, pre_hx, pre_cx, post_hx, post_cx
        x, pre_hx, pre_cx = self.pre_rnn(x, pre_hx, pre_cx)
        x = self.stack_time(x, x_lens)
        x, post_hx, post_cx = self.post_rnn(x, post_hx, post_cx)
        return x, pre_hx, pre_cx, post_hx, post_cx


class Prediction(torch.nn.Module):
    def __init__(self, enable_bf16, **kwargs):
        super().__init__()
        self.sos = RNNTParam.SOS
        hidden = RNNTParam.pred_hidden_size
        self.embed = torch.nn.Embedding(RNNTParam.num_labels - 1, hidden)
        self.pred_rnn = torch.nn.LSTM(hidden, hidden, RNNTParam.pred_num_layers)
        self.enable_bf16 = enable_bf16

    def prepack_weights(self):
        """modeling_rnnt.py:161-181: bf16 weights + summed bias [w_ih, w_hh, b_ih, b_ih + b_hh] per layer."""
        packed = []
        for layer in range(self.pred_rnn.num_layers):
            w_ih, w_hh, b_ih, b_hh = [t.detach() for t in self.pred_rnn.all_weights[layer][:4]]
            if self.enable_bf16:
                packed.append([transpose_tile_weight_bf16(w_ih.to(torch.bfloat16).transpose(1, 0)),
                               transpose_tile_weight_bf16(w_hh.to(torch.bfloat16).transpose(1, 0)), b_ih, b_hh + b_ih])
            else:
                packed.append([*P.prepack_lstm_weights(w_ih, w_hh), b_ih, b_hh])
        self.pred_rnn.weights = packed
        if self.enable_bf16:
            self.embed.weight = Parameter(self.embed.weight.to(torch.bfloat16), requires_grad=False)

    def _apply(self, fn, recurse=True):
        out = super()._apply(fn, recurse)
        if hasattr(self.pred_rnn, "weights"):
            self.pred_rnn.weights = [[fn(t) for t in layer] for layer in self.pred_rnn.weights]
        return out

    def forward(self, pre_g: Tensor, pre_hg: List[Tensor], pre_cg: List[Tensor]) -> Tuple[Tensor, List[Tensor], List[Tensor]]:
        """pre_g {U=1, N} token ids (SOS = -1 -> zero embedding); returns g {1, N, 320} and NEW hg/cg lists."""
        sos_mask = pre_g.eq(self.sos)
        g = self.embed(pre_g.masked_fill(sos_mask, 0).long())
        g = g.masked_fill_(sos_mask.unsqueeze(2), 0.0)
        if self.enable_bf16:
            return P.lstm_amx_bf16(g, pre_hg, pre_cg, self.pred_rnn.weights)
        return P.lstm(g, pre_hg, pre_cg, self.pred_rnn.weights)


class Joint(torch.nn.Module):
    def __init__(self, enable_bf16, **kwargs):
        super().__init__()
        self.linear1_trans = Linear(RNNTParam.trans_hidden_size, RNNTParam.joint_hidden_size)
        self.linear1_pred = torch.nn.Linear(RNNTParam.pred_hidden_size, RNNTParam.joint_hidden_size)
        self.relu = torch.nn.ReLU()
        self.linear2 = torch.nn.Linear(RNNTParam.joint_hidden_size, RNNTParam.num_labels)
        self.enable_bf16 = enable_bf16

    def prepack_weights(self):
        """modeling_rnnt.py:223-257: one summed fc1 bias, bf16 weights, fc2 padded 29 -> 32 rows with zero weights
        and zero bias (so the argmax of decoder.py:193 runs over 32 columns, 29..31 identically 0)."""
        if not self.enable_bf16:
            return
        self.linear1_bias = Parameter(self.linear1_trans.bias + self.linear1_pred.bias, requires_grad=False)
        for lin, pad in ((self.linear1_trans, False), (self.linear1_pred, False), (self.linear2, True)):
            lin.weight = Parameter(transpose_tile_weight_bf16(lin.weight.to(torch.bfloat16).transpose(1, 0), pad),
                                   requires_grad=False)
        b2 = self.linear2.bias
        if b2.shape[-1] % 32 != 0:
            b2 = F.pad(b2, (0, 32 - b2.shape[-1] % 32), "constant", 0)
        self.linear2.bias = Parameter(b2, requires_grad=False)

    def hoist_trans(self, f: Tensor, tile_steps=None, exact_order=False) -> Tensor:
        """y1 = bf16(f . W_trans^T) for every encoder frame at once.  The reference rounds this product to bf16 in
        memory each iteration (amx_linear.cpp:137-150), so it depends on the frame only and is computed once — on the
        tensor cores (tcgen05 bf16 GEMM).  exact_order=True uses the CUDA-core operator whose fp32 accumulation order is
        the CPU oracle's (bit-exact cross-checks); the two agree to <= 1 bf16 ulp."""
        if exact_order:
            return P.linear_bf16_bf16(f, self.linear1_trans.weight, tile_steps)
        return P.gemm_bf16_bf16(f, self.linear1_trans.weight, tile_steps)

    def forward(self, f: Tensor, g: Tensor, padded_batch_size: int = 1) -> Tensor:
        """f {N, 1024}, g {N, 320} -> logits {N, 32} fp32 (bf16 mode) or {N, 29} (fp32 mode)."""
        if self.enable_bf16:
            y = P.amx_linear_bf16_accum_relu(f, self.linear1_trans.weight, g, self.linear1_pred.weight, self.linear1_bias)
            return P.amx_linear_i16o32(y, self.linear2.weight, self.linear2.bias)
        y = self.linear1_trans(f)
        y = y + self.linear1_pred(g)
        return self.linear2(self.relu(y))


class StackTime(torch.nn.Module):
    def __init__(self, factor, run_mode="f32"):
        super().__init__()
        self.factor = factor
        self.run_mode = run_mode

    def forward(self, x: Tensor, x_lens: Tensor) -> Tensor:
        """{T, N, C} -> {ceil(T/factor), N, C*factor}; frames at or beyond x_lens[n] are zeroed."""
        if self.run_mode == "quant":
            return self.forward_quant(x, x_lens)
        return self.forward_f32(x, x_lens)

    def forward_f32(self, x: Tensor, x_lens: Tensor) -> Tensor:
        t, n, c = x.shape
        keep = torch.arange(t, device=x.device).unsqueeze(1) < x_lens.to(device=x.device, dtype=torch.int64).unsqueeze(0)
        x = x * keep.unsqueeze(2).to(x.dtype)
        pad_t = (-t) % self.factor
        if pad_t:
            x = torch.cat([x, x.new_zeros(pad_t, n, c)], 0)
        # [T', N, C] -> [T'/f, f, N, C] -> [T'/f, N, f*C]
        return x.view(-1, self.factor, n, c).permute(0, 2, 1, 3).reshape(-1, n, c * self.factor).contiguous()

    def forward_quant(self, x: Tensor, x_lens: Tensor) -> Tensor:
        return P.stack_time(x, x_lens, self.factor)


class GreedyDecoderUpdate(torch.nn.Module):
    def forward(self, symbols: Tensor, symbols_added: Tensor, res: Tensor, res_idx: Tensor, f: Tensor, f_lens: Tensor,
                time_idx: Tensor, fi: Tensor, pre_g: Tensor, pre_hg: List[Tensor], pre_cg: List[Tensor], hg: List[Tensor],
                cg: List[Tensor]) -> bool:
        return P.greedy_decode_update(symbols, symbols_added, res, res_idx, f, f_lens, time_idx, fi, pre_g, pre_hg, pre_cg,
                                      hg, cg)
