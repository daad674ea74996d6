"""Layer-by-layer comparison of the int8 encoder (GPU ops vs CPU oracle) on real pipeline data."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from oracle import oracle as O, pipeline as OP
from rnnt_inference_b200 import synthetic as S, _lib, ops
from rnnt_inference_b200.modeling_rnnt import RNNT

torch.cuda.set_device(0)
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 7
lens = [int(v) for v in (sys.argv[2].split(",") if len(sys.argv) > 2 else "40,33,21,8".split(","))]
first_id = int(sys.argv[3]) if len(sys.argv) > 3 else 0
sd = S.make_checkpoint(seed=seed)
rnnt = RNNT(sd, "quant", True).eval()
params = OP.prepare_params(sd)
feats = [S.make_features(n, first_id + i) for i, n in enumerate(lens)]
x, xl = OP.pad_batch(feats, pad_to=32)
n = x.shape[1]
tr = rnnt.transcription
stacks = [(tr.pre_rnn, 0), (tr.pre_rnn, 1), (tr.post_rnn, 0), (tr.post_rnn, 1), (tr.post_rnn, 2)]
enc = params["enc"]
cur_o = O.quantize_pad(x, enc[0]["in_scale"], enc[0]["k_pad"])
cur_g = ops.quantize_pad(torch.from_numpy(x).cuda(), float(tr.pre_rnn.in_scale[0]), 256)
print("seed", seed, "lens", lens, "| quantised input equal:", np.array_equal(cur_g.cpu().numpy(), cur_o))
for i, ((rnn, l), lp) in enumerate(zip(stacks, enc)):
    if i == 2:
        cur_o = O.stack_time(cur_o, xl.astype(np.int32), 2)
        cur_g = ops.stack_time(cur_g, torch.from_numpy(xl.astype(np.int32)).cuda(), 2)
        print("stack_time equal:", np.array_equal(cur_g.cpu().numpy(), cur_o))
    w_ih, w_hh, _b, bias = rnn.weights[l]
    last = i == 4
    print(f"layer {i}: weights equal {np.array_equal(w_ih.cpu().numpy(), lp['w_ih'])} {np.array_equal(w_hh.cpu().numpy(), lp['w_hh'])} "
          f"bias equal {np.array_equal(bias.cpu().numpy().view(np.uint32), lp['bias'].view(np.uint32))} scales "
          f"{float(rnn.rb_scale[l]) == lp['rb_scale']} {float(rnn.in_scale[l]) == lp['in_scale']} {float(rnn.out_scale[l])} vs {lp['out_scale']}")
    hx_o, cx_o = np.zeros((n, 1024), np.int8), np.zeros((n, 1024), np.uint16)
    y_o, g_o = O.lstm_layer_i8(cur_o, hx_o, cx_o, lp["w_ih"], lp["w_hh"], lp["bias"], lp["rb_scale"], lp["in_scale"], lp["out_scale"], last, dump_gates=True)
    hx = torch.zeros((n, 1024), dtype=torch.int8, device="cuda")
    cx = torch.zeros((n, 1024), dtype=torch.float16, device="cuda")
    # feed the GPU layer with the ORACLE's input so that every layer is judged on its own
    xin = torch.from_numpy(cur_o).cuda()
    y = ops.lstm_layer_i8(xin, hx, cx, w_ih, w_hh, bias, float(rnn.rb_scale[l]), float(rnn.in_scale[l]),
                          float(rnn.out_scale[l]) if not last else lp["out_scale"], last)
    _lib.check_device_status()
    y_g = y.view(torch.int16).cpu().numpy().view(np.uint16) if last else y.cpu().numpy()
    diff = np.argwhere(y_g != y_o)
    print(f"   y differing {len(diff)} of {y_o.size}; hx equal {np.array_equal(hx.cpu().numpy(), hx_o)}; cx equal {np.array_equal(cx.cpu().numpy().view(np.uint16), cx_o)}")
    if len(diff):
        t0, n0, j0 = diff[0]
        print(f"   first at (t={t0}, n={n0}, j={j0}): gpu {y_g[t0, n0, j0]} oracle {y_o[t0, n0, j0]}; oracle gates i,f,g,o = "
              f"{[float(g_o[t0, n0, q * 1024 + j0]) for q in range(4)]}")
        print("   t histogram of differences:", np.bincount(diff[:, 0], minlength=y_o.shape[0])[:20].tolist())
        # does the GPU's own projection GEMM match?
        acc = ops.gemm_i8_i32(xin[t0], w_ih).cpu().numpy()
        print("   projection of that timestep equal:", np.array_equal(acc, O.gemm_i8_i32(cur_o[t0], lp["w_ih"])))
    cur_g = y
    cur_o = y_o
