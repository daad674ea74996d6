#!/bin/bash
# Regenerates tests/golden/postop_fp16_hw.npz on a CPU with AVX512-FP16 (+AVX512-BF16).
# Test infrastructure; the vectors come from real silicon, not from the oracle.
set -e
cd "$(dirname "$0")"
TMP=$(mktemp -d)
/usr/bin/g++ -O1 -ffp-contract=off -mavx512fp16 -mavx512f -mavx512bw -mavx512vl -mavx512bf16 -mavx512dq \
    ../../oracle/hwprobe/gen_postop_golden.cpp -o "$TMP/gen"
"$TMP/gen" "$TMP"
python - "$TMP" <<'PY'
import sys, numpy as np
d = sys.argv[1] + "/"
ld = lambda n, dt: np.fromfile(d + n + ".bin", dtype=dt)
N, H, M = 48, 64, 16384
np.savez_compressed("postop_fp16_hw.npz",
    gates=ld("postop_gates_f32", np.float32).reshape(N, 4 * H), c_in=ld("postop_c_in_f16", np.uint16).reshape(N, H),
    scales=ld("postop_scales_f32", np.float32), c_out=ld("postop_c_out_f16", np.uint16).reshape(N, H),
    ybf=ld("postop_ybf_bf16", np.uint16).reshape(N, H), yq=ld("postop_yq_i8", np.int8).reshape(N, H),
    hq=ld("postop_hq_i8", np.int8).reshape(N, H), tanh_all=ld("tanh_all_f16", np.uint16),
    sigmoid_all=ld("sigmoid_all_f16", np.uint16), fma_a=ld("fma_a_f16", np.uint16)[:M], fma_b=ld("fma_b_f16", np.uint16)[:M],
    fma_c=ld("fma_c_f16", np.uint16)[:M], fma_out=ld("fma_out_f16", np.uint16)[:M])
PY
rm -rf "$TMP"
ls -la postop_fp16_hw.npz
