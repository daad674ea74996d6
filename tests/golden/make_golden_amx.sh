#!/bin/bash
# Regenerates tests/golden/amx_bf16_hw.npz on a CPU with AMX-BF16 + AVX512-BF16 (this container's Xeon has both).
# Test infrastructure; the vectors come from real silicon (TDPBF16PS, VRCP14PS), not from the oracle.
set -e
cd "$(dirname "$0")"
TMP=$(mktemp -d)
/usr/bin/g++ -O1 -ffp-contract=off -mamx-tile -mamx-bf16 -mavx512f -mavx512bw -mavx512vl -mavx512bf16 -mavx512dq \
    ../../oracle/hwprobe/gen_amx_bf16_golden.cpp -o "$TMP/gen"
"$TMP/gen" "$TMP"
python - "$TMP" <<'PY'
import sys, numpy as np
d = sys.argv[1] + "/"
ld = lambda n, dt: np.fromfile(d + n + ".bin", dtype=dt)
N, P, J, V = 32, 320, 512, 32
np.savez_compressed("amx_bf16_hw.npz",
    lstm_g_ih=ld("lstm_g_ih_f32", np.float32).reshape(N, 4 * P), lstm_gates=ld("lstm_gates_f32", np.float32).reshape(N, 4 * P),
    lstm_hy=ld("lstm_hy_bf16", np.uint16).reshape(N, P), lstm_cy=ld("lstm_cy_f32", np.float32).reshape(N, P),
    fc1_acc_trans=ld("fc1_acc_trans_f32", np.float32).reshape(N, J), fc1_y1=ld("fc1_y1_bf16", np.uint16).reshape(N, J),
    fc1_y=ld("fc1_y_bf16", np.uint16).reshape(N, J), fc2_logits=ld("fc2_logits_f32", np.float32).reshape(N, V))
PY
rm -rf "$TMP"
ls -la amx_bf16_hw.npz
