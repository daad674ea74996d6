#!/bin/bash
# ncu evidence of one bench step, run on the GPU box AFTER the same bench command has exited 0 without ncu:
#   bash profiles/ncu_capture.sh          -> gpurun_out/launches.csv, gpurun_out/prof_{rec,gemm,dec,gemm_bf16}.ncu-rep
# Numbers printed by bench.py under ncu are never bench values.  Summaries: python profiles/ncu_summarize.py
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-microbench --no-parity"
mkdir -p gpurun_out
# launch list of the whole command (warm-up pass, timed pass, instrumented pass): per-launch durations, cold-cache and serialised
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
# one full capture per kernel, a launch of the 2048-utterance batch in the second pass (-s = launches of that kernel skipped)
cap() { ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c 1 -f -o gpurun_out/prof_$3 $B > gpurun_out/ncu_$3.log 2>&1; }
cap lstm_recurrent_kernel 10 rec            # first pre-RNN layer (K = 256 projection, T = 500)
cap gemm_i8_i32_pair_kernel 11 gemm         # second pre-RNN layer (K = 1024)
cap greedy_decode_kernel 2 dec
cap gemm_bf16_bf16_kernel 2 gemm_bf16
ls -la gpurun_out/*.ncu-rep
