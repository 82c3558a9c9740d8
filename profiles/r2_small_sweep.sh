#!/usr/bin/env bash
# rms_norm geometry and the warp-vs-block row reduction threshold (ms per launch)
for C in 2 4; do for KB in 64 96 128; do
  echo -n "rms C=$C inflight_kb=$KB "; PDB200_RMS_PIPE_C=$C PDB200_PIPE_INFLIGHT_KB=$KB timeout 60 python profiles/run_rms.py 20 | tail -1
done; done
for c in argmax_fp32 sum_rows_fp32 max_rows_bf16; do for B in 16384 32768 65536; do
  echo -n "$c warp_max=$B "; PDB200_ROWS_WARP_MAX_BYTES=$B timeout 60 python profiles/run_case.py $c 20 | tail -1
done; done
