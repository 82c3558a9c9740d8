#!/usr/bin/env bash
# One gpurun call: ncu --set full on one launch of each kernel named in VERDICT r1 item 7 (each after a plain run of the
# same command has exited 0).  usage (on the GPU box): bash profiles/r2_ncu_batch.sh
set -u
run() {  # name, case, kernel regex, extra env
  local name="$1" c="$2" k="$3"; shift 3
  env "$@" timeout 120 python profiles/run_case.py "$c" 3 > "gpurun_out/${name}.log" 2>&1 &&
  env "$@" timeout 300 ncu --set full --clock-control none -k "regex:$k" -s 2 -c 1 -f -o "gpurun_out/${name}" \
      python profiles/run_case.py "$c" 3 > "gpurun_out/${name}.ncu.log" 2>&1
  local rc=$?
  # the reports are tens of MB each and gpurun_out/ travels back only below 64 MiB: keep the raw metric page as CSV
  if [ -f "gpurun_out/${name}.ncu-rep" ]; then
    ncu -i "gpurun_out/${name}.ncu-rep" --page raw --csv > "gpurun_out/${name}.raw.csv" 2>/dev/null
    rm -f "gpurun_out/${name}.ncu-rep"
  fi
  echo "$name rc=$rc $(cat gpurun_out/${name}.log | tail -1)"
}
run r2_add_bf16 add_bf16 ew_inner A=1
run r2_sum_full_fp32 sum_full_fp32 reduce_rows_block A=1
run r2_sum_full_fp32_nointerleave sum_full_fp32 reduce_rows_block PDB200_NO_INTERLEAVE=1
run r2_softmax_bf16 softmax_bf16 softmax_pipe A=1
run r2_softmax_grad_bf16 softmax_grad_bf16 softmax_grad_pipe A=1
run r2_mulgrad_bf16 mulgrad_bf16 mulgrad_inner_tma A=1
run r2_ew_general_bf16 ew_general_bf16 ew_general A=1
run r2_ew_outer_bf16 ew_outer_bf16 ew_outer A=1
run r2_softmax_cols_bf16 softmax_cols_bf16 softmax_cols_tile A=1
run r2_softmax_vocab_bf16 softmax_vocab_bf16 softmax_rows_cluster A=1
run r2_ln_vocab_bf16 ln_vocab_bf16 ln_rows_cluster A=1
run r2_softmax_mask_bf16 softmax_mask_bf16 softmax_mask_pipe A=1
run r2_softmax_tri_bf16 softmax_tri_bf16 softmax_mask_pipe A=1
run r2_ln_grad_bf16 ln_grad_bf16 ln_grad_pipe A=1
