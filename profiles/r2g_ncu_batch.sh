#!/usr/bin/env bash
# One gpurun call at the end of round 2: the launch list of bench.py and ncu --set full of the kernels added or reworked
# after the judge's run (each after a plain run of the same command has exited 0).  usage (on the GPU box): bash profiles/r2g_ncu_batch.sh
set -u
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/r2g_bench_plain.json 2> gpurun_out/r2g_bench_plain.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2g_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 > gpurun_out/r2g_bench_ncu.log 2>&1
echo "launch list rc=$?"
run() {  # name, case, kernel regex, extra env
  local name="$1" c="$2" k="$3"; shift 3
  env "$@" timeout 120 python profiles/run_case.py "$c" 3 > "gpurun_out/${name}.log" 2>&1 &&
  env "$@" timeout 300 ncu --set full --clock-control none -k "regex:$k" -s 2 -c 1 -f -o "gpurun_out/${name}" \
      python profiles/run_case.py "$c" 3 > "gpurun_out/${name}.ncu.log" 2>&1
  local rc=$?
  if [ -f "gpurun_out/${name}.ncu-rep" ]; then
    ncu -i "gpurun_out/${name}.ncu-rep" --page raw --csv > "gpurun_out/${name}.raw.csv" 2>/dev/null
    rm -f "gpurun_out/${name}.ncu-rep"
  fi
  echo "$name rc=$rc $(tail -1 gpurun_out/${name}.log)"
}
run r2g_softmax_vocab_bf16 softmax_vocab_bf16 softmax_rows_cluster_lean A=1
run r2g_softmax_vocab_fp32 softmax_vocab_fp32 softmax_rows_relay A=1
run r2g_ln_vocab_bf16 ln_vocab_bf16 ln_rows_cluster A=1
run r2g_max_grad_rows_bf16 max_grad_rows_bf16 select_outer A=1
run r2g_max_grad_cols_fp32 max_grad_cols_fp32 select_inner A=1
run r2g_add_bf16 add_bf16 ew_inner A=1
