#!/usr/bin/env bash
# knob sweep for the masked softmax kernels (bf16 [4,32,2048,2048]); prints ms per launch
for C in 1 2 4; do for KB in 64 128; do
  echo -n "mask C=$C inflight_kb=$KB "; PDB200_SOFTMAX_PIPE_C=$C PDB200_PIPE_INFLIGHT_KB=$KB timeout 60 python profiles/run_case.py softmax_mask_bf16 20 | tail -1
done; done
for C in 2 4; do for KB in 64 128; do
  echo -n "tri C=$C inflight_kb=$KB "; PDB200_SOFTMAX_PIPE_C=$C PDB200_PIPE_INFLIGHT_KB=$KB timeout 60 python profiles/run_case.py softmax_tri_bf16 20 | tail -1
done; done
for R in 2 16; do echo -n "tri C=4 rows_per_cta=$R "; PDB200_SOFTMAX_PIPE_C=4 PDB200_PIPE_ROWS_PER_CTA=$R timeout 60 python profiles/run_case.py softmax_tri_bf16 20 | tail -1; done
