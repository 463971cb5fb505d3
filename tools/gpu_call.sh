#!/bin/bash
mkdir -p gpurun_out
for d in "" 1; do echo "== phases, depth=${d:-0}"; SR_PROBE_DEPTH=$d SR_LIB=scratch/libsr_cuda_tl.so python tools/micro/py_pdl_probe.py torch 2>&1 | grep -E "^view|k_setup" | tail -9; done
