#!/bin/bash
mkdir -p gpurun_out
echo "== alone, fresh process"; python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/na_alone.log 2>&1; echo "rc $?"; tail -15 gpurun_out/na_alone.log
echo "== after the other tests (pytest -x, whole suite)"; python -m pytest tests -m gpu -x -q > gpurun_out/na_suite.log 2>&1; echo "rc $?"; tail -4 gpurun_out/na_suite.log
python -m pytest tests -m gpu -q --collect-only 2>/dev/null | tail -4
