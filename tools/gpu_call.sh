#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "depth or fuzz or parity or golden or bands" > gpurun_out/c11_pytest.log 2>&1; tail -8 gpurun_out/c11_pytest.log
SR_ZPASS=1 python -m pytest tests -m gpu -x -q -k "depth or fuzz or parity or golden or bands" > gpurun_out/c11_pytest_z1.log 2>&1; tail -8 gpurun_out/c11_pytest_z1.log
for z in 0 1; do SR_ZPASS=$z python bench.py --steps 50 --warmup 20 --depth-test 1 --quick 1 > gpurun_out/c11_depth_z$z.json 2>/dev/null; done
python bench.py --steps 50 --warmup 20 --depth-test 1 --quick 1 > gpurun_out/c11_depth_auto.json 2>/dev/null
for f in 0.06 0.2; do SR_ZFRAC=$f python bench.py --steps 50 --warmup 20 --depth-test 1 --quick 1 > gpurun_out/c11_depth_f$f.json 2>/dev/null; done
python profiles/show.py gpurun_out/c11_depth_*.json
SR_LIB=scratch/libsr_cuda_tl.so python tools/micro/py_pdl_probe.py torch 2>&1 | grep "^view" | tail -1
