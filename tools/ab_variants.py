#!/usr/bin/env python
"""A/B harness for kernel experiments (used for the v8..v12 measurements in DESIGN.md §4).

Builds one libsr_cuda variant per `name:FLAGS` argument into scratch/ (git-ignored, but it travels with gpurun) and
prints the gpurun command line that benches every variant back to back on one B200 through `SR_LIB=...` and prints
frames/s plus the per-kernel split of each.  The default build is always included as `base`.

  python tools/ab_variants.py m6:-DSR_TILE_MINB=6 t128:-DSR_SETUP_THREADS=128 -- --workload soup1m
  python tools/ab_variants.py --run ...        # also calls gpurun (about 40 s of box time per variant and workload)
"""
import os
import shlex
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "software_renderer_b200", "csrc")
FLAGS = ("-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true "
         "-Xcompiler -fPIC -Xptxas -v")            # keep in step with csrc/Makefile


def main():
    args = sys.argv[1:]
    run = "--run" in args
    args = [a for a in args if a != "--run"]
    bench_args = ""
    if "--" in args:
        k = args.index("--")
        bench_args = " ".join(shlex.quote(a) for a in args[k + 1:])
        args = args[:k]
    os.makedirs(os.path.join(ROOT, "scratch"), exist_ok=True)
    names = ["base"]
    procs = []
    for spec in args:
        name, _, flags = spec.partition(":")
        out = os.path.join(ROOT, "scratch", f"libsr_cuda_{name}.so")
        cmd = f"nvcc {FLAGS} {flags} -shared -o {shlex.quote(out)} sr_api.cu"
        procs.append((name, subprocess.Popen(cmd, shell=True, cwd=CSRC, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        names.append(name)
    for name, p in procs:
        log = p.communicate()[0]
        if p.returncode != 0:
            sys.exit(f"build of variant {name} failed:\n{log}")
        use = [l.split(":", 1)[1].strip() for l in log.splitlines() if "Used" in l and "registers" in l]
        spill = [l.strip() for l in log.splitlines() if "spill stores" in l and not l.strip().startswith("0 bytes stack frame, 0 bytes spill")]
        print(f"{name}: built; {len(spill)} of {len(use)} kernels spill")
    subprocess.check_call(["make", "-C", CSRC], stdout=subprocess.DEVNULL)
    steps = []
    for name in names:
        env = "" if name == "base" else f"SR_LIB=scratch/libsr_cuda_{name}.so "
        steps.append(f"{env}python bench.py {bench_args} > gpurun_out/ab_{name}.json 2>/dev/null; "
                     f"python -c \"import json; d=json.load(open('gpurun_out/ab_{name}.json')); "
                     f"print('{name}', round(d['value']), d['frame_ms_rank0']['median'], d['kernel_ms'])\"")
    inner = "; ".join(steps)
    line = f"/usr/local/graft/bin/gpurun --timeout {60 + 45 * len(names)} -- {shlex.quote(inner)}"
    print(line)
    if run:
        subprocess.call(line, shell=True, cwd=ROOT)


if __name__ == "__main__":
    main()
