"""CPU checks of the BUILT library's machine code (cuobjdump on libsr_cuda.so; no GPU): the static facts DESIGN.md §5 and
profiles/r02_final_{sass_opcodes,resource_usage}.md state about the shipped kernels.  Skipped where cuobjdump is absent."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "software_renderer_b200", "libsr_cuda.so")
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"

pytestmark = pytest.mark.skipif(not os.path.exists(CUOBJDUMP), reason="cuobjdump not found")


def demangled(name):
    out = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    out = out.replace("(anonymous namespace)::", "").replace("void ", "").replace("sr::", "")
    return re.sub(r"\(.*\)$", "", out)


@pytest.fixture(scope="module")
def resources():
    text = subprocess.run([CUOBJDUMP, "--dump-resource-usage", LIB], capture_output=True, text=True, check=True).stdout
    res, name = {}, None
    for line in text.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            name = demangled(m.group(1))
        elif name and "REG:" in line:
            res[name] = {k: int(v) for k, v in re.findall(r"([A-Z]+):(\d+)", line)}
            name = None
    return res


@pytest.fixture(scope="module")
def opcodes():
    text = subprocess.run([CUOBJDUMP, "-sass", LIB], capture_output=True, text=True, check=True).stdout
    ops, name = {}, None
    for line in text.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = demangled(m.group(1))
            ops[name] = []
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            ops[name].append(m.group(1))
    return ops


def test_library_holds_sm_100a_code_only():
    text = subprocess.run([CUOBJDUMP, "--list-elf", LIB], capture_output=True, text=True, check=True).stdout
    archs = set(re.findall(r"sm_\d+a?", text))
    assert archs == {"sm_100a"}, archs


def test_every_frame_kernel_is_in_the_binary(resources):
    for k in ("k_transform", "k_late_boxes", "k_tile<false, false>", "k_tile<false, true>", "k_tile<true, false>", "k_tile<true, true>",
              "k_resolve<false, 1>", "k_resolve<false, 4>", "k_resolve<true, 1>", "k_resolve<true, 4>", "k_points_scatter",
              "k_points_resolve", "k_scan", "k_bin_fill", "k_clear_depth", "k_convert_vertices", "k_convert_indices",
              "k2d_rect", "k2d_circle", "k2d_line", "k2d_triangle", "k_na_clear", "k_na_scatter", "k_na_resolve"):
        assert k in resources, k
    assert sum(k.startswith("k_setup<") for k in resources) == 8


def test_occupancy_budgets_of_the_hot_kernels(resources):
    """k_tile: 40 registers = 6 resident 256-thread CTAs per SM (+1.9 % / +7 % when adopted); the binning
    k_setup: 48 = 5 CTAs; k_late_boxes must stay at full occupancy (<= 40 at 256 threads would be 6 CTAs; it uses no shared
    memory)."""
    for k, r in resources.items():
        if k.startswith("k_tile<"):
            assert r["REG"] <= 40, (k, r)
            assert r["SHARED"] <= 4096, (k, r)
        if k.startswith("k_setup<"):
            assert r["REG"] <= 48, (k, r)
            assert r["SHARED"] <= 40 * 1024, (k, r)           # what the shipped instances use at most (the box-word instance: 37,864 B)
    assert resources["k_late_boxes"]["REG"] <= 40 and resources["k_late_boxes"]["SHARED"] == 0
    assert resources["k_tile<false, false>"]["STACK"] == 0    # the first-batch tile kernel does not spill


def test_no_fp64_arithmetic_in_the_tile_and_resolve_kernels(opcodes):
    """the (triangle, row) path of k_tile carried DMUL / DFMA and F2I.F64; the 32-bit quotient with one
    remainder step (sr_math.cuh lerp_f32, proved in test_host_math.py) removed them."""
    fp64 = re.compile(r"^D(MUL|FMA|ADD|SETP|MNMX)")
    for k, o in opcodes.items():
        if k.startswith(("k_tile<", "k_resolve<", "k_late_boxes", "k_transform", "k_points", "k_na_")):
            bad = [x for x in o if fp64.match(x) or (re.match(r"(F2I|I2F|F2F)", x) and "F64" in x)]
            assert not bad, (k, bad[:5])
