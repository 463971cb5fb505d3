"""The C++ host mirror of the reference interface (include/sr_renderer.hpp) compiles and links against the C ABI;
on a GPU box it renders KAT-1 through `sr::CudaRenderer::render(const RenderPass&)`."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "software_renderer_b200")


@pytest.fixture(scope="module")
def demo(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("cpp") / "host_demo")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "host_demo.cpp"),
                    "-L" + LIBDIR, "-lsr_cuda", "-Wl,-rpath," + LIBDIR, "-o", out], check=True)
    return out


def test_cpp_host_links(demo):
    r = subprocess.run([demo, "--link-only"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "abi 1000"


def test_cpp_host_fails_loudly_without_gpu(demo):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([demo], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_cpp_host_renders_kat1(demo):
    r = subprocess.run([demo], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "covered=65537 ok=1" in r.stdout
