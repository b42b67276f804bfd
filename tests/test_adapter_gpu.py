"""GPU: the reference-side C++ drop-in check (tests/cpp/adapter_test.cpp) — ggq::GpuBRWT / ggq::GpuColMajor
(include/ggq/gpu_matrix.hpp), subclasses of the reference's BinaryMatrix, against the unmodified BRWT /
ColMajorCompressed through BinaryMatrix::load(std::istream&) / get / get_row / get_column / serialize.
The binary is prebuilt by oracle/Makefile (needs /root/reference headers) and travels in oracle/_ref/."""
import os
import subprocess

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_cpp_adapters_match_reference():
    exe = os.path.join(ROOT, "oracle", "_ref", "adapter_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/adapter_test not built (needs /root/reference at build time)")
    p = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "all checks passed" in p.stdout
