"""GPU: the reference-side C++ drop-in check (tests/cpp/adapter_test.cpp) — ggq::GpuBRWT / ggq::GpuColMajor
(include/ggq/gpu_matrix.hpp), subclasses of the reference's BinaryMatrix, against the unmodified BRWT /
ColMajorCompressed through BinaryMatrix::load(std::istream&) / get / get_row / get_column / serialize.
The binary is prebuilt by oracle/Makefile (needs /root/reference headers) and travels in oracle/_ref/."""
import os
import subprocess

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def layout_env(wide):
    """GGQ_WIDE_SLOTS=1: the process creates its matrices in the capacity layout — same answers."""
    env = dict(os.environ)
    env["GGQ_WIDE_SLOTS"] = "1" if wide else "0"
    return env


@pytest.mark.parametrize("wide", [False, True], ids=["default_layout", "capacity_layout"])
def test_cpp_adapters_match_reference(wide):
    exe = os.path.join(ROOT, "oracle", "_ref", "adapter_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/adapter_test not built (needs /root/reference at build time)")
    p = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=layout_env(wide))
    assert p.returncode == 0, p.stdout[-3000:]
    assert "all checks passed" in p.stdout


@pytest.mark.parametrize("wide", [False, True], ids=["default_layout", "capacity_layout"])
def test_reference_annotator_instantiated_over_the_gpu_matrix(wide):
    """annotate::StaticBinRelAnnotator<ggq::GpuBRWT, std::string> — the reference's own caller template over the GPU
    matrix (tests/cpp/annotator_test.cpp): loads the .brwt.annodbg the reference wrote, and answers the known-answer
    cases of the reference's tests/test_annotation_BRWT.cpp (get / has_labels / get_top_labels = Label2:4, Label8:3,
    Label1:2, Label0:1 / get_labels at every ratio) plus a random annotation against the reference annotator."""
    exe = os.path.join(ROOT, "oracle", "_ref", "annotator_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/annotator_test not built (needs /root/reference at build time)")
    p = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=layout_env(wide))
    assert p.returncode == 0, p.stdout[-3000:]
    assert "all checks passed" in p.stdout


def test_label_batcher_hands_every_caller_its_own_answer():
    """ggq::LabelBatcher (include/ggq/label_batcher.hpp): 8 threads x 200 query sequences through one batcher, every
    answer equal to the caller's own ggq_get_labels call (tests/cpp/batcher_test.cpp; also prints the throughput)."""
    exe = os.path.join(ROOT, "tests", "cpp", "bin", "batcher_test")
    if not os.path.exists(exe):
        pytest.skip("tests/cpp/bin/batcher_test not built (make -C genome_graph_annotation_b200/csrc)")
    p = subprocess.run([exe, "300000", "400", "0.01", "8", "120", "300"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "all checks passed" in p.stdout
