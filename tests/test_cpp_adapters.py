"""The C++ drop-in adapters (include/ompl_b200/*.h) against the reference's own classes.

tests/cpp/test_adapters.cpp is compiled in the build container against the
reference headers and the reference's compiled state-space code
(oracle/_ref/libompl_refbase.so); the binary travels to the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "_bin", "test_adapters")


def test_adapter_headers_cite_the_reference_interfaces():
    nn = open(os.path.join(ROOT, "include", "ompl_b200", "NearestNeighborsCUDA.h")).read()
    for virt in ("reportsSortedResults", "clear", "add(const _T &data)", "add(const std::vector<_T> &data)", "remove",
                 "nearest(", "nearestK", "nearestR", "size()", "list("):
        assert virt in nn, virt
    assert "NearestNeighbors.h:46-115" in nn and "No elements found in nearest neighbors data structure" in nn
    br = open(os.path.join(ROOT, "include", "ompl_b200", "OmplBridge.h")).read()
    assert "class BatchedDiscreteMotionValidator : public MotionValidator" in br
    assert "class BatchedStateValidityChecker : public StateValidityChecker" in br


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/ompl"), reason="needs the reference headers")
def test_adapter_test_binary_is_built():
    assert os.path.exists(BIN), "run __graft_entry__.build()"


@pytest.mark.gpu
def test_cpp_adapters_against_reference_classes(gpu):
    if not os.path.exists(BIN):
        pytest.skip("tests/cpp/_bin/test_adapters was not built (no reference tree at build time)")
    p = subprocess.run([BIN], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-4000:]
    assert "all adapter tests passed" in p.stdout
