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


BIN2 = os.path.join(ROOT, "tests", "cpp", "_bin", "test_adapters2")
BENCH = os.path.join(ROOT, "tests", "cpp", "_bin", "bench_adapter")


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/ompl"), reason="needs the reference headers")
def test_adapter_test_binaries_are_built():
    for b in (BIN, BIN2, BENCH):
        assert os.path.exists(b), f"{b} missing: run __graft_entry__.build()"


def _run(binary, marker, *args):
    # The binaries need the reference headers to compile, so they are built where /root/reference exists
    # and travel to the GPU box with the snapshot.  A GPU box without them is a broken snapshot, not a
    # reason to skip: fail.
    assert os.path.exists(binary), f"{binary} did not travel to the GPU box (build it with __graft_entry__.build())"
    p = subprocess.run([binary, *args], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-4000:]
    assert marker in p.stdout, p.stdout[-2000:]
    return p.stdout


@pytest.mark.gpu
def test_cpp_adapters_against_reference_classes(gpu):
    _run(BIN, "all adapter tests passed")


@pytest.mark.gpu
def test_cpp_adapters_element_types_subclasses_sharding_threads(gpu):
    out = _run(BIN2, "all adapter tests (part 2) passed",
               os.path.join(ROOT, "tests", "golden", "benchmark_log_sample.log"))
    for part in ("unsigned long (PRM Vertex)", "Motion* (FMT)", "std::shared_ptr<Vertex> (BIT*)",
                 "subclass / explicit description", "sharded structure", "getMotionStates", "concurrent queries", "default structure selector"):
        assert part in out, part


@pytest.mark.gpu
def test_cpp_adapter_timing_binary_runs(gpu):
    # 20000 queries: the adapter's host work runs on several threads; the binary also checks nearestKBatchFlat == nearestKBatch
    out = _run(BENCH, "queries_per_s", "50000", "20000", "10", "2")
    assert '"neighbours": 200000' in out and "flat_queries_per_s" in out
