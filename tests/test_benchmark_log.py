"""The tools::Benchmark log writer (ompl_b200/benchmark_log.py, include/ompl_b200/BenchmarkLog.h) against the
reference's own parser, scripts/ompl_benchmark_statistics.py (run where /root/reference exists), and against a
committed golden log whose text follows tools/benchmark/src/Benchmark.cpp:226-352 line by line."""
import io
import os
import sqlite3
import subprocess
import sys

import pytest

import helpers
from ompl_b200 import benchmark_log as bl

REF_PARSER = "/root/reference/scripts/ompl_benchmark_statistics.py"


def _sample_log():
    log = bl.BenchmarkLog("kNN_on_B200", seed=7, max_time=5.0, max_mem=4096.0, run_count=3)
    log.host, log.start_time = "b200-box", "2026-01-01 00:00:00"
    log.parameters = {"n_states INTEGER": "1000000", "k INTEGER": "10"}
    log.setup_info = "RealVectorStateSpace(7)\n"
    log.cpu_info = "16 host threads\n"
    log.total_duration = 1.5
    a = log.add_planner("geometric_PRM_NearestNeighborsCUDA", {"nearest_neighbors STRING": "NearestNeighborsCUDA"})
    for i in range(3):
        a.add_run(time=("REAL", 0.25 + i), solved=("BOOLEAN", i != 1), status=("ENUM", 6 if i != 1 else 4),
                  graph_states=("INTEGER", 1000 + i))
    b = log.add_planner("geometric_PRM_GNAT")
    b.add_run(time=("REAL", 2.5), solved=("BOOLEAN", True), status=("ENUM", 6))
    b.runs.append({"time REAL": "3.5", "memory REAL": "12.25"})  # ragged: a property missing in other runs
    b.progress_property_names = ["time REAL", "best cost REAL"]
    b.runs_progress_data = [[{"time REAL": "0.1", "best cost REAL": "9.0"}, {"time REAL": "0.2", "best cost REAL": "8.0"}],
                            [{"time REAL": "0.1", "best cost REAL": "7.5"}]]
    return log


def test_log_text_matches_the_golden_layout():
    out = io.StringIO()
    assert _sample_log().save(out)
    golden = open(os.path.join(helpers.GOLDEN, "benchmark_log_sample.log")).read()
    assert out.getvalue() == golden
    assert not bl.BenchmarkLog("empty").save(io.StringIO())  # "There is no experimental data to save"


@pytest.mark.skipif(not os.path.exists(REF_PARSER), reason="needs the reference's scripts/")
def test_reference_statistics_script_parses_the_log(tmp_path):
    logf, db = tmp_path / "b200.log", tmp_path / "b200.db"
    with open(logf, "w") as f:
        assert _sample_log().save(f)
    p = subprocess.run([sys.executable, REF_PARSER, str(logf), "-d", str(db)], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=120)
    assert p.returncode == 0, p.stdout[-3000:]
    c = sqlite3.connect(str(db))
    assert c.execute("SELECT name, totaltime, seed FROM experiments").fetchall() == [("kNN_on_B200", 1.5, "7")]
    names = [r[0] for r in c.execute("SELECT name FROM plannerConfigs ORDER BY id")]
    assert names == ["geometric_PRM_NearestNeighborsCUDA", "geometric_PRM_GNAT"]
    rows = c.execute("SELECT plannerid, time, solved, status, graph_states, memory FROM runs ORDER BY id").fetchall()
    assert len(rows) == 5
    assert rows[0][1:5] == (0.25, 1, 6, 1000) and rows[1][2] == 0 and rows[1][3] == 4
    assert rows[4][1] == 3.5 and rows[4][2] is None and rows[4][5] == 12.25
    assert c.execute("SELECT COUNT(*) FROM progress").fetchone()[0] == 3
    en = dict(c.execute("SELECT value, description FROM enums WHERE name='status'").fetchall())
    assert en[6] == "Exact solution" and en[4] == "Timeout"
