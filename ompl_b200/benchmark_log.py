"""Writer of the tools::Benchmark text log (tools/benchmark/src/Benchmark.cpp:226-352), so results measured
with the B200 structures can be read by the reference's scripts/ompl_benchmark_statistics.py next to the
logs of demos/HypercubeBenchmark and demos/KinematicChainBenchmark.

The record mirrors tools::Benchmark::CompleteExperiment / PlannerExperiment; run properties are keyed
"name TYPE" ("time REAL", "solved BOOLEAN", "status ENUM" ...), values are strings as the reference stores
them.  The C++ twin is include/ompl_b200/BenchmarkLog.h.
"""
import socket
import time

# base::PlannerStatus::asString for StatusType 0..8 (base/PlannerStatus.h:51-73, PlannerStatus.cpp:39-60)
STATUS_NAMES = ["Unknown status", "Invalid start", "Invalid goal", "Unrecognized goal type", "Timeout",
                "Approximate solution", "Exact solution", "Crash", "Unknown status"]
OMPL_VERSION = "1.6.0"  # CMakeLists.txt:2 of the reference


def _num(v):
    """Numbers as the reference's std::to_string / operator<< print them (plain repr is parsed the same)."""
    if isinstance(v, bool):
        return "1" if v else "0"
    return str(v)


class PlannerExperiment:
    def __init__(self, name, common=None):
        self.name = name
        self.common = dict(common or {})      # "name TYPE" -> value
        self.runs = []                        # list of {"name TYPE": value}
        self.progress_property_names = []
        self.runs_progress_data = []          # per run: list of {"name TYPE": value} per time point

    def add_run(self, **props):
        """add_run(time=("REAL", 0.12), solved=("BOOLEAN", True), ...) or {"time REAL": "0.12"} via runs.append."""
        self.runs.append({f"{k.replace('_', ' ')} {t}": _num(v) for k, (t, v) in props.items()})


class BenchmarkLog:
    def __init__(self, name, seed=0, max_time=0.0, max_mem=0.0, run_count=0):
        self.name = name
        self.parameters = {}
        self.host = socket.gethostname()
        self.start_time = time.strftime("%Y-%m-%d %H:%M:%S")
        self.setup_info = ""
        self.cpu_info = ""
        self.seed = seed
        self.max_time = max_time
        self.max_mem = max_mem
        self.run_count = run_count
        self.total_duration = 0.0
        self.planners = []

    def add_planner(self, name, common=None):
        p = PlannerExperiment(name, common)
        self.planners.append(p)
        return p

    def save(self, out):
        """saveResultsToStream: `out` is a text file object.  Returns False if there is nothing to save."""
        if not self.planners:
            return False
        w = out.write
        w(f"OMPL version {OMPL_VERSION}\n")
        w(f"Experiment {self.name or 'NO_NAME'}\n")
        w(f"{len(self.parameters)} experiment properties\n")
        for k in sorted(self.parameters):
            w(f"{k} = {self.parameters[k]}\n")
        w(f"Running on {self.host or 'UNKNOWN'}\n")
        w(f"Starting at {self.start_time}\n")
        w(f"<<<|\n{self.setup_info}|>>>\n")
        w(f"<<<|\n{self.cpu_info}|>>>\n")
        w(f"{self.seed} is the random seed\n")
        w(f"{_num(self.max_time)} seconds per run\n")
        w(f"{_num(self.max_mem)} MB per run\n")
        w(f"{self.run_count} runs per planner\n")
        w(f"{_num(self.total_duration)} seconds spent to collect the data\n")
        w("1 enum type\n")
        w("status" + "".join("|" + s for s in STATUS_NAMES) + "\n")
        w(f"{len(self.planners)} planners\n")
        for pl in self.planners:
            w(pl.name + "\n")
            w(f"{len(pl.common)} common properties\n")
            for k in sorted(pl.common):
                w(f"{k} = {pl.common[k]}\n")
            props = sorted({k for run in pl.runs for k in run})
            w(f"{len(props)} properties for each run\n")
            for p in props:
                w(p + "\n")
            w(f"{len(pl.runs)} runs\n")
            for run in pl.runs:
                w("".join((run.get(p, "") + "; ") for p in props) + "\n")
            if pl.runs_progress_data:
                w(f"{len(pl.progress_property_names)} progress properties for each run\n")
                for n in sorted(pl.progress_property_names):  # a std::map in the reference: name order
                    w(n + "\n")
                w(f"{len(pl.runs_progress_data)} runs\n")
                for r in pl.runs_progress_data:
                    w("".join("".join(_num(t[k]) + "," for k in sorted(t)) + ";" for t in r) + "\n")
            w(".\n")
        return True
