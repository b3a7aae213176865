/*
 * BenchmarkLog.h -- writer of the tools::Benchmark text log, so that runs made with the B200 structures
 * land in the same files scripts/ompl_benchmark_statistics.py reads.
 *
 * Same record and the same text as tools::Benchmark::saveResultsToStream
 * (tools/benchmark/src/Benchmark.cpp:226-352, record: tools/benchmark/Benchmark.h, CompleteExperiment /
 * PlannerExperiment): header lines, experiment properties, the status enum, then per planner its common
 * properties, the sorted union of run property names ("name TYPE"), one "v; v; ...; " line per run, the
 * optional progress block and a closing ".".  Header-only, no Boost; the Python twin is
 * ompl_b200/benchmark_log.py.
 */
#ifndef OMPL_B200_BENCHMARK_LOG_H
#define OMPL_B200_BENCHMARK_LOG_H

#include <algorithm>
#include <map>
#include <ostream>
#include <string>
#include <vector>

namespace ompl
{
    namespace b200
    {
        struct BenchmarkLog
        {
            using RunProperties = std::map<std::string, std::string>;  // "name TYPE" -> value, e.g. "time REAL"
            using RunProgressData = std::vector<std::map<std::string, std::string>>;

            struct PlannerExperiment
            {
                std::string name;
                std::vector<RunProperties> runs;
                std::vector<std::string> progressPropertyNames;
                std::vector<RunProgressData> runsProgressData;
                RunProperties common;
            };

            std::string version{"1.6.0"};  // OMPL_VERSION of the reference (CMakeLists.txt:2)
            std::string name, host, startTime, setupInfo, cpuInfo;
            std::map<std::string, std::string> parameters;
            unsigned long seed{0};
            double maxTime{0.0}, maxMem{0.0}, totalDuration{0.0};
            unsigned int runCount{0};
            std::vector<PlannerExperiment> planners;

            /** base::PlannerStatus::asString for every StatusType in order (base/PlannerStatus.h:51-73,
                base/src/PlannerStatus.cpp:39-60; UNKNOWN and ABORT fall into the default branch). */
            static const std::vector<std::string> &statusNames()
            {
                static const std::vector<std::string> n = {"Unknown status",         "Invalid start",
                                                           "Invalid goal",           "Unrecognized goal type",
                                                           "Timeout",                "Approximate solution",
                                                           "Exact solution",         "Crash",
                                                           "Unknown status"};
                return n;
            }

            bool saveResultsToStream(std::ostream &out) const
            {
                if (planners.empty() || !out.good())
                    return false;
                out << "OMPL version " << version << std::endl;
                out << "Experiment " << (name.empty() ? "NO_NAME" : name) << std::endl;
                out << parameters.size() << " experiment properties" << std::endl;
                for (const auto &p : parameters)
                    out << p.first << " = " << p.second << std::endl;
                out << "Running on " << (host.empty() ? "UNKNOWN" : host) << std::endl;
                out << "Starting at " << startTime << std::endl;
                out << "<<<|" << std::endl << setupInfo << "|>>>" << std::endl;
                out << "<<<|" << std::endl << cpuInfo << "|>>>" << std::endl;
                out << seed << " is the random seed" << std::endl;
                out << maxTime << " seconds per run" << std::endl;
                out << maxMem << " MB per run" << std::endl;
                out << runCount << " runs per planner" << std::endl;
                out << totalDuration << " seconds spent to collect the data" << std::endl;
                out << "1 enum type" << std::endl;
                out << "status";
                for (const auto &s : statusNames())
                    out << '|' << s;
                out << std::endl;
                out << planners.size() << " planners" << std::endl;
                for (const auto &pl : planners)
                {
                    out << pl.name << std::endl;
                    out << pl.common.size() << " common properties" << std::endl;
                    for (const auto &c : pl.common)  // std::map: already sorted by name
                        out << c.first << " = " << c.second << std::endl;
                    std::vector<std::string> props;
                    for (const auto &run : pl.runs)
                        for (const auto &kv : run)
                            props.push_back(kv.first);
                    std::sort(props.begin(), props.end());
                    props.erase(std::unique(props.begin(), props.end()), props.end());
                    out << props.size() << " properties for each run" << std::endl;
                    for (const auto &p : props)
                        out << p << std::endl;
                    out << pl.runs.size() << " runs" << std::endl;
                    for (const auto &run : pl.runs)
                    {
                        for (const auto &p : props)
                        {
                            auto it = run.find(p);
                            if (it != run.end())
                                out << it->second;
                            out << "; ";
                        }
                        out << std::endl;
                    }
                    if (!pl.runsProgressData.empty())
                    {
                        out << pl.progressPropertyNames.size() << " progress properties for each run" << std::endl;
                        // the values of a time point come out in name order (a std::map, as in the reference, where
                        // the names are collected from the same map): list the names in that order too
                        std::vector<std::string> names(pl.progressPropertyNames);
                        std::sort(names.begin(), names.end());
                        for (const auto &n : names)
                            out << n << std::endl;
                        out << pl.runsProgressData.size() << " runs" << std::endl;
                        for (const auto &r : pl.runsProgressData)
                        {
                            for (const auto &t : r)
                            {
                                for (const auto &kv : t)
                                    out << kv.second << ",";
                                out << ";";
                            }
                            out << std::endl;
                        }
                    }
                    out << '.' << std::endl;
                }
                return true;
            }
        };
    }  // namespace b200
}  // namespace ompl

#endif
