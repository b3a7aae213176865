// bench_adapter.cpp -- what a C++ OMPL caller gets from the adapter at the headline configuration
// (RealVectorStateSpace(7), 1M states, 65536 queries, k = 10): nearestKBatch(std::vector<State*>)
// INCLUDING the host work around the C-ABI call -- copyToReals of every query through the row function,
// the call with pageable std::vector buffers, and the id -> element mapping into std::vector<State*>.
// Usage: bench_adapter [n_states] [n_queries] [k] [reps] [ndevices]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>

#include "ompl/base/spaces/RealVectorStateSpace.h"
#include "ompl_b200/OmplBridge.h"

namespace ob = ompl::base;

int main(int argc, char **argv)
{
    const std::size_t n = argc > 1 ? std::atol(argv[1]) : 1000000, nq = argc > 2 ? std::atol(argv[2]) : 65536,
                      k = argc > 3 ? std::atol(argv[3]) : 10;
    const int reps = argc > 4 ? std::atoi(argv[4]) : 5, nd = argc > 5 ? std::atoi(argv[5]) : 1;
    int ndev = 0;
    if (omplcu_device_count(&ndev) != OMPLCU_OK || ndev == 0)
    {
        std::printf("no CUDA device: %s\n", omplcu_last_error());
        return 2;
    }
    auto rv = std::make_shared<ob::RealVectorStateSpace>(7);
    ob::RealVectorBounds b(7);
    b.setLow(0);
    b.setHigh(1);
    rv->setBounds(b);
    rv->setup();
    ob::StateSpacePtr sp = rv;
    std::mt19937_64 gen(1);
    std::uniform_real_distribution<double> u(0, 1);
    auto make = [&](std::size_t m) {
        std::vector<ob::State *> v(m);
        for (auto &s : v)
        {
            s = sp->allocState();
            for (int i = 0; i < 7; ++i)
                s->as<ob::RealVectorStateSpace::StateType>()->values[i] = u(gen);
        }
        return v;
    };
    std::vector<ob::State *> states = make(n), queries = make(nq);
    ompl::NearestNeighborsCUDA<ob::State *> nn;
    if (nd > 1)
    {
        std::vector<int> devs;
        for (int i = 0; i < nd; ++i)
            devs.push_back(i % ndev);
        nn.setDevices(devs);
    }
    ompl::b200::bind<ob::State *>(nn, sp, [](ob::State *const &s) { return (const ob::State *)s; });
    auto t0 = std::chrono::steady_clock::now();
    nn.add(states);
    const double add_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::vector<std::vector<ob::State *>> out;
    double best = 1e30;
    for (int r = 0; r < reps + 1; ++r)
    {
        t0 = std::chrono::steady_clock::now();
        nn.nearestKBatch(queries, k, out);
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (r > 0 && s < best)
            best = s;
    }
    // the flat variant (no vector per query)
    std::vector<ob::State *> flat;
    double best_flat = 1e30;
    std::size_t kk = 0;
    for (int r = 0; r < reps + 1; ++r)
    {
        t0 = std::chrono::steady_clock::now();
        kk = nn.nearestKBatchFlat(queries, k, flat);
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (r > 0 && s < best_flat)
            best_flat = s;
    }
    std::size_t check = 0;
    bool same = flat.size() == nq * kk;
    for (std::size_t i = 0; i < out.size(); ++i)
    {
        check += out[i].size();
        for (std::size_t j = 0; same && j < out[i].size(); ++j)
            same = flat[i * kk + j] == out[i][j];
    }
    if (!same)
    {
        std::printf("nearestKBatchFlat and nearestKBatch disagree\n");
        return 1;
    }
    std::printf("{\"adapter\": \"NearestNeighborsCUDA<State*>::nearestKBatch\", \"n_states\": %zu, \"n_queries\": %zu, "
                "\"k\": %zu, \"devices\": %d, \"add_s\": %.3f, \"batch_ms\": %.3f, \"queries_per_s\": %.0f, "
                "\"flat_batch_ms\": %.3f, \"flat_queries_per_s\": %.0f, \"neighbours\": %zu}\n",
                n, nq, k, nd, add_s, best * 1e3, (double)nq / best, best_flat * 1e3, (double)nq / best_flat, check);
    return check == nq * k ? 0 : 1;
}
