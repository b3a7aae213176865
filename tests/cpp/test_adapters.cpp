// test_adapters.cpp -- the C++ drop-in adapters against the reference's own classes.
//
// Built only where /root/reference exists (tests/cpp/Makefile); the binary
// travels to the GPU box.  Links the reference's compiled state spaces and
// DiscreteMotionValidator (oracle/_ref/libompl_refbase.so) and libomplcu.so.
//
//   1. NearestNeighborsCUDA<State*> vs NearestNeighborsLinear<State*> on SE3 and
//      RealVector(7): the checks of tests/datastructures/nearestneighbors.cpp:117-206
//      (sizes, list, nearestK, nearestR(inf), nearestK(2n), remove, empty -> exception).
//   2. BatchedDiscreteMotionValidator vs DiscreteMotionValidator on SE2 with a circle
//      world: both checkMotion overloads, counters, batch entry point.
//   3. BatchedStateValidityChecker::isValidBatch vs the CPU predicate.
//   4. BatchPlanning.h (FMT* / PRM batch call sites) vs the sequential loops of FMT.cpp:169-240 and
//      PRM.cpp:562-596 run with the reference's NearestNeighborsLinear and DiscreteMotionValidator.
#include <cmath>
#include <cstdio>
#include <random>

#include "ompl/base/DiscreteMotionValidator.h"
#include "ompl/base/spaces/RealVectorStateSpace.h"
#include "ompl/base/spaces/SE2StateSpace.h"
#include "ompl/base/spaces/SE3StateSpace.h"
#include "ompl/datastructures/NearestNeighborsLinear.h"
#include "ompl_b200/BatchPlanning.h"
#include "ompl_b200/OmplBridge.h"

namespace ob = ompl::base;

static int failures = 0;
#define CHECK(cond)                                                              \
    do                                                                           \
    {                                                                            \
        if (!(cond))                                                             \
        {                                                                        \
            std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond);          \
            ++failures;                                                          \
        }                                                                        \
    } while (0)

static std::mt19937_64 gen(7);
static double uni(double a, double b)
{
    return std::uniform_real_distribution<double>(a, b)(gen);
}

static void randomSE3(const ob::StateSpacePtr &sp, ob::State *s, double lo, double hi)
{
    double q[4], n = 0;
    for (double &v : q)
    {
        v = std::normal_distribution<double>()(gen);
        n += v * v;
    }
    n = std::sqrt(n);
    std::vector<double> r = {uni(lo, hi), uni(lo, hi), uni(lo, hi), q[0] / n, q[1] / n, q[2] / n, q[3] / n};
    sp->copyFromReals(s, r);
}

template <typename Fill>
static void nnTest(const ob::StateSpacePtr &space, Fill fill, const char *name)
{
    const unsigned n = 200, k = 10;  // nearestneighbors.cpp:59-61
    std::vector<ob::State *> states(n);
    for (auto &s : states)
    {
        s = space->allocState();
        fill(s);
    }
    auto dist = [&space](ob::State *const &a, ob::State *const &b) { return space->distance(a, b); };
    ompl::NearestNeighborsLinear<ob::State *> lin;
    lin.setDistanceFunction(dist);
    ompl::NearestNeighborsCUDA<ob::State *> nn;
    ompl::b200::bind<ob::State *>(nn, space, [](ob::State *const &s) { return (const ob::State *)s; });
    nn.setDistanceFunction(dist);
    CHECK(nn.reportsSortedResults());

    bool threw = false;
    try
    {
        nn.nearest(states[0]);
    }
    catch (const ompl::Exception &e)
    {
        threw = std::string(e.what()).find("No elements found") == 0;
    }
    CHECK(threw);

    lin.add(states);
    nn.add(std::vector<ob::State *>(states.begin(), states.begin() + n / 2));  // bulk
    for (unsigned i = n / 2; i < n; ++i)
        nn.add(states[i]);  // one by one
    CHECK(nn.size() == n);
    std::vector<ob::State *> l1, l2;
    nn.list(l1);
    lin.list(l2);
    CHECK(l1 == l2);

    std::vector<ob::State *> a, b;
    for (unsigned i = 0; i < n; ++i)
    {
        nn.nearestK(states[i], k, a);
        lin.nearestK(states[i], k, b);
        CHECK(a.size() == k);
        CHECK(a == b);  // continuous data: no ties, identity must agree
        CHECK(nn.nearest(states[i]) == states[i]);
    }
    nn.nearestR(states[0], std::numeric_limits<double>::infinity(), a);
    lin.nearestR(states[0], std::numeric_limits<double>::infinity(), b);
    CHECK(a.size() == n && a[0] == states[0] && a == b);
    nn.nearestR(states[3], 0.5, a);
    lin.nearestR(states[3], 0.5, b);
    CHECK(a == b);
    nn.nearestK(states[0], 2 * n, a);
    CHECK(a.size() == n);

    // batch entry point == per-query calls
    std::vector<std::vector<ob::State *>> batch;
    std::vector<double> d;
    nn.nearestKBatch(states, k, batch, &d);
    for (unsigned i = 0; i < n; ++i)
    {
        lin.nearestK(states[i], k, b);
        CHECK(batch[i] == b);
        for (unsigned j = 0; j < k; ++j)
            CHECK(d[i * k + j] == space->distance(b[j], states[i]) ||
                  std::fabs(d[i * k + j] - space->distance(b[j], states[i])) < 1e-15);
    }

    for (unsigned i = 0; i < n; ++i)
    {
        CHECK(nn.remove(states[i]));
        lin.remove(states[i]);
        CHECK(nn.size() == n - i - 1);
        if (i % 40 == 0 && i + k + 1 < n)
        {
            nn.nearestK(states[n - 1], k, a);
            lin.nearestK(states[n - 1], k, b);
            CHECK(a == b);
        }
    }
    CHECK(!nn.remove(states[0]));
    nn.list(l1);
    CHECK(l1.empty());
    for (auto *s : states)
        space->freeState(s);
    std::printf("nn adapter (%s): done\n", name);
}

struct Circles
{
    std::vector<double> c;  // x, y, r
    bool valid(double x, double y) const
    {
        for (std::size_t i = 0; i < c.size(); i += 3)
        {
            const double dx = c[i] - x, dy = c[i + 1] - y;
            if (dx * dx + dy * dy < c[i + 2] * c[i + 2])  // tests/resources/circles2D.h:139-149
                return false;
        }
        return true;
    }
};

class CpuCircleChecker : public ob::StateValidityChecker
{
public:
    CpuCircleChecker(ob::SpaceInformation *si, const Circles *c) : ob::StateValidityChecker(si), c_(c)
    {
    }
    bool isValid(const ob::State *s) const override
    {
        const auto *se2 = s->as<ob::SE2StateSpace::StateType>();
        return c_->valid(se2->getX(), se2->getY());
    }
    const Circles *c_;
};

static void motionTest()
{
    auto space = std::make_shared<ob::SE2StateSpace>();
    ob::RealVectorBounds bounds(2);
    bounds.setLow(0);
    bounds.setHigh(1);
    space->setBounds(bounds);
    space->setup();
    Circles circ{{0.5, 0.5, 0.2, 0.2, 0.8, 0.1, 0.8, 0.2, 0.15}};

    auto siRef = std::make_shared<ob::SpaceInformation>(space);
    siRef->setStateValidityChecker(std::make_shared<CpuCircleChecker>(siRef.get(), &circ));
    ob::DiscreteMotionValidator ref(siRef.get());

    auto si = std::make_shared<ob::SpaceInformation>(space);
    auto handle = ompl::b200::CheckerHandle::circles2D(circ.c);
    auto svc = std::make_shared<ob::BatchedStateValidityChecker>(si, handle);
    si->setStateValidityChecker(svc);
    ob::BatchedDiscreteMotionValidator mv(si, handle);

    const unsigned n = 2000;
    std::vector<ob::State *> a(n), b(n);
    for (unsigned i = 0; i < n; ++i)
    {
        a[i] = space->allocState();
        b[i] = space->allocState();
        std::vector<double> ra = {uni(0, 1), uni(0, 1), uni(-M_PI, M_PI)}, rb = {uni(0, 1), uni(0, 1), uni(-M_PI, M_PI)};
        space->copyFromReals(a[i], ra);
        space->copyFromReals(b[i], rb);
    }
    // state validity: single and batch
    std::vector<std::uint8_t> sv(n);
    svc->isValidBatch(a.data(), n, sv.data());
    unsigned nvalid = 0;
    for (unsigned i = 0; i < n; ++i)
    {
        CHECK((sv[i] != 0) == siRef->isValid(a[i]));
        nvalid += sv[i];
    }
    CHECK(nvalid > 0 && nvalid < n);
    CHECK(si->isValid(a[0]) == siRef->isValid(a[0]));

    // motion validity: batch, then both single overloads on a subset
    std::vector<std::uint8_t> v(n);
    std::vector<double> t(n);
    mv.checkMotionBatch(a.data(), b.data(), n, v.data(), t.data());
    ob::State *scratch = space->allocState(), *scratch2 = space->allocState();
    unsigned nv = 0;
    for (unsigned i = 0; i < n; ++i)
    {
        std::pair<ob::State *, double> lv(scratch, -1.0);
        const bool ok = ref.checkMotion(a[i], b[i], lv);
        CHECK(ok == (v[i] != 0));
        if (!ok)
            CHECK(lv.second == t[i]);
        nv += ok;
        if (i < 200)
        {
            CHECK(mv.checkMotion(a[i], b[i]) == ok);
            std::pair<ob::State *, double> lv2(scratch2, -1.0);
            CHECK(mv.checkMotion(a[i], b[i], lv2) == ok);
            if (!ok)
            {
                CHECK(lv2.second == lv.second);
                CHECK(space->equalStates(lv2.first, lv.first));
            }
            else
                CHECK(lv2.second == -1.0);  // untouched on success (MotionValidator.h:93-95)
        }
    }
    CHECK(nv > 0 && nv < n);
    CHECK(mv.getCheckedMotionCount() == n + 400);
    CHECK(ref.getValidMotionCount() == nv);
    space->freeState(scratch);
    space->freeState(scratch2);
    for (unsigned i = 0; i < n; ++i)
    {
        space->freeState(a[i]);
        space->freeState(b[i]);
    }
    std::printf("validity adapters (SE2 circles): %u of %u motions valid\n", nv, n);
}

static void batchPlanningTest()
{
    auto space = std::make_shared<ob::SE2StateSpace>();
    ob::RealVectorBounds bounds(2);
    bounds.setLow(0);
    bounds.setHigh(1);
    space->setBounds(bounds);
    space->setup();
    Circles circ{{0.5, 0.5, 0.2, 0.2, 0.8, 0.1, 0.8, 0.2, 0.15, 0.15, 0.3, 0.08}};
    auto siRef = std::make_shared<ob::SpaceInformation>(space);
    siRef->setStateValidityChecker(std::make_shared<CpuCircleChecker>(siRef.get(), &circ));
    ob::DiscreteMotionValidator ref(siRef.get());
    auto si = std::make_shared<ob::SpaceInformation>(space);
    auto handle = ompl::b200::CheckerHandle::circles2D(circ.c);
    auto svc = std::make_shared<ob::BatchedStateValidityChecker>(si, handle);
    si->setStateValidityChecker(svc);
    ob::BatchedDiscreteMotionValidator mv(si, handle);

    // ---- FMT::sampleFree: same random stream, sequential vs batched ----
    const std::size_t wanted = 500;
    auto drawWith = [&space](std::mt19937_64 &g) {
        return [&space, &g](ob::State *s) {
            std::uniform_real_distribution<double> u(0, 1), a(-M_PI, M_PI);
            std::vector<double> r = {u(g), u(g), a(g)};
            space->copyFromReals(s, r);
        };
    };
    std::mt19937_64 g1(99), g2(99);
    std::vector<ob::State *> seq;
    std::size_t seqAttempts = 0;
    {
        auto draw = drawWith(g1);
        ob::State *s = space->allocState();
        while (seq.size() < wanted)
        {
            draw(s);
            ++seqAttempts;
            if (siRef->isValid(s))
            {
                seq.push_back(s);
                s = space->allocState();
            }
        }
        space->freeState(s);
    }
    std::vector<ob::State *> acc;
    const std::size_t attempts = ompl::b200::sampleFreeBatch(si, *svc, drawWith(g2), wanted, acc, 192);
    CHECK(acc.size() == wanted && attempts == seqAttempts && attempts > wanted);
    for (std::size_t i = 0; i < wanted && i < acc.size(); ++i)
        CHECK(space->equalStates(acc[i], seq[i]));

    // ---- FMT::saveNeighborhood for every node ----
    auto dist = [&space](ob::State *const &a, ob::State *const &b) { return space->distance(a, b); };
    ompl::NearestNeighborsLinear<ob::State *> lin;
    lin.setDistanceFunction(dist);
    ompl::NearestNeighborsCUDA<ob::State *> nn;
    ompl::b200::bind<ob::State *>(nn, space, [](ob::State *const &s) { return (const ob::State *)s; });
    const std::size_t roadmap = 300;
    std::vector<ob::State *> first(acc.begin(), acc.begin() + roadmap), rest(acc.begin() + roadmap, acc.end());
    lin.add(first);
    nn.add(first);
    std::vector<std::vector<ob::State *>> nbhK, nbhR;
    ompl::b200::saveNeighborhoodsBatch<ob::State *>(nn, first, 8, 0.0, nbhK);
    ompl::b200::saveNeighborhoodsBatch<ob::State *>(nn, first, 0, 0.25, nbhR);
    std::vector<ob::State *> want;
    for (std::size_t i = 0; i < roadmap; ++i)
    {
        lin.nearestK(first[i], 8, want);
        want.erase(want.begin());  // FMT.cpp:181-183
        CHECK(nbhK[i] == want);
        lin.nearestR(first[i], 0.25, want);
        want.erase(want.begin());
        CHECK(nbhR[i] == want);
    }

    // ---- PRM::addMilestone for a batch (semantics stated in BatchPlanning.h) ----
    const std::size_t k = 6;
    std::vector<ompl::b200::CandidateEdge<ob::State *>> edges;
    ompl::b200::connectMilestonesBatch<ob::State *>(
        nn, mv, [](ob::State *const &s) { return (const ob::State *)s; }, rest, k, edges);
    lin.add(rest);
    CHECK(nn.size() == lin.size());
    std::size_t e = 0, nvalid = 0;
    for (ob::State *m : rest)
    {
        lin.nearestK(m, k + 1, want);
        std::size_t taken = 0;
        for (ob::State *n : want)
        {
            if (n == m || taken == k)
                continue;
            ++taken;
            CHECK(e < edges.size() && edges[e].from == n && edges[e].to == m);
            const bool ok = ref.checkMotion(n, m);  // PRM.cpp:582
            if (e < edges.size())
                CHECK(edges[e].valid == ok);
            nvalid += ok;
            ++e;
        }
    }
    CHECK(e == edges.size() && nvalid > 0 && nvalid < e);

    // ---- lazy edge wavefront, explicit state sequence, path check ----
    std::vector<std::pair<const ob::State *, const ob::State *>> pairs;
    for (std::size_t i = 0; i + 7 < acc.size(); i += 3)
        pairs.emplace_back(acc[i], acc[i + 7]);
    std::vector<std::uint8_t> flags;
    ompl::b200::checkEdgesBatch(mv, pairs, flags);
    for (std::size_t i = 0; i < pairs.size(); ++i)
        CHECK((flags[i] != 0) == ref.checkMotion(pairs[i].first, pairs[i].second));

    // ---- BIT*: updateSamples loop, nearestSamples for a set of vertices, bulk prune ----
    {
        std::mt19937_64 h1(7), h2(7);
        const std::size_t have = 10, required = 260, factor = 100;
        std::vector<ob::State *> seqS;
        std::size_t seqTries = 0, n = have;
        {
            auto draw = drawWith(h1);  // ImplicitGraph.cpp:965-993
            for (; seqTries < factor * required && n < required; ++seqTries)
            {
                ob::State *s = space->allocState();
                draw(s);
                if (siRef->isValid(s))
                {
                    seqS.push_back(s);
                    ++n;
                }
                else
                    space->freeState(s);
            }
        }
        std::vector<ob::State *> batS;
        const std::size_t tries = ompl::b200::updateSamplesBatch(si, *svc, drawWith(h2), have, required, factor, batS, 64);
        CHECK(tries == seqTries && batS.size() == seqS.size() && batS.size() == required - have);
        for (std::size_t i = 0; i < batS.size() && i < seqS.size(); ++i)
            CHECK(space->equalStates(batS[i], seqS[i]));

        ompl::NearestNeighborsLinear<ob::State *> linS;
        linS.setDistanceFunction(dist);
        ompl::NearestNeighborsCUDA<ob::State *> samples;
        ompl::b200::bind<ob::State *>(samples, space, [](ob::State *const &s) { return (const ob::State *)s; });
        linS.add(batS);
        samples.add(batS);  // addToSamples: one bulk add
        std::vector<ob::State *> vertices(acc.begin(), acc.begin() + 40);
        std::vector<std::vector<ob::State *>> nsK, nsR;
        ompl::b200::nearestSamplesBatch<ob::State *>(samples, vertices, 7, 0.0, nsK);
        ompl::b200::nearestSamplesBatch<ob::State *>(samples, vertices, 0, 0.2, nsR);
        for (std::size_t i = 0; i < vertices.size(); ++i)
        {
            linS.nearestK(vertices[i], 7, want);
            CHECK(nsK[i] == want);
            linS.nearestR(vertices[i], 0.2, want);
            CHECK(nsR[i] == want);
        }
        std::vector<ob::State *> pruned;
        for (std::size_t i = 0; i < batS.size(); i += 3)
            pruned.push_back(batS[i]);
        for (ob::State *p : pruned)
            linS.remove(p);
        CHECK(ompl::b200::pruneSamplesBatch<ob::State *>(samples, pruned) == pruned.size());
        CHECK(ompl::b200::pruneSamplesBatch<ob::State *>(samples, pruned) == 0);  // already gone
        CHECK(samples.size() == linS.size());
        ompl::b200::nearestSamplesBatch<ob::State *>(samples, vertices, 7, 0.0, nsK);
        for (std::size_t i = 0; i < vertices.size(); ++i)
        {
            linS.nearestK(vertices[i], 7, want);
            CHECK(nsK[i] == want);
        }
    }

    std::vector<ob::State *> mixed(acc.begin(), acc.begin() + 40);
    ob::State *bad = space->allocState();
    std::vector<double> inside = {0.5, 0.5, 0.0};
    space->copyFromReals(bad, inside);
    mixed.insert(mixed.begin() + 17, bad);
    unsigned int firstBad = 1000, firstBadRef = 2000;
    CHECK(!ompl::b200::checkStatesBatch(*svc, mixed, (unsigned int)mixed.size(), firstBad));
    // the loop of SpaceInformation::checkMotion(states, count, firstInvalid), SpaceInformation.cpp:274-285
    // (SpaceInformation.cpp itself is not part of oracle/_ref: it needs the sampler / planner tree)
    for (unsigned int i = 0; i < mixed.size(); ++i)
        if (!siRef->isValid(mixed[i]))
        {
            firstBadRef = i;
            break;
        }
    CHECK(firstBad == 17 && firstBadRef == 17);
    CHECK(ompl::b200::checkStatesBatch(*svc, mixed, 17, firstBad));

    std::vector<ob::State *> pathOk, pathBad(acc.begin(), acc.begin() + 30);
    for (int i = 0; i < 20; ++i)
    {
        ob::State *s = space->allocState();
        std::vector<double> r = {0.02 + 0.002 * i, 0.02, 0.1 * i};
        space->copyFromReals(s, r);
        pathOk.push_back(s);
    }
    auto refPath = [&](const std::vector<ob::State *> &p) {  // PathGeometric.cpp:159-180
        if (p.empty())
            return true;
        if (!siRef->isValid(p[0]))
            return false;
        for (std::size_t j = 0; j + 1 < p.size(); ++j)
            if (!ref.checkMotion(p[j], p[j + 1]))
                return false;
        return true;
    };
    CHECK(ompl::b200::checkPathBatch(*svc, mv, pathOk) == refPath(pathOk) && refPath(pathOk));
    CHECK(ompl::b200::checkPathBatch(*svc, mv, pathBad) == refPath(pathBad) && !refPath(pathBad));
    CHECK(ompl::b200::checkPathBatch(*svc, mv, {}) && !ompl::b200::checkPathBatch(*svc, mv, {bad}));

    for (auto *s : pathOk)
        space->freeState(s);
    space->freeState(bad);
    for (auto *s : seq)
        space->freeState(s);
    for (auto *s : acc)
        space->freeState(s);
    std::printf("batch planning call sites: %zu samples in %zu attempts, %zu candidate edges (%zu valid)\n", wanted,
                attempts, edges.size(), nvalid);
}

int main()
{
    int ndev = 0;
    if (omplcu_device_count(&ndev) != OMPLCU_OK || ndev == 0)
    {
        std::printf("no CUDA device: %s\n", omplcu_last_error());
        return 2;
    }
    {
        auto se3 = std::make_shared<ob::SE3StateSpace>();
        ob::RealVectorBounds b(3);
        b.setLow(0);
        b.setHigh(1);
        se3->setBounds(b);
        se3->setup();
        nnTest(se3, [&](ob::State *s) { randomSE3(se3, s, 0, 1); }, "SE3");
    }
    {
        auto rv = std::make_shared<ob::RealVectorStateSpace>(7);
        ob::RealVectorBounds b(7);
        b.setLow(0);
        b.setHigh(1);
        rv->setBounds(b);
        rv->setup();
        nnTest(rv, [&](ob::State *s) {
            for (int i = 0; i < 7; ++i)
                s->as<ob::RealVectorStateSpace::StateType>()->values[i] = uni(0, 1);
        }, "RealVector(7)");
    }
    motionTest();
    batchPlanningTest();
    std::printf(failures ? "FAILED (%d)\n" : "all adapter tests passed\n", failures);
    return failures ? 1 : 0;
}
