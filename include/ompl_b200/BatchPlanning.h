/* BatchPlanning.h -- batch-submit call sites for sampling-based planners.
 *
 * The reference planners call the hot path one element at a time:
 *   FMT*   sampleFree          one isValid per sample          geometric/planners/fmt/src/FMT.cpp:212-240
 *          saveNeighborhood    one nearestK / nearestR per node                          FMT.cpp:169-191
 *          expandTreeFromNode  one checkMotion per candidate edge                        FMT.cpp:498-575
 *   PRM    addMilestone        connectionStrategy_(m) then one checkMotion per neighbour
 *                                                      geometric/planners/prm/src/PRM.cpp:562-596
 *          KStrategy           nn->nearestK(m, k, neighbors)   geometric/planners/prm/ConnectionStrategy.h:77-81
 *
 * The functions below are those loops with the per-element calls replaced by ONE batch call on
 * the adapters of OmplBridge.h.  They hold no planner state (graphs, trees and queues stay the
 * planner's); each states the exact sequential loop it is equivalent to, and
 * tests/cpp/test_adapters.cpp runs that loop with the reference's own NearestNeighborsLinear and
 * DiscreteMotionValidator beside it.
 */
#ifndef OMPL_B200_BATCH_PLANNING_H
#define OMPL_B200_BATCH_PLANNING_H

#include <algorithm>
#include <cstdint>
#include <functional>
#include <utility>
#include <vector>

#include "ompl_b200/OmplBridge.h"

namespace ompl
{
    namespace b200
    {
        /** FMT::sampleFree (FMT.cpp:212-240) for a batch: `draw(state)` is called exactly as the
            sequential loop would call sampler_->sampleUniform, `chunk` states at a time, one
            isValidBatch per chunk; valid states are kept IN DRAWING ORDER until `wanted` are found.
            Returns the number of draws up to and including the last accepted one -- the
            `sampleAttempts` of the reference loop -- so accepted set, order and attempt count are
            those of the sequential loop run on the same random stream.  Unused states of the last
            chunk are freed; `maxAttempts` bounds the loop (the reference uses the termination
            condition). */
        inline std::size_t sampleFreeBatch(const base::SpaceInformationPtr &si,
                                           const base::BatchedStateValidityChecker &svc,
                                           const std::function<void(base::State *)> &draw, std::size_t wanted,
                                           std::vector<base::State *> &accepted, std::size_t chunk = 65536,
                                           std::size_t maxAttempts = ~std::size_t(0))
        {
            std::size_t attempts = 0, attemptsAtLastAccept = 0;
            std::vector<base::State *> cand;
            std::vector<std::uint8_t> ok;
            while (accepted.size() < wanted && attempts < maxAttempts)
            {
                const std::size_t n = std::min(chunk, maxAttempts - attempts);
                cand.resize(n);
                ok.resize(n);
                for (auto &s : cand)
                {
                    s = si->allocState();
                    draw(s);
                }
                svc.isValidBatch(cand.data(), n, ok.data());
                for (std::size_t i = 0; i < n; ++i)
                {
                    if (accepted.size() < wanted)
                    {
                        ++attempts;
                        if (ok[i] != 0)
                        {
                            accepted.push_back(cand[i]);
                            attemptsAtLastAccept = attempts;
                            continue;
                        }
                    }
                    si->freeState(cand[i]);
                }
            }
            return accepted.size() == wanted ? attemptsAtLastAccept : attempts;
        }

        /** FMT::saveNeighborhood (FMT.cpp:169-191) for all nodes at once.  neighborhoods[i] is what the
            reference stores for nodes[i]: nearestK(m, k) (or nearestR(m, r) when k == 0) WITHOUT its
            first element, which is m itself.  One batch query instead of nodes.size() queries. */
        template <typename _T>
        void saveNeighborhoodsBatch(const NearestNeighborsCUDA<_T> &nn, const std::vector<_T> &nodes, std::size_t k,
                                    double r, std::vector<std::vector<_T>> &neighborhoods)
        {
            if (k != 0)
                nn.nearestKBatch(nodes, k, neighborhoods);
            else
                nn.nearestRBatch(nodes, std::vector<double>(nodes.size(), r), neighborhoods);
            for (auto &nbh : neighborhoods)
                if (!nbh.empty())
                    nbh.erase(nbh.begin());
        }

        /** One candidate edge of a roadmap / tree. */
        template <typename _T>
        struct CandidateEdge
        {
            _T from, to;
            bool valid{false};
        };

        /** PRM::addMilestone (PRM.cpp:562-596) for a BATCH of new milestones.

            Batch semantics (stated, because the sequential planner connects milestone i to the
            milestones that existed before i): all milestones of the batch are added to `nn` first;
            every one is then connected to its k nearest among the roadmap INCLUDING the whole batch,
            itself excluded -- i.e. the loop
                for m in batch: nn.add(m)
                for m in batch: for n in nearestK(m, k + 1) minus m: edge(n, m) iff checkMotion(n, m)
            which is sPRM / PRM* construction order.  `state(x)` maps an element to its State.
            Edges come out grouped by milestone, neighbours in nearestK order; checkMotion(n, m) keeps
            the reference's argument order (neighbour first).  Two batch calls replace
            batch.size() * (1 + k) single calls. */
        template <typename _T>
        void connectMilestonesBatch(NearestNeighborsCUDA<_T> &nn, const base::BatchedDiscreteMotionValidator &mv,
                                    const std::function<const base::State *(const _T &)> &state,
                                    const std::vector<_T> &batch, std::size_t k,
                                    std::vector<CandidateEdge<_T>> &edges)
        {
            edges.clear();
            if (batch.empty() || k == 0)
                return;
            nn.add(batch);
            std::vector<std::vector<_T>> nbh;
            nn.nearestKBatch(batch, k + 1, nbh);
            std::vector<const base::State *> s1, s2;
            for (std::size_t i = 0; i < batch.size(); ++i)
            {
                std::size_t taken = 0;
                for (const _T &n : nbh[i])
                {
                    if (n == batch[i] || taken == k)
                        continue;
                    ++taken;
                    edges.push_back(CandidateEdge<_T>{n, batch[i], false});
                    s1.push_back(state(n));
                    s2.push_back(state(batch[i]));
                }
            }
            std::vector<std::uint8_t> ok(edges.size());
            mv.checkMotionBatch(s1.data(), s2.data(), edges.size(), ok.data(), nullptr);
            for (std::size_t e = 0; e < edges.size(); ++e)
                edges[e].valid = ok[e] != 0;
        }

        /** The lazy edge checks of FMT::expandTreeFromNode (FMT.cpp:559-573) and BITstar::checkEdge
            (informedtrees/src/BITstar.cpp:799-817) for a wavefront: every (from, to) pair in one
            call.  flags[i] = checkMotion(from_i, to_i). */
        inline void checkEdgesBatch(const base::BatchedDiscreteMotionValidator &mv,
                                    const std::vector<std::pair<const base::State *, const base::State *>> &pairs,
                                    std::vector<std::uint8_t> &flags)
        {
            std::vector<const base::State *> s1(pairs.size()), s2(pairs.size());
            for (std::size_t i = 0; i < pairs.size(); ++i)
            {
                s1[i] = pairs[i].first;
                s2[i] = pairs[i].second;
            }
            flags.resize(pairs.size());
            mv.checkMotionBatch(s1.data(), s2.data(), pairs.size(), flags.data(), nullptr);
        }

        /** BIT*: the sampling loop of ImplicitGraph::updateSamples
            (geometric/planners/informedtrees/bitstar/src/ImplicitGraph.cpp:965-993)

                for (tries = 0; tries < factor * numRequired && numSamples < numRequired; ++tries)
                    draw a state; if isValid: keep it, ++numSamples

            with isValid replaced by isValidBatch over chunks of the same random stream.  States are accepted
            in drawing order, so the kept states and the returned number of tries (= state collision checks)
            equal the sequential loop's.  `draw` fills one state; `accepted` receives the kept states (owned by
            the caller, allocated with si->allocState()); the caller adds them to the sample structure with
            one add(vector) (addToSamples, :996). */
        inline std::size_t updateSamplesBatch(const base::SpaceInformationPtr &si,
                                              const base::BatchedStateValidityChecker &svc,
                                              const std::function<void(base::State *)> &draw, std::size_t numHave,
                                              std::size_t numRequired, std::size_t allowedFailedFactor,
                                              std::vector<base::State *> &accepted, std::size_t chunk = 16384)
        {
            const std::size_t limit = allowedFailedFactor * numRequired;
            std::size_t tries = 0;
            while (tries < limit && numHave < numRequired)
            {
                const std::size_t n = std::min(chunk, limit - tries);
                std::vector<base::State *> cand(n);
                for (auto &s : cand)
                {
                    s = si->allocState();
                    draw(s);
                }
                std::vector<std::uint8_t> ok(n);
                svc.isValidBatch(cand.data(), n, ok.data());
                std::size_t used = n;
                for (std::size_t i = 0; i < n; ++i)
                {
                    if (numHave < numRequired && ok[i] != 0)
                    {
                        accepted.push_back(cand[i]);
                        cand[i] = nullptr;
                        if (++numHave == numRequired)
                            used = i + 1;
                    }
                }
                for (auto *s : cand)
                    if (s != nullptr)
                        si->freeState(s);
                tries += used;
            }
            return tries;
        }

        /** BIT*: ImplicitGraph::nearestSamples (ImplicitGraph.cpp:303-321) for every vertex about to be
            expanded: nearestK(vertex, k) if k != 0, else nearestR(vertex, r), in one batched call. */
        template <typename _T>
        void nearestSamplesBatch(const NearestNeighborsCUDA<_T> &samples, const std::vector<_T> &vertices,
                                 std::size_t k, double r, std::vector<std::vector<_T>> &neighbourSamples)
        {
            if (k != 0)
                samples.nearestKBatch(vertices, k, neighbourSamples);
            else
                samples.nearestRBatch(vertices, std::vector<double>(vertices.size(), r), neighbourSamples);
        }

        /** BIT*: the samples_->remove(sample) calls of ImplicitGraph::pruneSamples as one bulk removal
            (tombstones on the device; compaction when more than half of the slots are dead). */
        template <typename _T>
        std::size_t pruneSamplesBatch(NearestNeighborsCUDA<_T> &samples, const std::vector<_T> &pruned)
        {
            return samples.removeBatch(pruned);
        }

        /** SpaceInformation::checkMotion(states, count, firstInvalidStateIndex)
            (base/src/SpaceInformation.cpp:274-285) with one isValidBatch. */
        inline bool checkStatesBatch(const base::BatchedStateValidityChecker &svc,
                                     const std::vector<base::State *> &states, unsigned int count,
                                     unsigned int &firstInvalidStateIndex)
        {
            std::vector<std::uint8_t> ok(count);
            svc.isValidBatch(states.data(), count, ok.data());
            for (unsigned int i = 0; i < count; ++i)
                if (ok[i] == 0)
                {
                    firstInvalidStateIndex = i;
                    return false;
                }
            return true;
        }

        /** PathGeometric::check (geometric/src/PathGeometric.cpp:159-180): the first state, then every
            consecutive pair of the path through the motion validator, in one call. */
        inline bool checkPathBatch(const base::BatchedStateValidityChecker &svc,
                                   const base::BatchedDiscreteMotionValidator &mv,
                                   const std::vector<base::State *> &path)
        {
            if (path.empty())
                return true;
            if (!svc.isValid(path[0]))
                return false;
            if (path.size() < 2)
                return true;
            std::vector<const base::State *> s1(path.begin(), path.end() - 1), s2(path.begin() + 1, path.end());
            std::vector<std::uint8_t> ok(s1.size());
            mv.checkMotionBatch(s1.data(), s2.data(), s1.size(), ok.data(), nullptr);
            for (auto f : ok)
                if (f == 0)
                    return false;
            return true;
        }

        /** SpaceInformation::getMotionStates (base/src/SpaceInformation.cpp:201-272) with the interpolation
            done as ONE batched device call: the same control flow (count + 1 segments, optional end points,
            optional allocation, never more states than `states` holds when alloc is false), the states in
            between are interpolate(s1, s2, j / count) evaluated by omplcu_space_interpolate with the
            reference's operations.  `space` describes si's state space (describe() or an explicit description). */
        inline unsigned int getMotionStates(const base::SpaceInformation *si, const SpaceHandle &space,
                                            const base::State *s1, const base::State *s2,
                                            std::vector<base::State *> &states, unsigned int count, bool endpoints,
                                            bool alloc, int device = 0)
        {
            count++;
            if (count < 2)
            {
                unsigned int added = 0;
                if (endpoints)
                {
                    if (alloc)
                    {
                        states.resize(2);
                        states[0] = si->allocState();
                        states[1] = si->allocState();
                    }
                    if (states.size() > 0)
                    {
                        si->copyState(states[0], s1);
                        added++;
                    }
                    if (states.size() > 1)
                    {
                        si->copyState(states[1], s2);
                        added++;
                    }
                }
                else if (alloc)
                    states.resize(0);
                return added;
            }
            if (alloc)
            {
                states.resize(count + (endpoints ? 1 : -1));
                if (endpoints)
                    states[0] = si->allocState();
            }
            unsigned int added = 0;
            if (endpoints && states.size() > 0)
            {
                si->copyState(states[0], s1);
                added++;
            }
            // the states in between: j = 1 .. count - 1 while there is room
            const std::size_t room = states.size() > added ? states.size() - added : 0;
            const std::size_t m = std::min<std::size_t>(count - 1, room), st = space.stride();
            if (m > 0)
            {
                std::vector<double> a, b, from(m * st), to(m * st), t(m), out(m * st);
                gatherRows(si->getStateSpace().get(), &s1, 1, st, a);
                gatherRows(si->getStateSpace().get(), &s2, 1, st, b);
                for (std::size_t j = 0; j < m; ++j)
                {
                    std::copy(a.begin(), a.end(), from.begin() + j * st);
                    std::copy(b.begin(), b.end(), to.begin() + j * st);
                    t[j] = (double)(j + 1) / (double)count;
                }
                check(omplcu_space_interpolate(space.get(), device, from.data(), to.data(), t.data(), m, out.data()));
                std::vector<double> row(st);
                for (std::size_t j = 0; j < m; ++j)
                {
                    if (alloc)
                        states[added] = si->allocState();
                    row.assign(out.begin() + j * st, out.begin() + (j + 1) * st);
                    si->getStateSpace()->copyFromReals(states[added], row);
                    added++;
                }
            }
            if (added < states.size() && endpoints)
            {
                if (alloc)
                    states[added] = si->allocState();
                si->copyState(states[added], s2);
                added++;
            }
            return added;
        }

        /** StateSpace::satisfiesBounds for a batch (RealVectorStateSpace.cpp:200-208 with its DBL_EPSILON slack,
            SO2 / SO3 / compound rules; valid-state samplers call it before isValid,
            base/samplers/src/UniformValidStateSampler.cpp:53).  low / high: one entry per stored value in
            copyToReals order (entries of SO2 / SO3 values are ignored). */
        inline void satisfiesBoundsBatch(const base::StateSpace *stateSpace, const SpaceHandle &space,
                                         const std::vector<double> &low, const std::vector<double> &high,
                                         const base::State *const *states, std::size_t n, std::uint8_t *inBounds,
                                         int device = 0)
        {
            if (low.size() != space.stride() || high.size() != space.stride())
                throw Exception("ompl::b200::satisfiesBoundsBatch: one low / high entry per stored value");
            std::vector<double> rows;
            gatherRows(stateSpace, states, n, space.stride(), rows);
            check(omplcu_space_satisfies_bounds(space.get(), device, low.data(), high.data(), rows.data(), n, inBounds));
        }

        /** StateSpace::getMaximumExtent of the described space (RealVectorStateSpace.cpp:167-176 and the
            SO2 / SO3 / compound rules). */
        inline double getMaximumExtent(const SpaceHandle &space, const std::vector<double> &low,
                                       const std::vector<double> &high)
        {
            double e = 0.0;
            check(omplcu_space_maximum_extent(space.get(), low.data(), high.data(), &e));
            return e;
        }
    }  // namespace b200
}  // namespace ompl

#endif
