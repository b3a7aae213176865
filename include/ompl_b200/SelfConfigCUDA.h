/*
 * SelfConfigCUDA.h -- default nearest-neighbour structure for planners: the B200 one where the state
 * space has a device metric, the reference's choice otherwise.
 *
 * Replaces the call sites of tools::SelfConfig::getDefaultNearestNeighbors<_T>(planner)
 * (tools/config/SelfConfig.h:105-117), e.g. prm/src/PRM.cpp:119-121, fmt/src/FMT.cpp:99,
 * informedtrees/bitstar/src/ImplicitGraph.cpp:112: a planner's setup() does
 *     if (!nn_) nn_.reset(ompl::b200::getDefaultNearestNeighbors<Motion *>(this, [](Motion *const &m) { return m->state; }));
 * and keeps everything else.  SURVEY 8(f) row 4.
 *
 * Rule:
 *   - the space is exactly RealVector / SO2 / SO3 / SE2 / SE3 / a flat CompoundStateSpace of those
 *     (describe() vouches for its metric) and a GPU is present -> NearestNeighborsCUDA<_T>, bound to the space
 *     and the accessor (sharded over `devices` if more than one is given);
 *   - anything else (custom metrics, Dubins, projections ...)     -> the reference's own rule: GNAT for
 *     metric spaces (thread-safe variant for multithreaded planners), SqrtApprox otherwise.
 * A caller that knows better passes an explicit SpaceDescription (e.g. the kinematic chain).
 */
#ifndef OMPL_B200_SELF_CONFIG_CUDA_H
#define OMPL_B200_SELF_CONFIG_CUDA_H

#include "ompl/base/Planner.h"
#include "ompl/datastructures/NearestNeighborsGNAT.h"
#include "ompl/datastructures/NearestNeighborsGNATNoThreadSafety.h"
#include "ompl/datastructures/NearestNeighborsSqrtApprox.h"
#include "ompl_b200/OmplBridge.h"

namespace ompl
{
    namespace b200
    {
        /** True if describe() accepts the space and a CUDA device is visible. */
        inline bool deviceMetricAvailable(const base::StateSpacePtr &space)
        {
            int n = 0;
            if (omplcu_device_count(&n) != OMPLCU_OK || n == 0)
                return false;
            try
            {
                describe(space);
            }
            catch (const Exception &)
            {
                return false;
            }
            return true;
        }

        /** The reference's rule, verbatim in effect (tools/config/SelfConfig.h:105-117). */
        template <typename _T>
        NearestNeighbors<_T> *referenceDefaultNearestNeighbors(const base::StateSpacePtr &space,
                                                               const base::PlannerSpecs &specs)
        {
            if (space->isMetricSpace())
            {
                if (specs.multithreaded)
                    return new NearestNeighborsGNAT<_T>();
                return new NearestNeighborsGNATNoThreadSafety<_T>();
            }
            return new NearestNeighborsSqrtApprox<_T>();
        }

        /** Default structure for elements of type _T living in `space`.  The caller owns the result and still
            calls setDistanceFunction() on it, as planners do for every structure. */
        template <typename _T>
        NearestNeighbors<_T> *getDefaultNearestNeighbors(const base::StateSpacePtr &space,
                                                         const base::PlannerSpecs &specs,
                                                         std::function<const base::State *(const _T &)> accessor,
                                                         const std::vector<int> &devices = std::vector<int>())
        {
            if (accessor && deviceMetricAvailable(space))
            {
                auto *nn = new NearestNeighborsCUDA<_T>();
                try
                {
                    if (devices.size() == 1)
                        nn->setDevice(devices[0]);
                    else if (devices.size() > 1)
                        nn->setDevices(devices);
                    bind<_T>(*nn, space, std::move(accessor));
                }
                catch (...)
                {
                    delete nn;
                    throw;
                }
                return nn;
            }
            return referenceDefaultNearestNeighbors<_T>(space, specs);
        }

        /** With an explicit description (spaces describe() refuses, e.g. SpaceDescription::kinematicChain). */
        template <typename _T>
        NearestNeighbors<_T> *getDefaultNearestNeighbors(const base::StateSpacePtr &space,
                                                         const SpaceDescription &description,
                                                         std::function<const base::State *(const _T &)> accessor,
                                                         const std::vector<int> &devices = std::vector<int>())
        {
            auto *nn = new NearestNeighborsCUDA<_T>();
            try
            {
                if (devices.size() == 1)
                    nn->setDevice(devices[0]);
                else if (devices.size() > 1)
                    nn->setDevices(devices);
                bind<_T>(*nn, space, std::move(accessor), description);
            }
            catch (...)
            {
                delete nn;
                throw;
            }
            return nn;
        }

        /** The planner-facing form, same signature idea as SelfConfig::getDefaultNearestNeighbors<_T>(planner). */
        template <typename _T>
        NearestNeighbors<_T> *getDefaultNearestNeighbors(const base::Planner *planner,
                                                         std::function<const base::State *(const _T &)> accessor,
                                                         const std::vector<int> &devices = std::vector<int>())
        {
            return getDefaultNearestNeighbors<_T>(planner->getSpaceInformation()->getStateSpace(), planner->getSpecs(),
                                                  std::move(accessor), devices);
        }
    }  // namespace b200
}  // namespace ompl

#endif
