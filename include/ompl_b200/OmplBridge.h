/*
 * OmplBridge.h -- glue between ompl::base objects and the C ABI (include/omplcu.h).
 *
 * Needs ompl/base (StateSpace, SpaceInformation), i.e. an OMPL build tree with
 * Boost; NearestNeighborsCUDA.h itself does not.
 *
 *   describe(space)                  base::StateSpacePtr -> SpaceDescription
 *                                    (RealVector, SO2, SO3, SE2, SE3, flat compounds of those)
 *   bind(nn, space, accessor)        wires a NearestNeighborsCUDA<_T> to a space
 *   BatchedStateValidityChecker      base::StateValidityChecker + isValidBatch
 *   BatchedDiscreteMotionValidator   base::MotionValidator (both checkMotion
 *                                    overloads of base/MotionValidator.h:83,99,
 *                                    semantics of base/src/DiscreteMotionValidator.cpp:48-145)
 *                                    + checkMotionBatch
 *
 * States cross the boundary as rows of doubles in StateSpace::copyToReals order
 * (base/src/StateSpace.cpp:329-335).
 */
#ifndef OMPL_B200_OMPL_BRIDGE_H
#define OMPL_B200_OMPL_BRIDGE_H

#include <atomic>
#include <memory>
#include <mutex>
#include <typeinfo>
#include <utility>
#include <vector>

#include "ompl/base/MotionValidator.h"
#include "ompl/base/SpaceInformation.h"
#include "ompl/base/StateSpace.h"
#include "ompl/base/StateValidityChecker.h"
#include "ompl/base/spaces/RealVectorStateSpace.h"
#include "ompl/base/spaces/SE2StateSpace.h"
#include "ompl/base/spaces/SE3StateSpace.h"
#include "ompl/base/spaces/SO2StateSpace.h"
#include "ompl/base/spaces/SO3StateSpace.h"
#include "ompl_b200/NearestNeighborsCUDA.h"

namespace ompl
{
    namespace b200
    {
        namespace detail
        {
            // The device evaluates the STOCK distance / interpolate of these classes.  A subclass may
            // override either (demos/KinematicChain.h:88-176 derives from RealVectorStateSpace and keeps
            // its type id; DubinsStateSpace / ReedsSheppStateSpace derive from SE2StateSpace), so only the
            // exact classes are accepted; everything else needs an explicit SpaceDescription.
            template <typename Stock>
            inline bool exactly(const base::StateSpace *s)
            {
                return typeid(*s) == typeid(Stock);
            }
            inline void leaf(const base::StateSpace *s, double weight, SpaceDescription &d)
            {
                const double lvs = s->getLongestValidSegmentLength();
                const unsigned f = s->getValidSegmentCountFactor();
                if (!(exactly<base::RealVectorStateSpace>(s) || exactly<base::SO2StateSpace>(s) ||
                      exactly<base::SO3StateSpace>(s)))
                    throw Exception("ompl::b200::describe: state space '" + s->getName() +
                                    "' is not exactly RealVectorStateSpace, SO2StateSpace or SO3StateSpace; a subclass "
                                    "may override distance() or interpolate(), which the device cannot know. Pass an "
                                    "explicit SpaceDescription (e.g. SpaceDescription::kinematicChain) instead");
                switch (s->getType())
                {
                    case base::STATE_SPACE_REAL_VECTOR:
                        d.components.push_back(SpaceDescription::component(OMPLCU_RV, (int)s->getDimension(), weight, lvs, f));
                        break;
                    case base::STATE_SPACE_SO2:
                        d.components.push_back(SpaceDescription::component(OMPLCU_SO2, 1, weight, lvs, f));
                        break;
                    case base::STATE_SPACE_SO3:
                        d.components.push_back(SpaceDescription::component(OMPLCU_SO3, 4, weight, lvs, f));
                        break;
                    default:
                        throw Exception("ompl::b200::describe: state space '" + s->getName() +
                                        "' has no device metric (RealVector, SO2, SO3 and compounds of them do)");
                }
            }
        }  // namespace detail

        /** Device description of a state space after setup() (longest valid segments are read from it). */
        inline SpaceDescription describe(const base::StateSpacePtr &space)
        {
            SpaceDescription d;
            if (space->isCompound())
            {
                if (!(detail::exactly<base::CompoundStateSpace>(space.get()) ||
                      detail::exactly<base::SE2StateSpace>(space.get()) ||
                      detail::exactly<base::SE3StateSpace>(space.get())))
                    throw Exception("ompl::b200::describe: compound state space '" + space->getName() +
                                    "' is not exactly CompoundStateSpace, SE2StateSpace or SE3StateSpace (Dubins, "
                                    "Reeds-Shepp and other subclasses override distance() / interpolate()); pass an "
                                    "explicit SpaceDescription if the device metric really is the stock one");
                // CompoundStateSpace::distance is the weighted SUM of component distances
                // (StateSpace.cpp:1068-1076); nested compounds are not flattened because
                // validSegmentCount takes the max per DIRECT component (:1085-1097).
                const auto *c = space->as<base::CompoundStateSpace>();
                d.compound = true;
                for (unsigned int i = 0; i < c->getSubspaceCount(); ++i)
                {
                    if (c->getSubspace(i)->isCompound())
                        throw Exception("ompl::b200::describe: nested compound state spaces are not supported");
                    detail::leaf(c->getSubspace(i).get(), c->getSubspaceWeight(i), d);
                }
                if (d.components.size() > OMPLCU_MAX_COMPONENTS)
                    throw Exception("ompl::b200::describe: more than 8 subspaces");
            }
            else
                detail::leaf(space.get(), 1.0, d);
            return d;
        }

        /** Row writer for elements whose state is reachable through `accessor`. */
        template <typename _T>
        typename NearestNeighborsCUDA<_T>::RowFunction rowsFrom(const base::StateSpacePtr &space,
                                                                std::function<const base::State *(const _T &)> accessor)
        {
            return [space, accessor](const _T &e, double *row) {
                thread_local std::vector<double> reals;
                space->copyToReals(reals, accessor(e));
                std::copy(reals.begin(), reals.end(), row);
            };
        }

        /** setStateSpace + setStateAccessor of SURVEY 8(b). */
        template <typename _T>
        void bind(NearestNeighborsCUDA<_T> &nn, const base::StateSpacePtr &space,
                  std::function<const base::State *(const _T &)> accessor)
        {
            nn.setSpace(describe(space));
            nn.setRowFunction(rowsFrom<_T>(space, std::move(accessor)));
        }

        /** The same with a description supplied by the caller, for spaces describe() refuses: e.g.
            bind(nn, chainSpace, accessor, SpaceDescription::kinematicChain(links, linkLength)). */
        template <typename _T>
        void bind(NearestNeighborsCUDA<_T> &nn, const base::StateSpacePtr &space,
                  std::function<const base::State *(const _T &)> accessor, const SpaceDescription &description)
        {
            if (description.stride() != space->getDimension() && !space->isCompound())
                throw Exception("ompl::b200::bind: the description's stride differs from the space's dimension");
            nn.setSpace(description);
            nn.setRowFunction(rowsFrom<_T>(space, std::move(accessor)));
        }

        /** KinematicChainSpace (demos/KinematicChain.h:88) described from the live space: links = dimension,
            longest valid segment and count factor read from it after setup(). */
        inline SpaceDescription describeKinematicChain(const base::StateSpacePtr &space, double linkLength)
        {
            return SpaceDescription::kinematicChain((int)space->getDimension(), linkLength,
                                                    space->getLongestValidSegmentLength(),
                                                    space->getValidSegmentCountFactor());
        }

        /** Owns an omplcu_checker_t. */
        class CheckerHandle
        {
        public:
            explicit CheckerHandle(omplcu_checker_t h) : h_(h)
            {
            }
            CheckerHandle(const CheckerHandle &) = delete;
            CheckerHandle &operator=(const CheckerHandle &) = delete;
            ~CheckerHandle()
            {
                omplcu_checker_destroy(h_);
            }
            omplcu_checker_t get() const
            {
                return h_;
            }

            static std::shared_ptr<CheckerHandle> hypercube(int dim, double edgeWidth = 0.1, int device = 0)
            {
                omplcu_checker_t h = nullptr;
                check(omplcu_checker_create_hypercube(device, dim, edgeWidth, &h));
                return std::make_shared<CheckerHandle>(h);
            }
            /** env: segments as (x0,y0,x1,y1) rows, demos/KinematicChain.h:279-318 */
            static std::shared_ptr<CheckerHandle> kinematicChain(int links, double linkLength,
                                                                  const std::vector<double> &env, int device = 0)
            {
                omplcu_checker_t h = nullptr;
                check(omplcu_checker_create_chain(device, links, linkLength, env.data(), (int)(env.size() / 4), &h));
                return std::make_shared<CheckerHandle>(h);
            }
            static std::shared_ptr<CheckerHandle> se3Obstacles(const std::vector<double> &robot,
                                                                const std::vector<double> &spheres,
                                                                const std::vector<double> &boxes, int device = 0)
            {
                omplcu_checker_t h = nullptr;
                check(omplcu_checker_create_se3(device, robot.data(), (int)(robot.size() / 4), spheres.data(),
                                                (int)(spheres.size() / 4), boxes.data(), (int)(boxes.size() / 6), &h));
                return std::make_shared<CheckerHandle>(h);
            }
            static std::shared_ptr<CheckerHandle> circles2D(const std::vector<double> &circles, int device = 0)
            {
                omplcu_checker_t h = nullptr;
                check(omplcu_checker_create_circles(device, circles.data(), (int)(circles.size() / 3), &h));
                return std::make_shared<CheckerHandle>(h);
            }
            static std::shared_ptr<CheckerHandle> allValid(int device = 0)
            {
                omplcu_checker_t h = nullptr;
                check(omplcu_checker_create_allvalid(device, &h));
                return std::make_shared<CheckerHandle>(h);
            }

        private:
            omplcu_checker_t h_;
        };
        using CheckerHandlePtr = std::shared_ptr<CheckerHandle>;

        inline void gatherRows(const base::StateSpace *space, const base::State *const *states, std::size_t n,
                               std::size_t stride, std::vector<double> &rows)
        {
            rows.resize(n * stride);
            std::vector<double> reals;
            for (std::size_t i = 0; i < n; ++i)
            {
                space->copyToReals(reals, states[i]);
                std::copy(reals.begin(), reals.end(), rows.begin() + i * stride);
            }
        }
    }  // namespace b200

    namespace base
    {
        /** StateValidityChecker whose predicate runs on the device.  Install with
            si->setStateValidityChecker(ptr) (SpaceInformation.h:146-153). */
        class BatchedStateValidityChecker : public StateValidityChecker
        {
        public:
            BatchedStateValidityChecker(const SpaceInformationPtr &si, b200::CheckerHandlePtr checker)
              : StateValidityChecker(si), checker_(std::move(checker))
            {
            }

            bool isValid(const State *state) const override
            {
                std::uint8_t v = 0;
                isValidBatch(&state, 1, &v);
                return v != 0;
            }

            /** valid[i] = isValid(states[i]) for the whole batch in one launch. */
            void isValidBatch(const State *const *states, std::size_t n, std::uint8_t *valid) const
            {
                if (n == 0)
                    return;
                const std::size_t stride = si_->getStateSpace()->getDimension() == 0 ? 0 : rowStride();
                std::vector<double> rows;
                b200::gatherRows(si_->getStateSpace().get(), states, n, stride, rows);
                b200::check(omplcu_check_states(checker_->get(), (int)stride, rows.data(), n, valid));
            }

            const b200::CheckerHandlePtr &checker() const
            {
                return checker_;
            }

        private:
            std::size_t rowStride() const
            {
                std::vector<double> reals;
                State *s = si_->allocState();
                si_->getStateSpace()->copyToReals(reals, s);
                si_->freeState(s);
                return reals.size();
            }
            b200::CheckerHandlePtr checker_;
        };

        /** DiscreteMotionValidator on the device.  Install with si->setMotionValidator(ptr)
            (SpaceInformation.h:165-172) AFTER si->setup() fixed the longest valid segments. */
        class BatchedDiscreteMotionValidator : public MotionValidator
        {
        public:
            BatchedDiscreteMotionValidator(const SpaceInformationPtr &si, b200::CheckerHandlePtr checker)
              : MotionValidator(si), checker_(std::move(checker)), space_(b200::describe(si->getStateSpace()))
            {
            }
            /** With an explicit description (spaces describe() refuses, e.g. the kinematic chain). */
            BatchedDiscreteMotionValidator(const SpaceInformationPtr &si, b200::CheckerHandlePtr checker,
                                           const b200::SpaceDescription &description)
              : MotionValidator(si), checker_(std::move(checker)), space_(description)
            {
            }
            /** One checker per GPU holding the same obstacle set: checkMotionBatch splits its segments over
                them (omplcu_check_motions_multi). */
            void setCheckers(std::vector<b200::CheckerHandlePtr> perDevice)
            {
                multi_ = std::move(perDevice);
                multiRaw_.clear();
                for (const auto &c : multi_)
                    multiRaw_.push_back(c->get());
            }

            bool checkMotion(const State *s1, const State *s2) const override
            {
                std::uint8_t v = 0;
                checkMotionBatch(&s1, &s2, 1, &v, nullptr);
                return v != 0;
            }

            bool checkMotion(const State *s1, const State *s2, std::pair<State *, double> &lastValid) const override
            {
                std::uint8_t v = 0;
                double t = 1.0;
                std::vector<double> st(space_.stride());
                const std::size_t stride = space_.stride();
                std::vector<double> a, b;
                b200::gatherRows(si_->getStateSpace().get(), &s1, 1, stride, a);
                b200::gatherRows(si_->getStateSpace().get(), &s2, 1, stride, b);
                b200::check(omplcu_check_motions(checker_->get(), space_.get(), a.data(), b.data(), 1, &v, &t,
                                                 lastValid.first != nullptr ? st.data() : nullptr));
                count(v != 0, 1);
                if (v == 0)
                {
                    lastValid.second = t;
                    if (lastValid.first != nullptr)
                        si_->getStateSpace()->copyFromReals(lastValid.first, st);
                }
                return v != 0;
            }

            /** valid[i] = checkMotion(s1[i], s2[i]); lastT (optional) = lastValid.second on failure, 1 on success. */
            void checkMotionBatch(const State *const *s1, const State *const *s2, std::size_t n, std::uint8_t *valid,
                                  double *lastT) const
            {
                if (n == 0)
                    return;
                const std::size_t stride = space_.stride();
                std::vector<double> a, b;
                b200::gatherRows(si_->getStateSpace().get(), s1, n, stride, a);
                b200::gatherRows(si_->getStateSpace().get(), s2, n, stride, b);
                if (multiRaw_.size() > 1)
                    b200::check(omplcu_check_motions_multi(multiRaw_.data(), (int)multiRaw_.size(), space_.get(),
                                                           a.data(), b.data(), n, valid, lastT, nullptr));
                else
                    b200::check(omplcu_check_motions(checker_->get(), space_.get(), a.data(), b.data(), n, valid,
                                                     lastT, nullptr));
                std::size_t nv = 0;
                for (std::size_t i = 0; i < n; ++i)
                    nv += valid[i] != 0;
                count(true, nv);
                count(false, n - nv);
            }

        private:
            // valid_/invalid_ are plain mutable unsigned in the reference (MotionValidator.h:136-139);
            // one increment per segment as DiscreteMotionValidator.cpp:139-142
            void count(bool ok, std::size_t k) const
            {
                std::lock_guard<std::mutex> lk(mu_);
                (ok ? valid_ : invalid_) += (unsigned int)k;
            }
            b200::CheckerHandlePtr checker_;
            std::vector<b200::CheckerHandlePtr> multi_;
            std::vector<omplcu_checker_t> multiRaw_;
            b200::SpaceHandle space_;
            mutable std::mutex mu_;
        };
    }  // namespace base
}  // namespace ompl

#endif
