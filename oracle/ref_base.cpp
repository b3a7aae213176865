// ref_base.cpp -- C wrapper around the reference's OWN state-space and
// DiscreteMotionValidator code.  TEST INFRASTRUCTURE ONLY.
//
// Compiled together with these files, taken UNMODIFIED from /root/reference/src/ompl
// (include path + a handful of stand-in headers under oracle/shims for the Boost /
// Eigen headers they name; nothing is copied into this repository):
//   base/src/StateSpace.cpp                 CompoundStateSpace distance / interpolate /
//                                           validSegmentCount, StateSpace::setup
//   base/spaces/src/RealVectorStateSpace.cpp, RealVectorBounds.cpp
//   base/spaces/src/SO2StateSpace.cpp, SO3StateSpace.cpp, SE2StateSpace.cpp, SE3StateSpace.cpp
//   base/src/DiscreteMotionValidator.cpp    both checkMotion overloads
//   base/src/StateSampler.cpp, GenericParam.cpp, util/src/Console.cpp, String.cpp (link closure)
//
// What is NOT the reference's: (1) the definitions in the "link stubs" section
// below -- members of ProjectionEvaluator / RealVector*ProjectionEvaluator /
// SpaceInformation whose real sources need Eigen's SVD or the whole planner tree.
// None of them does arithmetic on the hot path: projections are never evaluated,
// and SpaceInformation only hands DiscreteMotionValidator its state space and
// validity checker through the header's inline accessors.  (2) the validity
// predicate, which is the oracle's restatement of the demo checkers (the demos
// are executables with Boost.ProgramOptions, not library code).
//
// So: distance / interpolate / validSegmentCount / longest valid segment and the
// DiscreteMotionValidator control flow are the reference's own machine code.
#include "ompl/base/DiscreteMotionValidator.h"
#include "ompl/base/SpaceInformation.h"
#include "ompl/base/spaces/RealVectorStateProjections.h"
#include "ompl/base/spaces/RealVectorStateSpace.h"
#include "ompl/base/spaces/SE2StateSpace.h"
#include "ompl/base/spaces/SE3StateSpace.h"
#include "ompl/base/spaces/SO2StateSpace.h"
#include "ompl/base/spaces/SO3StateSpace.h"
#include "ompl_oracle.h"

#include <cstring>
#include <thread>

namespace ob = ompl::base;

// ---------------------------------------------------------------------------
// link stubs (never executed on the path; see header comment)
// ---------------------------------------------------------------------------
static void off_path(const char *what)
{
    throw ompl::Exception(std::string("oracle/ref_base.cpp stub reached: ") + what);
}

// util/src/RandomNumbers.cpp needs Boost.Random; samplers are not on the path
ompl::RNG::RNG() : localSeed_(1), generator_(1)
{
}
void ompl::RNG::quaternion(double *)
{
    off_path("RNG::quaternion");
}

ob::ProjectionEvaluator::ProjectionEvaluator(const StateSpace *space)
  : space_(space), bounds_(0), estimatedBounds_(0), defaultCellSizes_(true), cellSizesWereInferred_(false)
{
}
ob::ProjectionEvaluator::ProjectionEvaluator(const StateSpacePtr &space)
  : space_(space.get()), bounds_(0), estimatedBounds_(0), defaultCellSizes_(true), cellSizesWereInferred_(false)
{
}
ob::ProjectionEvaluator::~ProjectionEvaluator() = default;
bool ob::ProjectionEvaluator::userConfigured() const
{
    return false;
}
void ob::ProjectionEvaluator::setCellSizes(const std::vector<double> &)
{
}
void ob::ProjectionEvaluator::defaultCellSizes()
{
}
void ob::ProjectionEvaluator::setup()
{
}
void ob::ProjectionEvaluator::printSettings(std::ostream &) const
{
}
void ob::ProjectionEvaluator::printProjection(const Eigen::Ref<Eigen::VectorXd> &, std::ostream &) const
{
}
ob::ProjectionMatrix::Matrix ob::ProjectionMatrix::ComputeRandom(unsigned int from, unsigned int to,
                                                                 const std::vector<double> &)
{
    return Matrix(to, from);
}
ob::RealVectorLinearProjectionEvaluator::RealVectorLinearProjectionEvaluator(const StateSpace *space,
                                                                             const ProjectionMatrix::Matrix &)
  : ProjectionEvaluator(space)
{
}
unsigned int ob::RealVectorLinearProjectionEvaluator::getDimension() const
{
    return 0;
}
void ob::RealVectorLinearProjectionEvaluator::project(const State *, Eigen::Ref<Eigen::VectorXd>) const
{
    off_path("RealVectorLinearProjectionEvaluator::project");
}
ob::RealVectorIdentityProjectionEvaluator::RealVectorIdentityProjectionEvaluator(const StateSpace *space)
  : ProjectionEvaluator(space)
{
}
unsigned int ob::RealVectorIdentityProjectionEvaluator::getDimension() const
{
    return 0;
}
void ob::RealVectorIdentityProjectionEvaluator::defaultCellSizes()
{
}
void ob::RealVectorIdentityProjectionEvaluator::setup()
{
}
void ob::RealVectorIdentityProjectionEvaluator::project(const State *, Eigen::Ref<Eigen::VectorXd>) const
{
    off_path("RealVectorIdentityProjectionEvaluator::project");
}

// SpaceInformation: a holder of (state space, validity checker); everything
// DiscreteMotionValidator touches is inline in the reference header.
ob::SpaceInformation::SpaceInformation(StateSpacePtr space) : stateSpace_(std::move(space)), setup_(false)
{
}
void ob::SpaceInformation::setup()
{
    setup_ = true;
}
unsigned int ob::SpaceInformation::getMotionStates(const State *, const State *, std::vector<State *> &, unsigned int,
                                                   bool, bool) const
{
    off_path("SpaceInformation::getMotionStates");
    return 0;
}
void ob::SpaceInformation::printSettings(std::ostream &) const
{
}
void ob::SpaceInformation::printProperties(std::ostream &) const
{
}

// ---------------------------------------------------------------------------
// wrapper
// ---------------------------------------------------------------------------
namespace
{
    // the oracle's predicate behind the reference's StateValidityChecker interface
    class OracleChecker : public ob::StateValidityChecker
    {
    public:
        OracleChecker(ob::SpaceInformation *si, const orc_checker *c) : ob::StateValidityChecker(si), c_(c)
        {
        }
        bool isValid(const ob::State *s) const override
        {
            std::vector<double> r;
            si_->getStateSpace()->copyToReals(r, s);
            return orc_is_valid(c_, r.data()) != 0;
        }

    private:
        const orc_checker *c_;
    };

    struct RefBase
    {
        ob::StateSpacePtr space;
        std::shared_ptr<ob::SpaceInformation> si;
        unsigned stride{0};
    };

    struct Scoped
    {
        const ob::StateSpacePtr &sp;
        ob::State *s;
        Scoped(const ob::StateSpacePtr &p, const double *row, unsigned n) : sp(p), s(p->allocState())
        {
            std::vector<double> r(row, row + n);
            sp->copyFromReals(s, r);
        }
        ~Scoped()
        {
            sp->freeState(s);
        }
    };
}  // namespace

extern "C" {

// kind: 0 RealVector(dim, low, high), 1 SO2, 2 SO3, 3 SE2 (low/high: 2), 4 SE3 (low/high: 3)
void *refb_create(int kind, int dim, const double *low, const double *high, double fraction, unsigned factor)
{
    auto *h = new RefBase;
    auto bounds = [&](int d) {
        ob::RealVectorBounds b(d);
        for (int i = 0; i < d; ++i)
        {
            b.setLow(i, low[i]);
            b.setHigh(i, high[i]);
        }
        return b;
    };
    if (kind == 0)
    {
        auto s = std::make_shared<ob::RealVectorStateSpace>(dim);
        s->setBounds(bounds(dim));
        h->space = s;
    }
    else if (kind == 1)
        h->space = std::make_shared<ob::SO2StateSpace>();
    else if (kind == 2)
        h->space = std::make_shared<ob::SO3StateSpace>();
    else if (kind == 3)
    {
        auto s = std::make_shared<ob::SE2StateSpace>();
        s->setBounds(bounds(2));
        h->space = s;
    }
    else
    {
        auto s = std::make_shared<ob::SE3StateSpace>();
        s->setBounds(bounds(3));
        h->space = s;
    }
    h->space->setLongestValidSegmentFraction(fraction);  // SpaceInformation::setStateValidityCheckingResolution
    h->space->setValidSegmentCountFactor(factor);
    h->space->setup();
    h->si = std::make_shared<ob::SpaceInformation>(h->space);
    std::vector<double> r;
    ob::State *tmp = h->space->allocState();
    h->space->copyToReals(r, tmp);
    h->space->freeState(tmp);
    h->stride = (unsigned)r.size();
    return h;
}

void refb_destroy(void *hv)
{
    delete static_cast<RefBase *>(hv);
}

unsigned refb_stride(void *hv)
{
    return static_cast<RefBase *>(hv)->stride;
}

double refb_max_extent(void *hv)
{
    return static_cast<RefBase *>(hv)->space->getMaximumExtent();
}

double refb_longest_valid_segment(void *hv)
{
    return static_cast<RefBase *>(hv)->space->getLongestValidSegmentLength();
}

void refb_distance(void *hv, const double *a, const double *b, size_t n, double *out)
{
    auto *h = static_cast<RefBase *>(hv);
    for (size_t i = 0; i < n; ++i)
    {
        Scoped x(h->space, a + i * h->stride, h->stride), y(h->space, b + i * h->stride, h->stride);
        out[i] = h->space->distance(x.s, y.s);
    }
}

void refb_interpolate(void *hv, const double *a, const double *b, const double *t, size_t n, double *out)
{
    auto *h = static_cast<RefBase *>(hv);
    std::vector<double> r;
    for (size_t i = 0; i < n; ++i)
    {
        Scoped x(h->space, a + i * h->stride, h->stride), y(h->space, b + i * h->stride, h->stride),
            z(h->space, a + i * h->stride, h->stride);
        h->space->interpolate(x.s, y.s, t[i], z.s);
        h->space->copyToReals(r, z.s);
        std::memcpy(out + i * h->stride, r.data(), h->stride * sizeof(double));
    }
}

void refb_segment_count(void *hv, const double *a, const double *b, size_t n, unsigned *out)
{
    auto *h = static_cast<RefBase *>(hv);
    for (size_t i = 0; i < n; ++i)
    {
        Scoped x(h->space, a + i * h->stride, h->stride), y(h->space, b + i * h->stride, h->stride);
        out[i] = h->space->validSegmentCount(x.s, y.s);
    }
}

// The reference DiscreteMotionValidator over a batch.  last_t / last_state may be
// NULL (then the two-argument overload runs).  counts[0..1] = valid_, invalid_.
void refb_check_motions(void *hv, const orc_checker *chk, const double *s1, const double *s2, size_t n,
                        unsigned char *valid, double *last_t, double *last_state, unsigned *counts)
{
    auto *h = static_cast<RefBase *>(hv);
    h->si->setStateValidityChecker(std::make_shared<OracleChecker>(h->si.get(), chk));
    ob::DiscreteMotionValidator mv(h->si.get());
    std::vector<double> r;
    for (size_t i = 0; i < n; ++i)
    {
        Scoped x(h->space, s1 + i * h->stride, h->stride), y(h->space, s2 + i * h->stride, h->stride);
        if (last_t == nullptr)
            valid[i] = mv.checkMotion(x.s, y.s) ? 1 : 0;
        else
        {
            Scoped z(h->space, s1 + i * h->stride, h->stride);
            std::pair<ob::State *, double> lv(last_state != nullptr ? z.s : nullptr, 1.0);
            valid[i] = mv.checkMotion(x.s, y.s, lv) ? 1 : 0;
            last_t[i] = valid[i] ? 1.0 : lv.second;
            if (last_state != nullptr && !valid[i])
            {
                h->space->copyToReals(r, z.s);
                std::memcpy(last_state + i * h->stride, r.data(), h->stride * sizeof(double));
            }
        }
    }
    if (counts != nullptr)
    {
        counts[0] = mv.getValidMotionCount();
        counts[1] = mv.getInvalidMotionCount();
    }
}
}
