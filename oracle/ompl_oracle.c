/*
 * ompl_oracle.c -- CPU restatement of the OMPL 1.6.0 hot path.
 * TEST INFRASTRUCTURE ONLY (see ompl_oracle.h).  Every function cites the
 * reference file:line (relative to /root/reference) it follows operation by
 * operation.  Compile with -O3 -ffp-contract=off: the reference x86-64 build
 * emits no FMA, so each multiply and add below rounds separately.
 */
#include "ompl_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* boost::math::constants::pi<double>() == nearest double to pi */
static const double ORC_PI = 3.14159265358979323846;
static const double MAX_QUATERNION_NORM_ERROR = 1e-9; /* SO3StateSpace.cpp:47 */

/* ------------------------------------------------------------------ */
/* space construction                                                  */
/* ------------------------------------------------------------------ */

/* RealVectorStateSpace::getMaximumExtent, RealVectorStateSpace.cpp:167-176 */
double orc_rv_max_extent(const double *low, const double *high, int d)
{
    double e = 0.0;
    for (int i = 0; i < d; ++i)
    {
        double df = high[i] - low[i];
        e += df * df;
    }
    return sqrt(e);
}

static void comp_set(orc_component *c, int type, int dim, double weight, double extent, double fraction,
                     double param)
{
    c->type = type;
    c->dim = dim;
    c->weight = weight;
    /* StateSpace::setup, StateSpace.cpp:237-240: longestValidSegment_ = maxExtent_ * fraction */
    c->lvs = extent * fraction;
    c->factor = 1; /* StateSpace.cpp:94 */
    c->param = param;
}

void orc_space_rv(orc_space *sp, int d, const double *low, const double *high, double fraction)
{
    memset(sp, 0, sizeof(*sp));
    sp->ncomp = 1;
    sp->compound = 0;
    sp->stride = d;
    comp_set(&sp->comp[0], ORC_RV, d, 1.0, orc_rv_max_extent(low, high, d), fraction, 0.0);
}

void orc_space_so2(orc_space *sp, double fraction)
{
    memset(sp, 0, sizeof(*sp));
    sp->ncomp = 1;
    sp->stride = 1;
    comp_set(&sp->comp[0], ORC_SO2, 1, 1.0, ORC_PI, fraction, 0.0); /* SO2StateSpace.cpp:70-73 */
}

void orc_space_so3(orc_space *sp, double fraction)
{
    memset(sp, 0, sizeof(*sp));
    sp->ncomp = 1;
    sp->stride = 4;
    comp_set(&sp->comp[0], ORC_SO3, 4, 1.0, .5 * ORC_PI, fraction, 0.0); /* SO3StateSpace.cpp:166-169 */
}

/* SE2 = R^2 (w 1.0) + SO2 (w 0.5): SE2StateSpace.h:108-112.  Each component's
 * longestValidSegment_ comes from its OWN extent (CompoundStateSpace::setup,
 * StateSpace.cpp:1217-1223; fraction propagated by :1078-1083). */
void orc_space_se2(orc_space *sp, const double *low2, const double *high2, double fraction)
{
    memset(sp, 0, sizeof(*sp));
    sp->ncomp = 2;
    sp->compound = 1;
    sp->stride = 3;
    comp_set(&sp->comp[0], ORC_RV, 2, 1.0, orc_rv_max_extent(low2, high2, 2), fraction, 0.0);
    comp_set(&sp->comp[1], ORC_SO2, 1, 0.5, ORC_PI, fraction, 0.0);
}

/* SE3 = R^3 (w 1.0) + SO3 (w 1.0): SE3StateSpace.h constructor */
void orc_space_se3(orc_space *sp, const double *low3, const double *high3, double fraction)
{
    memset(sp, 0, sizeof(*sp));
    sp->ncomp = 2;
    sp->compound = 1;
    sp->stride = 7;
    comp_set(&sp->comp[0], ORC_RV, 3, 1.0, orc_rv_max_extent(low3, high3, 3), fraction, 0.0);
    comp_set(&sp->comp[1], ORC_SO3, 4, 1.0, .5 * ORC_PI, fraction, 0.0);
}

/* KinematicChainSpace: a RealVectorStateSpace(links) with bounds [-pi,pi]
 * (demos/KinematicChain.h:91-99); extent is the RealVector one. */
void orc_space_chain(orc_space *sp, int links, double link_length, double fraction)
{
    memset(sp, 0, sizeof(*sp));
    sp->ncomp = 1;
    sp->stride = links;
    double e = 0.0;
    for (int i = 0; i < links; ++i)
    {
        double df = ORC_PI - (-ORC_PI);
        e += df * df;
    }
    comp_set(&sp->comp[0], ORC_CHAIN, links, 1.0, sqrt(e), fraction, link_length);
}

/* ------------------------------------------------------------------ */
/* per-component math                                                  */
/* ------------------------------------------------------------------ */

/* RealVectorStateSpace::distance, RealVectorStateSpace.cpp:230-242 */
static double rv_distance(const double *s1, const double *s2, int d)
{
    double dist = 0.0;
    for (int i = 0; i < d; ++i)
    {
        double diff = s1[i] - s2[i];
        dist += diff * diff;
    }
    return sqrt(dist);
}

/* RealVectorStateSpace::interpolate, RealVectorStateSpace.cpp:257-265 */
static void rv_interpolate(const double *from, const double *to, double t, double *out, int d)
{
    for (int i = 0; i < d; ++i)
        out[i] = from[i] + (to[i] - from[i]) * t;
}

/* SO2StateSpace::distance, SO2StateSpace.cpp:116-128 */
static double so2_distance(double a, double b)
{
    double d = fabs(a - b);
    return (d > ORC_PI) ? 2.0 * ORC_PI - d : d;
}

/* SO2StateSpace::interpolate, SO2StateSpace.cpp:136-155 */
static double so2_interpolate(double from, double to, double t)
{
    double diff = to - from;
    if (fabs(diff) <= ORC_PI)
        return from + diff * t;
    double v;
    if (diff > 0.0)
        diff = 2.0 * ORC_PI - diff;
    else
        diff = -2.0 * ORC_PI - diff;
    v = from - diff * t;
    if (v > ORC_PI)
        v -= 2.0 * ORC_PI;
    else if (v < -ORC_PI)
        v += 2.0 * ORC_PI;
    return v;
}

/* arcLength, SO3StateSpace.cpp:254-262; storage order x,y,z,w (:371-374) */
static double so3_arc_length(const double *q1, const double *q2)
{
    double dq = fabs(q1[0] * q2[0] + q1[1] * q2[1] + q1[2] * q2[2] + q1[3] * q2[3]);
    if (dq > 1.0 - MAX_QUATERNION_NORM_ERROR)
        return 0.0;
    return acos(dq);
}

/* SO3StateSpace::interpolate (slerp), SO3StateSpace.cpp:289-318 */
static void so3_interpolate(const double *from, const double *to, double t, double *out)
{
    double theta = so3_arc_length(from, to);
    if (theta > DBL_EPSILON)
    {
        double d = 1.0 / sin(theta);
        double s0 = sin((1.0 - t) * theta);
        double s1 = sin(t * theta);
        double dq = from[0] * to[0] + from[1] * to[1] + from[2] * to[2] + from[3] * to[3];
        if (dq < 0)
            s1 = -s1;
        double x = (from[0] * s0 + to[0] * s1) * d;
        double y = (from[1] * s0 + to[1] * s1) * d;
        double z = (from[2] * s0 + to[2] * s1) * d;
        double w = (from[3] * s0 + to[3] * s1) * d;
        out[0] = x;
        out[1] = y;
        out[2] = z;
        out[3] = w;
    }
    else if (out != from)
    {
        out[0] = from[0];
        out[1] = from[1];
        out[2] = from[2];
        out[3] = from[3];
    }
}

/* KinematicChainSpace::distance, demos/KinematicChain.h:105-121 */
static double chain_distance(const double *s1, const double *s2, int d, double link_length)
{
    double theta1 = 0., theta2 = 0., dx = 0., dy = 0., dist = 0.;
    for (int i = 0; i < d; ++i)
    {
        theta1 += s1[i];
        theta2 += s2[i];
        dx += cos(theta1) - cos(theta2);
        dy += sin(theta1) - sin(theta2);
        dist += sqrt(dx * dx + dy * dy);
    }
    return dist * link_length;
}

/* KinematicChainSpace::interpolate, demos/KinematicChain.h:150-176 */
static void chain_interpolate(const double *from, const double *to, double t, double *out, int d)
{
    for (int i = 0; i < d; ++i)
    {
        double diff = to[i] - from[i];
        if (fabs(diff) <= ORC_PI)
            out[i] = from[i] + diff * t;
        else
        {
            if (diff > 0.0)
                diff = 2.0 * ORC_PI - diff;
            else
                diff = -2.0 * ORC_PI - diff;
            double v = from[i] - diff * t;
            if (v > ORC_PI)
                v -= 2.0 * ORC_PI;
            else if (v < -ORC_PI)
                v += 2.0 * ORC_PI;
            out[i] = v;
        }
    }
}

static double comp_distance(const orc_component *c, const double *a, const double *b)
{
    switch (c->type)
    {
        case ORC_RV:
            return rv_distance(a, b, c->dim);
        case ORC_SO2:
            return so2_distance(a[0], b[0]);
        case ORC_SO3:
            return so3_arc_length(a, b);
        default:
            return chain_distance(a, b, c->dim, c->param);
    }
}

static void comp_interpolate(const orc_component *c, const double *a, const double *b, double t, double *out)
{
    switch (c->type)
    {
        case ORC_RV:
            rv_interpolate(a, b, t, out, c->dim);
            break;
        case ORC_SO2:
            out[0] = so2_interpolate(a[0], b[0], t);
            break;
        case ORC_SO3:
            so3_interpolate(a, b, t, out);
            break;
        default:
            chain_interpolate(a, b, t, out, c->dim);
            break;
    }
}

/* StateSpace::validSegmentCount, StateSpace.cpp:851-854 */
static unsigned comp_segment_count(const orc_component *c, const double *a, const double *b)
{
    return c->factor * (unsigned int)ceil(comp_distance(c, a, b) / c->lvs);
}

/* ------------------------------------------------------------------ */
/* whole-space math                                                    */
/* ------------------------------------------------------------------ */

/* CompoundStateSpace::distance, StateSpace.cpp:1068-1076 (sum from 0.0 in
 * component order); a plain space returns its own distance unweighted. */
double orc_distance(const orc_space *sp, const double *a, const double *b)
{
    if (!sp->compound)
        return comp_distance(&sp->comp[0], a, b);
    double dist = 0.0;
    int off = 0;
    for (int i = 0; i < sp->ncomp; ++i)
    {
        dist += sp->comp[i].weight * comp_distance(&sp->comp[i], a + off, b + off);
        off += sp->comp[i].dim;
    }
    return dist;
}

/* CompoundStateSpace::interpolate, StateSpace.cpp:1109-1116 */
void orc_interpolate(const orc_space *sp, const double *from, const double *to, double t, double *out)
{
    int off = 0;
    for (int i = 0; i < sp->ncomp; ++i)
    {
        comp_interpolate(&sp->comp[i], from + off, to + off, t, out + off);
        off += sp->comp[i].dim;
    }
}

/* CompoundStateSpace::validSegmentCount, StateSpace.cpp:1085-1097: the max
 * over components, each with its own longestValidSegment_ and factor. */
unsigned orc_valid_segment_count(const orc_space *sp, const double *a, const double *b)
{
    if (!sp->compound)
        return comp_segment_count(&sp->comp[0], a, b);
    unsigned sc = 0;
    int off = 0;
    for (int i = 0; i < sp->ncomp; ++i)
    {
        unsigned sci = comp_segment_count(&sp->comp[i], a + off, b + off);
        if (sci > sc)
            sc = sci;
        off += sp->comp[i].dim;
    }
    return sc;
}

/* ------------------------------------------------------------------ */
/* validity checkers                                                   */
/* ------------------------------------------------------------------ */

/* isStateValid, demos/HypercubeBenchmark.cpp:57-72 */
static int hypercube_valid(const double *s, int ndim, double edgeWidth)
{
    int foundMaxDim = 0;
    for (int i = ndim - 1; i >= 0; i--)
        if (!foundMaxDim)
        {
            if (s[i] > edgeWidth)
                foundMaxDim = 1;
        }
        else if (s[i] < (1. - edgeWidth))
            return 0;
    return 1;
}

/* KinematicChainValidityChecker::intersectionTest, demos/KinematicChain.h:251-276.
 * Segment layout (x0,y0,x1,y1). */
static int seg_intersect(const double *s0, const double *s1)
{
    double s10_x = s0[2] - s0[0];
    double s10_y = s0[3] - s0[1];
    double s32_x = s1[2] - s1[0];
    double s32_y = s1[3] - s1[1];
    double denom = s10_x * s32_y - s32_x * s10_y;
    if (fabs(denom) < DBL_EPSILON)
        return 0; /* collinear */
    int denomPositive = denom > 0;

    double s02_x = s0[0] - s1[0];
    double s02_y = s0[1] - s1[1];
    double s_numer = s10_x * s02_y - s10_y * s02_x;
    if ((s_numer < (double)FLT_EPSILON) == denomPositive)
        return 0;
    double t_numer = s32_x * s02_y - s32_y * s02_x;
    if ((t_numer < (double)FLT_EPSILON) == denomPositive)
        return 0;
    if (((s_numer - denom > -(double)FLT_EPSILON) == denomPositive) ||
        ((t_numer - denom > (double)FLT_EPSILON) == denomPositive))
        return 0;
    return 1;
}

/* KinematicChainValidityChecker::isValidImpl, demos/KinematicChain.h:209-249 */
static int chain_valid(const orc_checker *c, const double *s)
{
    int n = c->nlinks;
    double *segs = (double *)malloc(sizeof(double) * 4 * (size_t)(n + 1));
    double linkLength = c->link_length;
    double theta = 0., x = 0., y = 0., xN, yN;
    for (int i = 0; i < n; ++i)
    {
        theta += s[i];
        xN = x + cos(theta) * linkLength;
        yN = y + sin(theta) * linkLength;
        segs[4 * i + 0] = x;
        segs[4 * i + 1] = y;
        segs[4 * i + 2] = xN;
        segs[4 * i + 3] = yN;
        x = xN;
        y = yN;
    }
    xN = x + cos(theta) * 0.001;
    yN = y + sin(theta) * 0.001;
    segs[4 * n + 0] = x;
    segs[4 * n + 1] = y;
    segs[4 * n + 2] = xN;
    segs[4 * n + 3] = yN;
    int ok = 1;
    /* selfIntersectionTest :233-240 */
    for (int i = 0; ok && i < n + 1; ++i)
        for (int j = i + 1; j < n + 1; ++j)
            if (seg_intersect(segs + 4 * i, segs + 4 * j))
            {
                ok = 0;
                break;
            }
    /* environmentIntersectionTest :242-249 */
    for (int i = 0; ok && i < n + 1; ++i)
        for (int j = 0; j < c->nenv; ++j)
            if (seg_intersect(segs + 4 * i, c->env + 4 * j))
            {
                ok = 0;
                break;
            }
    free(segs);
    return ok;
}

/* SE3 rigid body vs analytic obstacles.  NOT in the reference (SURVEY.md 8a
 * row a21: demos/RigidBodyPlanning.cpp:48-62 is always true); specified here,
 * following the reference's overlap idiom Circles2D::noOverlap
 * (tests/resources/circles2D.h:139-149: obstacle minus point, strict '<').
 * State = (x,y,z, qx,qy,qz,qw).  Robot = body-fixed spheres (ox,oy,oz,r).
 * World centre c = p + R(q) o with R(q) o = o + w*t + q_v x t, t = 2 (q_v x o).
 * Sphere obstacle collides iff |s - c|^2 < (sr + r)^2; axis-aligned box
 * collides iff sum_i max(|c_i - b_i| - h_i, 0)^2 < r^2. */
static int se3_valid(const orc_checker *c, const double *st)
{
    const double px = st[0], py = st[1], pz = st[2];
    const double qx = st[3], qy = st[4], qz = st[5], qw = st[6];
    for (int i = 0; i < c->nrobot; ++i)
    {
        const double ox = c->robot[4 * i + 0], oy = c->robot[4 * i + 1], oz = c->robot[4 * i + 2];
        const double rr = c->robot[4 * i + 3];
        double tx = 2.0 * (qy * oz - qz * oy);
        double ty = 2.0 * (qz * ox - qx * oz);
        double tz = 2.0 * (qx * oy - qy * ox);
        double cx = px + ((ox + qw * tx) + (qy * tz - qz * ty));
        double cy = py + ((oy + qw * ty) + (qz * tx - qx * tz));
        double cz = pz + ((oz + qw * tz) + (qx * ty - qy * tx));
        for (int j = 0; j < c->nsph; ++j)
        {
            const double *s = c->sph + 4 * j;
            double dx = s[0] - cx;
            double dy = s[1] - cy;
            double dz = s[2] - cz;
            double R = s[3] + rr;
            if (dx * dx + dy * dy + dz * dz < R * R)
                return 0;
        }
        for (int j = 0; j < c->nbox; ++j)
        {
            const double *b = c->box + 6 * j;
            double ex = fabs(cx - b[0]) - b[3];
            double ey = fabs(cy - b[1]) - b[4];
            double ez = fabs(cz - b[2]) - b[5];
            if (ex < 0.0)
                ex = 0.0;
            if (ey < 0.0)
                ey = 0.0;
            if (ez < 0.0)
                ez = 0.0;
            if (ex * ex + ey * ey + ez * ez < rr * rr)
                return 0;
        }
    }
    return 1;
}

/* Circles2D::noOverlap, tests/resources/circles2D.h:139-149 (r2_ = r*r) */
static int circles_valid(const orc_checker *c, const double *s)
{
    for (int i = 0; i < c->ncirc; ++i)
    {
        double dx = c->circ[3 * i + 0] - s[0];
        double dy = c->circ[3 * i + 1] - s[1];
        double r2 = c->circ[3 * i + 2] * c->circ[3 * i + 2];
        if (dx * dx + dy * dy < r2)
            return 0;
    }
    return 1;
}

int orc_is_valid(const orc_checker *c, const double *state)
{
    switch (c->kind)
    {
        case ORC_CHK_HYPERCUBE:
            return hypercube_valid(state, c->dim, c->edge_width);
        case ORC_CHK_CHAIN:
            return chain_valid(c, state);
        case ORC_CHK_SE3:
            return se3_valid(c, state);
        case ORC_CHK_CIRCLES:
            return circles_valid(c, state);
        default:
            return 1; /* AllValidStateValidityChecker, StateValidityChecker.h:165-182 */
    }
}

/* ------------------------------------------------------------------ */
/* DiscreteMotionValidator                                             */
/* ------------------------------------------------------------------ */

/* DiscreteMotionValidator::checkMotion(s1,s2), DiscreteMotionValidator.cpp:93-145.
 * The reference uses a FIFO of (first,last) index ranges (breadth-first
 * bisection); reproduced with an array queue. */
int orc_check_motion(const orc_space *sp, const orc_checker *c, const double *s1, const double *s2)
{
    if (!orc_is_valid(c, s2))
        return 0;
    int result = 1;
    int nd = (int)orc_valid_segment_count(sp, s1, s2);
    if (nd >= 2)
    {
        int cap = nd + 2;
        int *qa = (int *)malloc(sizeof(int) * 2 * (size_t)cap);
        double *test = (double *)malloc(sizeof(double) * (size_t)sp->stride);
        int head = 0, tail = 0;
        qa[0] = 1;
        qa[1] = nd - 1;
        tail = 1;
        while (head < tail)
        {
            int first = qa[2 * head], second = qa[2 * head + 1];
            int mid = (first + second) / 2;
            orc_interpolate(sp, s1, s2, (double)mid / (double)nd, test);
            if (!orc_is_valid(c, test))
            {
                result = 0;
                break;
            }
            head++;
            if (first < mid)
            {
                qa[2 * tail] = first;
                qa[2 * tail + 1] = mid - 1;
                tail++;
            }
            if (second > mid)
            {
                qa[2 * tail] = mid + 1;
                qa[2 * tail + 1] = second;
                tail++;
            }
        }
        free(qa);
        free(test);
    }
    return result;
}

/* DiscreteMotionValidator::checkMotion(s1,s2,lastValid), DiscreteMotionValidator.cpp:48-91 */
int orc_check_motion_last(const orc_space *sp, const orc_checker *c, const double *s1, const double *s2,
                          double *last_t, double *last_state)
{
    int result = 1;
    int nd = (int)orc_valid_segment_count(sp, s1, s2);
    if (nd > 1)
    {
        double *test = (double *)malloc(sizeof(double) * (size_t)sp->stride);
        for (int j = 1; j < nd; ++j)
        {
            orc_interpolate(sp, s1, s2, (double)j / (double)nd, test);
            if (!orc_is_valid(c, test))
            {
                *last_t = (double)(j - 1) / (double)nd;
                if (last_state != NULL)
                    orc_interpolate(sp, s1, s2, *last_t, last_state);
                result = 0;
                break;
            }
        }
        free(test);
    }
    if (result)
        if (!orc_is_valid(c, s2))
        {
            *last_t = (double)(nd - 1) / (double)nd;
            if (last_state != NULL)
                orc_interpolate(sp, s1, s2, *last_t, last_state);
            result = 0;
        }
    return result;
}

/* ------------------------------------------------------------------ */
/* batched helpers (pthreads; used for the timed CPU baseline)         */
/* ------------------------------------------------------------------ */

typedef struct
{
    const orc_space *sp;
    const orc_checker *c;
    const double *a, *b;
    size_t lo, hi;
    int stride;
    uint8_t *out;
    double *last_t;
    /* knn */
    const double *pts, *q;
    size_t n, k;
    uint64_t *idx;
    double *dist;
} orc_job;

static void run_threads(void *(*fn)(void *), orc_job *proto, size_t n, int nthreads)
{
    if (nthreads < 1)
        nthreads = 1;
    if ((size_t)nthreads > n)
        nthreads = n ? (int)n : 1;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
    orc_job *jobs = (orc_job *)malloc(sizeof(orc_job) * (size_t)nthreads);
    for (int t = 0; t < nthreads; ++t)
    {
        jobs[t] = *proto;
        jobs[t].lo = n * (size_t)t / (size_t)nthreads;
        jobs[t].hi = n * (size_t)(t + 1) / (size_t)nthreads;
        if (nthreads == 1)
            fn(&jobs[t]);
        else
            pthread_create(&th[t], NULL, fn, &jobs[t]);
    }
    if (nthreads > 1)
        for (int t = 0; t < nthreads; ++t)
            pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}

static void *states_worker(void *p)
{
    orc_job *j = (orc_job *)p;
    for (size_t i = j->lo; i < j->hi; ++i)
        j->out[i] = (uint8_t)orc_is_valid(j->c, j->a + i * (size_t)j->stride);
    return NULL;
}

void orc_check_states_batch(const orc_checker *c, int stride, const double *states, size_t n, uint8_t *out,
                            int nthreads)
{
    orc_job j;
    memset(&j, 0, sizeof(j));
    j.c = c;
    j.a = states;
    j.stride = stride;
    j.out = out;
    run_threads(states_worker, &j, n, nthreads);
}

static void *motions_worker(void *p)
{
    orc_job *j = (orc_job *)p;
    size_t st = (size_t)j->sp->stride;
    for (size_t i = j->lo; i < j->hi; ++i)
    {
        if (j->last_t)
        {
            double t = 1.0;
            int ok = orc_check_motion_last(j->sp, j->c, j->a + i * st, j->b + i * st, &t, NULL);
            j->out[i] = (uint8_t)ok;
            j->last_t[i] = ok ? 1.0 : t;
        }
        else
            j->out[i] = (uint8_t)orc_check_motion(j->sp, j->c, j->a + i * st, j->b + i * st);
    }
    return NULL;
}

void orc_check_motions_batch(const orc_space *sp, const orc_checker *c, const double *s1, const double *s2,
                             size_t n, uint8_t *valid, double *last_t, int nthreads)
{
    orc_job j;
    memset(&j, 0, sizeof(j));
    j.sp = sp;
    j.c = c;
    j.a = s1;
    j.b = s2;
    j.out = valid;
    j.last_t = last_t;
    run_threads(motions_worker, &j, n, nthreads);
}

/* ------------------------------------------------------------------ */
/* nearest neighbours: NearestNeighborsLinear semantics                */
/* ------------------------------------------------------------------ */
/* NearestNeighborsLinear::nearestK (NearestNeighborsLinear.h:119-132) sorts
 * elements by distFun_(a,q) < distFun_(b,q) with an unstable std sort; the
 * north star fixes the tie rule as insertion index, so the oracle orders by
 * the total order (distance, index).  tests/test_oracle_ref.py checks that
 * the real Linear agrees wherever no tie touches the k-th boundary. */

typedef struct
{
    double d;
    uint64_t i;
} orc_di;

static int di_less(const orc_di *a, const orc_di *b)
{
    return a->d < b->d || (a->d == b->d && a->i < b->i);
}

static void heap_sift_down(orc_di *h, size_t n, size_t i)
{
    for (;;)
    {
        size_t l = 2 * i + 1, r = l + 1, m = i;
        if (l < n && di_less(&h[m], &h[l]))
            m = l;
        if (r < n && di_less(&h[m], &h[r]))
            m = r;
        if (m == i)
            return;
        orc_di t = h[i];
        h[i] = h[m];
        h[m] = t;
        i = m;
    }
}

static int di_cmp(const void *a, const void *b)
{
    const orc_di *x = (const orc_di *)a, *y = (const orc_di *)b;
    if (di_less(x, y))
        return -1;
    if (di_less(y, x))
        return 1;
    return 0;
}

size_t orc_knn(const orc_space *sp, const double *pts, size_t n, const double *q, size_t k, uint64_t *idx,
               double *dist)
{
    size_t kk = k < n ? k : n;
    if (kk == 0)
        return 0;
    orc_di *h = (orc_di *)malloc(sizeof(orc_di) * kk);
    size_t hs = 0;
    size_t st = (size_t)sp->stride;
    for (size_t i = 0; i < n; ++i)
    {
        /* argument order distFun_(element, query): NearestNeighborsLinear.h:168 */
        orc_di e = {orc_distance(sp, pts + i * st, q), (uint64_t)i};
        if (hs < kk)
        {
            h[hs++] = e;
            if (hs == kk)
                for (size_t j = kk / 2; j-- > 0;)
                    heap_sift_down(h, kk, j);
        }
        else if (di_less(&e, &h[0]))
        {
            h[0] = e;
            heap_sift_down(h, kk, 0);
        }
    }
    qsort(h, kk, sizeof(orc_di), di_cmp);
    for (size_t j = 0; j < kk; ++j)
    {
        idx[j] = h[j].i;
        if (dist)
            dist[j] = h[j].d;
    }
    free(h);
    return kk;
}

/* NearestNeighborsLinear::nearestR, NearestNeighborsLinear.h:135-142 (inclusive <=) */
size_t orc_radius(const orc_space *sp, const double *pts, size_t n, const double *q, double radius, uint64_t *idx,
                  double *dist, size_t cap)
{
    size_t st = (size_t)sp->stride;
    orc_di *h = (orc_di *)malloc(sizeof(orc_di) * (n ? n : 1));
    size_t c = 0;
    for (size_t i = 0; i < n; ++i)
    {
        double d = orc_distance(sp, pts + i * st, q);
        if (d <= radius)
        {
            h[c].d = d;
            h[c].i = (uint64_t)i;
            c++;
        }
    }
    qsort(h, c, sizeof(orc_di), di_cmp);
    for (size_t j = 0; j < c && j < cap; ++j)
    {
        idx[j] = h[j].i;
        if (dist)
            dist[j] = h[j].d;
    }
    free(h);
    return c;
}

static void *knn_worker(void *p)
{
    orc_job *j = (orc_job *)p;
    size_t st = (size_t)j->sp->stride;
    size_t kk = j->k < j->n ? j->k : j->n;
    for (size_t i = j->lo; i < j->hi; ++i)
        orc_knn(j->sp, j->pts, j->n, j->q + i * st, j->k, j->idx + i * kk, j->dist ? j->dist + i * kk : NULL);
    return NULL;
}

/* output rows have min(k,n) entries each */
void orc_knn_batch(const orc_space *sp, const double *pts, size_t n, const double *q, size_t nq, size_t k,
                   uint64_t *idx, double *dist, int nthreads)
{
    orc_job j;
    memset(&j, 0, sizeof(j));
    j.sp = sp;
    j.pts = pts;
    j.n = n;
    j.q = q;
    j.k = k;
    j.idx = idx;
    j.dist = dist;
    run_threads(knn_worker, &j, nq, nthreads);
}

/* ------------------------------------------------------------------ */
/* createHornEnvironment, demos/KinematicChain.h:279-318 (file output omitted) */
/* ------------------------------------------------------------------ */
int orc_horn_environment(unsigned d, double eps, double *segs)
{
    int n = 0;
    double w = 1. / (double)d, x = w, y = -eps, xN, yN, theta = 0., scale = w * (1. + ORC_PI * eps);
    for (unsigned int i = 0; i < d - 1; ++i)
    {
        theta += ORC_PI / (double)d;
        xN = x + cos(theta) * scale;
        yN = y + sin(theta) * scale;
        segs[4 * n + 0] = x;
        segs[4 * n + 1] = y;
        segs[4 * n + 2] = xN;
        segs[4 * n + 3] = yN;
        n++;
        x = xN;
        y = yN;
    }
    theta = 0.;
    x = w;
    y = eps;
    scale = w * (1.0 - ORC_PI * eps);
    for (unsigned int i = 0; i < d - 1; ++i)
    {
        theta += ORC_PI / d;
        xN = x + cos(theta) * scale;
        yN = y + sin(theta) * scale;
        segs[4 * n + 0] = x;
        segs[4 * n + 1] = y;
        segs[4 * n + 2] = xN;
        segs[4 * n + 3] = yN;
        n++;
        x = xN;
        y = yN;
    }
    return n;
}
