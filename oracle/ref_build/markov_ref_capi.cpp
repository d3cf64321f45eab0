// C wrapper around the reference's MarkovModel, compiled together with the reference's own, unmodified
// include/mantis3/Markov.cpp (from /root/reference) into oracle/_ref/libmarkov_ref.so.  TEST INFRASTRUCTURE ONLY:
// it pins oracle/legacy.c's restatement of the yaw filter.
//
// Markov.h pulls in "Mantis3Types.h" (ROS, tf, OpenCV) for one thing: Hypothesis::getW2C().getBasis().getRPY() and
// Hypothesis::error.  The include guard of that header is defined here so that its body is skipped, and a Hypothesis
// that carries a yaw angle stands in.  opencv2/plot.hpp and highgui.hpp come from ref_build/shim.
#define MANTIS_INCLUDE_MANTIS3_MANTIS3TYPES_H_
#define ROS_DEBUG_STREAM(x) do { } while (0)   /* rosconsole logging (came in through the skipped header) */
#include <vector>

struct RefBasis {
    double yaw;
    void getRPY(double& r, double& p, double& y) const { r = 0; p = 0; y = yaw; }
};
struct RefTransform {
    double yaw;
    RefBasis getBasis() const { return RefBasis{yaw}; }
};
struct Hypothesis {
    double yaw = 0, error = 0;
    RefTransform getW2C() const { return RefTransform{yaw}; }
};

#include "Markov.cpp"

static markovPlane load(const double* p)
{
    markovPlane m;
    for (int i = 0; i < DEGREES; ++i) m[i] = p[i];
    return m;
}
static void store(const markovPlane& m, double* p)
{
    for (int i = 0; i < DEGREES; ++i) p[i] = m[i];
}

extern "C" {

// MarkovModel::updateWeights, Markov.cpp:76-124
void markov_ref_update_weights(double* p360, double stddev)
{
    markovPlane zero{};
    MarkovModel m(zero);
    markovPlane y = load(p360);
    m.updateWeights(y, stddev);
    store(y, p360);
}

// MarkovModel::MarkovModel(Hypothesis), Markov.cpp:14-26
void markov_ref_init(double* p360, double yaw_rad)
{
    Hypothesis h;
    h.yaw = yaw_rad;
    MarkovModel m(h);
    store(m.getDistrbution(), p360);
}

// MarkovModel::convolve, Markov.cpp:227-258
void markov_ref_convolve(double* p360, double dtheta, double dt)
{
    MarkovModel m(load(p360));
    m.convolve(dtheta, dt);
    store(m.getDistrbution(), p360);
}

// MarkovModel::senseFusion(markovPlane), Markov.cpp:164-183.  (senseFusion(Hypothesis), :185-205, cannot be run: its
// zeroing loop `for(int i=0; sense.size(); ++i)` never ends.  The caller builds the sense plane the way that function
// means to -- a one-hot bin smoothed by updateWeights(sense, 3.5) -- with the reference's own updateWeights.)
void markov_ref_sense_plane(double* p360, const double* sense360)
{
    MarkovModel m(load(p360));
    m.senseFusion(load(sense360));
    store(m.getDistrbution(), p360);
}

// MarkovModel::updateHypothesis, Markov.cpp:207-220
void markov_ref_update_errors(const double* p360, const double* yaw_rad, double* err, int n)
{
    MarkovModel m(load(p360));
    std::vector<Hypothesis> hyps((size_t)n);
    for (int i = 0; i < n; ++i) { hyps[i].yaw = yaw_rad[i]; hyps[i].error = err[i]; }
    m.updateHypothesis(hyps);
    for (int i = 0; i < n; ++i) err[i] = hyps[i].error;
}

// MarkovModel::getYaw, Markov.cpp:265-274
double markov_ref_get_yaw(const double* p360)
{
    MarkovModel m(load(p360));
    return m.getYaw();
}

}  // extern "C"
