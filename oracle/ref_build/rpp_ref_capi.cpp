// C wrapper around the reference's RPP::Rpp, compiled together with the reference's own
// RobustPlanarPose/RPP.cpp + Rpoly.cpp (from /root/reference) into oracle/_ref/librpp_ref.so.
// TEST INFRASTRUCTURE ONLY.
#include "RPP.h"

extern "C" int rpp_ref(const double* model3xn, const double* iprts3xn, int n, double* R9, double* t3, int* iters,
                       double* obj_err, double* img_err)
{
    cv::Mat model(3, n, CV_64F), ip(3, n, CV_64F);
    for (int i = 0; i < 3 * n; ++i) { model.data[i] = model3xn[i]; ip.data[i] = iprts3xn[i]; }
    cv::Mat R, t;
    int it = 0;
    double oe = 0, ie = 0;
    bool ok = RPP::Rpp(model, ip, R, t, it, oe, ie);
    if (!ok || !R.data || !t.data) return 0;
    for (int i = 0; i < 9; ++i) R9[i] = R.data[i];
    for (int i = 0; i < 3; ++i) t3[i] = t.data[i];
    *iters = it; *obj_err = oe; *img_err = ie;
    return 1;
}
