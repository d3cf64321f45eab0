// Shim (written for this repo, TEST INFRASTRUCTURE ONLY): the cv::plot surface include/mantis3/Markov.cpp names in
// its two plotMarkovPlane() functions, so that the file compiles unmodified without OpenCV.  Nothing is drawn.
#pragma once
#include "core/core.hpp"

namespace cv {
struct Scalar {
    double v[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) : v{a, b, c, d} {}
};
template <typename T>
struct Ptr {
    T obj;
    T* operator->() { return &obj; }
};
namespace plot {
struct Plot2d {
    void setPlotBackgroundColor(Scalar) {}
    void setPlotLineColor(Scalar) {}
    void render(Mat&) {}
};
inline Ptr<Plot2d> createPlot2d(const Mat&) { return Ptr<Plot2d>(); }
}  // namespace plot
}  // namespace cv
