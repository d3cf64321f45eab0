// Shim (written for this repo, TEST INFRASTRUCTURE ONLY): imshow / waitKey as no-ops for include/mantis3/Markov.cpp.
#pragma once
#include "core/core.hpp"

namespace cv {
inline void imshow(const char*, const Mat&) {}
inline int waitKey(int = 0) { return -1; }
}  // namespace cv
