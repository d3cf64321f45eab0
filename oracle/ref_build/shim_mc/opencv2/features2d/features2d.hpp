// forwarding header of the MonteCarlo.cpp shim (see mc_ros.hpp / mc_cv.hpp): written for this repository, test infrastructure only
#pragma once
#include "mc_ros.hpp"
