// oracle/cv_compat -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Forwarding header: the reference's
// LK sources include OpenCV by this name; everything they use lives in cv_compat.hpp.
#pragma once
#include "../../cv_compat.hpp"
