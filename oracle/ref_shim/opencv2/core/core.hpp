// TEST INFRASTRUCTURE ONLY — forwards to the OpenCV stand-in (see vs_cv_standin.hpp).
#include "vs_cv_standin.hpp"
