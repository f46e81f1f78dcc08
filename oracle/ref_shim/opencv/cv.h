// forwards to the from-scratch OpenCV sliver (oracle/ref_shim/cv_shim.h); test infrastructure only
#pragma once
#include "cv_shim.h"
