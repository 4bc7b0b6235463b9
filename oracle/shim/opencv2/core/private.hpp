// oracle/shim: forwards to the OpenCV-core stand-in (test infrastructure)
#include "opencv2/core.hpp"
