// oracle/shim: only the two hot-path headers of the reference module are pulled in (test infrastructure)
#include "opencv2/core.hpp"
#include "opencv2/ximgproc/edge_filter.hpp"
#include "opencv2/ximgproc/disparity_filter.hpp"
