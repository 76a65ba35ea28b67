/* cv-shim: stand-in for <opencv2/core/core.hpp>; everything lives in cvshim.hpp */
#include "cvshim.hpp"
