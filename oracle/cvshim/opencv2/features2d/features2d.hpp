/* cv-shim: stand-in for <opencv2/features2d/features2d.hpp>; everything lives in cvshim.hpp */
#include "cvshim.hpp"
