/* cv-shim: stand-in for <opencv2/imgproc/imgproc.hpp>; everything lives in cvshim.hpp */
#include "cvshim.hpp"
