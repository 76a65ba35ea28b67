/* cv-shim: stand-in for <opencv2/opencv.hpp>; everything lives in cvshim.hpp */
#include "cvshim.hpp"
