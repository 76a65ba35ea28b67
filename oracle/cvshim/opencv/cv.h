/* cv-shim: stand-in for <opencv/cv.h>; everything lives in cvshim.hpp */
#include "cvshim.hpp"
