/* cv-shim: stand-in for <opencv2/highgui/highgui.hpp>; everything lives in cvshim.hpp */
#include "cvshim.hpp"
