/* oracle/refshim: every OpenCV header of the reference resolves to the one stand-in (see opencv2/opencv.hpp) */
#include "../../opencv2/opencv.hpp"
