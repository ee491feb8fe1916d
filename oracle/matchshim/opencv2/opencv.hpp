// oracle/matchshim forwarding header (test infrastructure): the reference includes <opencv2/opencv.hpp>.
#include "../../cvshim/cvshim.h"
