// oracle/cvshim forwarding header (test infrastructure): the reference includes <opencv2/imgproc/imgproc.hpp>.
#include "../../cvshim.h"
