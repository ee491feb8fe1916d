// oracle/cvshim forwarding header (test infrastructure): the reference includes <opencv2/features2d/features2d.hpp>.
#include "../../cvshim.h"
