// oracle/cvshim forwarding header (test infrastructure): the reference includes <opencv2/highgui/highgui.hpp>.
#include "../../cvshim.h"
