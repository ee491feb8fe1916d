// oracle/cvshim forwarding header (test infrastructure): the reference includes <opencv/cv.h>.
#include "../cvshim.h"
