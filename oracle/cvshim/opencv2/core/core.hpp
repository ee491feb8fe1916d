// oracle/cvshim forwarding header (test infrastructure): the reference includes <opencv2/core/core.hpp>.
#include "../../cvshim.h"
