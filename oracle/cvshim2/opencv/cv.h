// oracle/cvshim2 forwarding header -- TEST INFRASTRUCTURE (see opencv2/core/core.hpp)
#include "opencv2/core/core.hpp"
