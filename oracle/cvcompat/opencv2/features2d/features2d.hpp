/* oracle/cvcompat: forwards to the single compat header (TEST INFRASTRUCTURE). */
#include "opencv2/core/core.hpp"
