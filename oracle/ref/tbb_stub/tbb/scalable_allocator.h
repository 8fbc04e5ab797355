/* oracle/ref/tbb_stub: see tbb.h beside this file (TEST INFRASTRUCTURE) */
#include "tbb.h"
