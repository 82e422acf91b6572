/* oracle/ttm_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle.h).  Filled in below. */
#include "oracle.h"
