// TEST INFRASTRUCTURE (oracle shim) -- not product code.  CRAM is not supported by the shim;
// the option enums live in htslib/sam.h.  Included by reference gclib/GSam.h:7.
#ifndef BRB_SHIM_CRAM_H
#define BRB_SHIM_CRAM_H
#include "htslib/sam.h"
#endif
