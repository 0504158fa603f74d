// Kernel group 5: long-read soft-clip rescue (ksw2 extension alignment).  Filled in below.
#pragma once
#include "dev_common.cuh"
#include "k_chain.cuh"
