#define LUX_LOGN 11
#include "br_inst.cuh"
