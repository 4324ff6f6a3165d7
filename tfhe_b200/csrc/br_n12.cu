#define LUX_LOGN 12
#include "br_inst.cuh"
