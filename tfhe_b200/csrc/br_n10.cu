#define LUX_LOGN 10
#include "br_inst.cuh"
