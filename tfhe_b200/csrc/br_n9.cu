#define LUX_LOGN 9
#include "br_inst.cuh"
