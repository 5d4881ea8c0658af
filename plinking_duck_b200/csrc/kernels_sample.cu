// K4-K7 live here (added below).
#include "kernels.cuh"
