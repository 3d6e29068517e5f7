// scan_gemv_kernel.cuh instantiated for int8 rows
#define ZVB_GEMV_DT ZVB200_DT_INT8
#include "scan_gemv_kernel.cuh"
