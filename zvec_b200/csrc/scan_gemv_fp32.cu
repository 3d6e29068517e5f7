// scan_gemv_kernel.cuh instantiated for fp32 rows
#define ZVB_GEMV_DT ZVB200_DT_FP32
#include "scan_gemv_kernel.cuh"
