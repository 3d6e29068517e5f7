// scan_gemv_kernel.cuh instantiated for fp16 rows
#define ZVB_GEMV_DT ZVB200_DT_FP16
#include "scan_gemv_kernel.cuh"
