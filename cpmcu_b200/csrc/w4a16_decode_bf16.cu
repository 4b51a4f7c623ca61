#include <cuda_bf16.h>
#define DECODE_T __nv_bfloat16
#define DECODE_FN launch_decode_bf16
#include "decode_launch.inc"
