#include <cuda_fp16.h>
#define DECODE_T __half
#define DECODE_FN launch_decode_fp16
#include "decode_launch.inc"
