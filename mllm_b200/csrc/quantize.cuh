#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
// Q8_0 quantisation into the Linear kernels' internal split layout (quantize.cu).
int launch_quantize_q8_0_split(const float *x, int8_t *qs, float *d, int rows, int k, int ldx, cudaStream_t st);
