// Packed fp16x2 code-block kernel (placeholder until the real kernel lands: reports "unsupported" for every size so
// that the lane kernel takes every code block).
#pragma once
#include "pdc_device.cuh"

extern "C" int pdc_h2_supported(int, int)
{
  return 0;
}
extern "C" size_t pdc_h2_smem_bytes(int, int, int)
{
  return 0;
}
extern "C" cudaError_t pdc_launch_h2(const pdc_launch_params&, int, size_t, cudaStream_t)
{
  return cudaErrorNotSupported;
}
static cudaError_t pdc_h2_init_device()
{
  return cudaSuccess;
}
