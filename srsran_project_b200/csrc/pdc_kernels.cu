// The single CUDA translation unit holding the device code of libpuschdec_cuda.so: constant tables, the two
// code-block kernels and their host launchers.
#define PDC_KERNEL_TU 1
#include "pdc_device.cuh"
#include "pdc_kernel_lane.cuh"
#include "pdc_kernel_h2.cuh"

extern "C" cudaError_t pdc_launch_lane(const pdc_launch_params& prm, int block, size_t smem, cudaStream_t stream)
{
  pdc_cb_kernel_lane<<<prm.n, block, smem, stream>>>(prm);
  return cudaGetLastError();
}

// Per-device one-time setup (constant tables, opt-in shared memory); called from pdc_create with the device current.
extern "C" cudaError_t pdc_kernels_init_device(const pdc_bg_rows* rows)
{
  cudaError_t e = cudaMemcpyToSymbol(c_bg_rows, rows, sizeof(pdc_bg_rows));
  if (e != cudaSuccess) {
    return e;
  }
  e = cudaFuncSetAttribute(pdc_cb_kernel_lane, cudaFuncAttributeMaxDynamicSharedMemorySize, PDC_MAX_DYN_SMEM);
  if (e != cudaSuccess) {
    return e;
  }
  return pdc_h2_init_device();
}
