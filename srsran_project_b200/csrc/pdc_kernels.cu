// The single CUDA translation unit holding the device code of libpuschdec_cuda.so: constant tables, the two
// code-block kernels and their host launchers.
#define PDC_KERNEL_TU 1
#include "pdc_device.cuh"
#include "pdc_kernel_lane.cuh"
#include "pdc_kernel_h2.cuh"
#include "pdc_kernel_h2g.cuh"
#include "pdc_kernel_tb.cuh"
#include "pdc_kernel_enc.cuh"
#include "pdc_probe.cuh"

extern "C" cudaError_t pdc_launch_lane(const pdc_launch_params& prm, int block, size_t smem, cudaStream_t stream)
{
  pdc_cb_kernel_lane<<<prm.n, block, smem, stream>>>(prm);
  return cudaGetLastError();
}

// Descriptor upload without the copy engine: the SMs read the page-locked host image (device-addressable: unified addressing)
// and write the device copy. A small host-to-device memcpy queues behind every large LLR transfer already handed to the copy
// engine (measured: 100 us for 20 KB with sixteen threads pushing 1.4 MB regions); these reads interleave with them on the link.
__global__ void __launch_bounds__(256) pdc_upload_kernel(uint32_t* __restrict__ dst0, const uint32_t* __restrict__ src0, uint32_t n0,
                                                         uint32_t* __restrict__ dst1, const uint32_t* __restrict__ src1, uint32_t n1)
{
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n0 + n1; i += stride) {
    if (i < n0) {
      dst0[i] = src0[i];
    } else {
      dst1[i - n0] = src1[i - n0];
    }
  }
}
extern "C" cudaError_t pdc_launch_upload(void* dst0, const void* src0, size_t bytes0, void* dst1, const void* src1, size_t bytes1,
                                         cudaStream_t stream)
{
  if (((bytes0 | bytes1) & 3) != 0) {
    return cudaErrorInvalidValue;
  }
  const uint32_t n0 = (uint32_t)(bytes0 >> 2), n1 = (uint32_t)(bytes1 >> 2);
  const uint32_t grid = (n0 + n1 + 1023) / 1024; // four words per thread in flight
  if (grid == 0) {
    return cudaSuccess;
  }
  pdc_upload_kernel<<<grid, 256, 0, stream>>>((uint32_t*)dst0, (const uint32_t*)src0, n0, (uint32_t*)dst1, (const uint32_t*)src1, n1);
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
  const void* h2_kernels[4] = {(const void*)pdc_cb_kernel_h2<1, 0>, (const void*)pdc_cb_kernel_h2<2, 0>,
                               (const void*)pdc_cb_kernel_h2<1, 384>, (const void*)pdc_cb_kernel_h2<2, 384>};
  for (const void* k : h2_kernels) {
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e != cudaSuccess) {
      return e;
    }
  }
  const void* h2g_kernels[2] = {(const void*)pdc_cb_kernel_h2g<1>, (const void*)pdc_cb_kernel_h2g<2>};
  for (const void* k : h2g_kernels) {
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, PDC_MAX_DYN_SMEM);
    if (e != cudaSuccess) {
      return e;
    }
  }
  // Structure the general packed kernel relies on: in every extension row the last edge is the row's own degree-1
  // parity column (Kb + m) and every other edge lies in the core columns.
  for (int bg = 0; bg != 2; ++bg) {
    const uint8_t* deg = bg ? pdc_bg2_row_deg : pdc_bg1_row_deg;
    const uint8_t* col = bg ? pdc_bg2_edge_col : pdc_bg1_edge_col;
    const int      M = bg ? PDC_BG2_ROWS : PDC_BG1_ROWS, kb = bg ? 10 : 22;
    int            e0 = 0;
    for (int m = 0; m != M; ++m) {
      if (m >= 4) {
        if (col[e0 + deg[m] - 1] != kb + m || deg[m] < 3 || deg[m] > 10) {
          return cudaErrorInvalidValue;
        }
        for (int k = 0; k + 1 < deg[m]; ++k) {
          if (col[e0 + k] >= kb + 4) {
            return cudaErrorInvalidValue;
          }
        }
      }
      e0 += deg[m];
    }
  }
  e = pdc_h2g_init_device();
  if (e != cudaSuccess) {
    return e;
  }
  return pdc_h2_init_device();
}

// Times one issue-rate probe (see pdc_probe.cuh) on `stream`; reports thread-level instructions per second.
extern "C" cudaError_t pdc_probe_run(int mode, int sm_count, cudaStream_t stream, double* thread_inst_per_s)
{
  uint32_t*   d_out = nullptr;
  cudaError_t e     = cudaMalloc((void**)&d_out, sizeof(uint32_t) * 256 * (size_t)sm_count * 8);
  if (e != cudaSuccess) {
    return e;
  }
  cudaEvent_t t0, t1;
  cudaEventCreate(&t0), cudaEventCreate(&t1);
  double winst = 0;
  const int blocks = sm_count * 8, iters = 2000;
  e = pdc_probe_launch(mode, blocks, 50, d_out, stream, &winst); // warm-up
  double best = 0;
  for (int rep = 0; rep < 3 && e == cudaSuccess; ++rep) {
    cudaEventRecord(t0, stream);
    e = pdc_probe_launch(mode, blocks, iters, d_out, stream, &winst);
    cudaEventRecord(t1, stream);
    cudaEventSynchronize(t1);
    float ms = 0;
    cudaEventElapsedTime(&ms, t0, t1);
    double rate = winst * 32.0 / (ms * 1e-3);
    best        = rate > best ? rate : best;
  }
  cudaEventDestroy(t0), cudaEventDestroy(t1);
  cudaFree(d_out);
  *thread_inst_per_s = best;
  return e;
}
