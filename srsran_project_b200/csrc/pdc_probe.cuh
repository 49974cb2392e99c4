// Issue-rate microbenchmarks used as the roofline denominator of the decoder kernels (bench.py "roofline.peak").
// The decoder is bound by the SM's integer / half2 ALU issue rate, not by HBM or tensor cores, so the peak is measured
// on the same device, in the same run, with straight-line chains of the instruction classes the decoder uses.
//   mode 0: LOP3 only            (alu pipe)
//   mode 1: IMAD only            (fma pipe)
//   mode 2: LOP3 + IMAD 1:1      (both pipes; the 32-bit integer peak)
//   mode 3: HMNMX2 + HFMA2 1:1   (both pipes, two 16-bit lanes per instruction: what the packed kernel issues)
//   mode 4: VIMNMX.S16x2 + VIADD.16x2
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define PDC_PROBE_CHAINS 8
#define PDC_PROBE_UNROLL 16

template <int MODE>
__global__ void __launch_bounds__(256) pdc_probe_kernel(uint32_t* out, int iters, uint32_t seed)
{
  uint32_t a[PDC_PROBE_CHAINS], b[PDC_PROBE_CHAINS];
#pragma unroll
  for (int k = 0; k < PDC_PROBE_CHAINS; ++k) {
    a[k] = seed + threadIdx.x * 31u + k * 7u;
    b[k] = seed ^ (threadIdx.x * 17u + k * 3u);
  }
  const uint32_t c = seed | 1u;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < PDC_PROBE_UNROLL; ++u) {
#pragma unroll
      for (int k = 0; k < PDC_PROBE_CHAINS; ++k) {
        if (MODE == 0) {
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[k]) : "r"(b[k]), "r"(c));
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[k]) : "r"(a[k]), "r"(c));
        } else if (MODE == 1) {
          asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(c), "r"(b[k]));
          asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(b[k]) : "r"(c), "r"(a[k]));
        } else if (MODE == 2) {
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[k]) : "r"(b[k]), "r"(c));
          asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(b[k]) : "r"(c), "r"(a[k]));
        } else if (MODE == 3) {
          asm volatile("min.f16x2 %0, %0, %1;" : "+r"(a[k]) : "r"(b[k]));
          asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(b[k]) : "r"(c), "r"(a[k]));
        } else {
          asm volatile("min.s16x2 %0, %0, %1;" : "+r"(a[k]) : "r"(b[k]));
          asm volatile("add.s16x2 %0, %0, %1;" : "+r"(b[k]) : "r"(a[k]));
        }
      }
    }
  }
  uint32_t r = 0;
#pragma unroll
  for (int k = 0; k < PDC_PROBE_CHAINS; ++k) {
    r ^= a[k] ^ b[k];
  }
  if (r == 0x12345678u) { // practically never: keeps the chains alive without a store per thread
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  }
}

// Launches one probe; returns the number of warp-level instructions of the measured class it executes in total.
static cudaError_t pdc_probe_launch(int mode, int blocks, int iters, uint32_t* d_out, cudaStream_t stream,
                                    double* warp_instructions)
{
  const int threads = 256;
  switch (mode) {
    case 0:
      pdc_probe_kernel<0><<<blocks, threads, 0, stream>>>(d_out, iters, 12345u);
      break;
    case 1:
      pdc_probe_kernel<1><<<blocks, threads, 0, stream>>>(d_out, iters, 12345u);
      break;
    case 2:
      pdc_probe_kernel<2><<<blocks, threads, 0, stream>>>(d_out, iters, 12345u);
      break;
    case 3:
      pdc_probe_kernel<3><<<blocks, threads, 0, stream>>>(d_out, iters, 12345u);
      break;
    case 4:
      pdc_probe_kernel<4><<<blocks, threads, 0, stream>>>(d_out, iters, 12345u);
      break;
    default:
      return cudaErrorInvalidValue;
  }
  *warp_instructions = (double)blocks * (threads / 32) * (double)iters * PDC_PROBE_UNROLL * PDC_PROBE_CHAINS * 2.0;
  return cudaGetLastError();
}
