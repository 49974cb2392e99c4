// Evidence for DESIGN.md 3.1: what the byte-SIMD intrinsics the north star names compile to on sm_100a, against the two-lane
// (16x2 / half2) instructions the kernels use instead. Build + disassemble: tools/sass_premise.sh -> profiles/.
#include <cuda_fp16.h>
#include <stdint.h>
extern "C" __global__ void k_vminu4(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = __vminu4(a[threadIdx.x], b[threadIdx.x]); }
extern "C" __global__ void k_vsubss4(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = __vsubss4(a[threadIdx.x], b[threadIdx.x]); }
extern "C" __global__ void k_vabsss4(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = __vabsss4(a[threadIdx.x]); }
extern "C" __global__ void k_vaddss4(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = __vaddss4(a[threadIdx.x], b[threadIdx.x]); }
extern "C" __global__ void k_vmins2(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = __vmins2(a[threadIdx.x], b[threadIdx.x]); }
extern "C" __global__ void k_vadd2(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = __vadd2(a[threadIdx.x], b[threadIdx.x]); }
extern "C" __global__ void k_vaddss2(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = __vaddss2(a[threadIdx.x], b[threadIdx.x]); }
extern "C" __global__ void k_hmin2(const __half2* a, const __half2* b, __half2* o) { o[threadIdx.x] = __hmin2(a[threadIdx.x], b[threadIdx.x]); }
extern "C" __global__ void k_hfma2_sat(const __half2* a, const __half2* b, __half2* o) { o[threadIdx.x] = __hfma2_sat(a[threadIdx.x], b[threadIdx.x], a[threadIdx.x]); }
extern "C" __global__ void k_empty(const uint32_t* a, const uint32_t* b, uint32_t* o) { o[threadIdx.x] = a[threadIdx.x] ^ b[threadIdx.x]; }
