// Transport-block assembly on the device: what pusch_decoder_impl::join_and_notify and
// concatenate_codeblocks (pusch_decoder_impl.cpp:396-505) do on the host after the code blocks of a TB have been decoded.
//
//   C == 1 : tb_crc_ok = the code block's CRC; the first tbs_bits message bits are the payload (:416-424).
//   C  > 1 : only if EVERY code block passed (:425): the D = K - F - 24 data bits of each code block are concatenated
//            (copy_offset, bit granular) up to tbs bits, the next 24 bits of the last code block are the TB checksum (what may
//            follow is zero padding, :485-487), and tb_crc_ok = CRC24A(payload) == checksum (:426-437). The payload is written
//            before the check, i.e. also when the check then fails.
//
// Two kernels. pdc_tb_assemble_kernel: one CTA per (code block, TB) writes the payload words that START inside its code block
// (pulling the first bits of the next one where a word straddles) and its share of the CRC: the checksum test is
// "remainder of payload || checksum == 0", the remainder is linear, so code block r < C - 1 contributes
// (sum over its set data bits i of x^(D-1-i)) * x^(Dl + D (C-2-r)) mod P and the last one the sum over its first Dl bits of
// x^(Dl-1-i), Dl = tbs + 24 - (C - 1) D. pdc_tb_finish_kernel: one warp per TB xors the shares.
#pragma once
#include "pdc_device.cuh"

#define PDC_CRC24A_POLY 0x1864CFBu

// a * b mod P for 24-bit residues of CRC24A (shift-and-add, 24 steps).
__device__ __forceinline__ uint32_t tb_mulmod(uint32_t a, uint32_t b)
{
  uint32_t r = 0;
#pragma unroll 4
  for (int i = 23; i >= 0; --i) {
    r <<= 1;
    r ^= (r & 0x1000000u) ? PDC_CRC24A_POLY : 0u;
    r ^= ((b >> i) & 1u) ? a : 0u;
  }
  return r;
}

// A code block counts as passed if its CRC matched in this submission or an earlier one (flags).
__device__ __forceinline__ bool tb_cb_ok(const pdc_dev_cb& cb)
{
  return (cb.flags & (PDC_F_DECODED | PDC_F_DEMATCH_ONLY)) ? true : (cb.res->crc_pass != 0);
}

// 32 message bits of a code block starting at bit o (MSB first: bit_buffer.h:106-136), zero beyond word nw.
__device__ __forceinline__ uint32_t tb_bits_at(const uint32_t* __restrict__ cb32, int o, int nw)
{
  const int      k  = o >> 5, s = o & 31;
  const uint32_t hi = (k < nw) ? __byte_perm(cb32[k], 0u, 0x0123) : 0u;
  const uint32_t lo = (k + 1 < nw) ? __byte_perm(cb32[k + 1], 0u, 0x0123) : 0u;
  return __funnelshift_l(lo, hi, s);
}

__global__ void __launch_bounds__(128) pdc_tb_assemble_kernel(const pdc_dev_tb* __restrict__ tbs,
                                                               const pdc_dev_cb* __restrict__ cbs,
                                                               const uint32_t* __restrict__ crc_pow)
{
  const pdc_dev_tb tb = tbs[blockIdx.y];
  const int        r = blockIdx.x, C = (int)tb.nof_cbs, tid = threadIdx.x, nth = blockDim.x;
  if (r >= C) {
    return;
  }
  int ok = 1;
  for (int c = tid; c < C; c += nth) {
    ok &= tb_cb_ok(cbs[tb.first_cb + c]) ? 1 : 0;
  }
  if (!__syncthreads_and(ok)) {
    return; // nothing is written: the transport block buffer stays as it is (:425, :447)
  }
  const pdc_dev_cb cb   = cbs[tb.first_cb + r];
  const int        K    = ((cb.bg == 1) ? 22 : 10) * (int)cb.Z, nw = (K + 31) >> 5;
  const uint32_t*  cb32 = reinterpret_cast<const uint32_t*>(cb.out);
  const int        tbs_bits = (int)tb.tbs_bits, nbytes = tbs_bits >> 3;
  uint32_t*        out32 = reinterpret_cast<uint32_t*>(tb.out);
  if (C == 1) {
    for (int w = tid; 32 * w < tbs_bits; w += nth) {
      const int left = nbytes - 4 * w; // payload bytes from this word on
      if (left >= 4) {
        out32[w] = cb32[w];
      } else {
        for (int b = 0; b < left; ++b) {
          tb.out[4 * w + b] = cb.out[4 * w + b];
        }
      }
    }
    return;
  }
  const int       D      = (int)tb.cb_bits, p0 = r * D, p1 = p0 + D; // TB bit range of this code block (checksum included)
  const uint32_t* next32 = (r + 1 < C) ? reinterpret_cast<const uint32_t*>(cbs[tb.first_cb + r + 1].out) : nullptr;
  for (int w = ((p0 + 31) >> 5) + tid; 32 * w < p1 && 32 * w < tbs_bits; w += nth) {
    const int o  = 32 * w - p0, n1 = min(32, D - o); // n1 bits from this code block, the rest from the next one
    uint32_t  v  = tb_bits_at(cb32, o, nw);
    if (n1 < 32) {
      v &= ~(0xffffffffu >> n1);
      if (next32) {
        v |= tb_bits_at(next32, 0, nw) >> n1;
      }
    }
    const int left = nbytes - 4 * w;
    if (left >= 4) {
      out32[w] = __byte_perm(v, 0u, 0x0123);
    } else {
      for (int b = 0; b < left; ++b) {
        tb.out[4 * w + b] = (uint8_t)(v >> (24 - 8 * b));
      }
    }
  }
  // CRC share: sum over the set bits i < Dr of x^(Dr - 1 - i) mod P (crc_pow holds x^j mod P of CRC24A in its second table).
  const uint32_t* pw  = crc_pow + PDC_CRC_TABLE_LEN;
  const int       Dl = (int)tb.last_bits, Dr = (r == C - 1) ? Dl : D;
  uint32_t        acc = 0;
  for (int k = tid; 32 * k < Dr; k += nth) {
    uint32_t v = __byte_perm(cb32[k], 0u, 0x0123);
    if (Dr - 32 * k < 32) {
      v &= ~(0xffffffffu >> (Dr - 32 * k));
    }
    while (v) {
      const int b = __clz(v);
      v &= ~(0x80000000u >> b);
      acc ^= __ldg(pw + (Dr - 1 - (32 * k + b)));
    }
  }
  __shared__ uint32_t s_part[4];
  acc = __reduce_xor_sync(0xffffffffu, acc);
  if ((tid & 31) == 0) {
    s_part[tid >> 5] = acc;
  }
  __syncthreads();
  if (tid == 0) {
    uint32_t share = 0;
    for (int w = 0; w < (nth >> 5); ++w) {
      share ^= s_part[w];
    }
    // times x^(Dl + D (C - 2 - r)) for r < C - 1: square-and-multiply on x^D mod P.
    if (r != C - 1) {
      uint32_t base = __ldg(pw + D), f = __ldg(pw + Dl);
      for (int e = C - 2 - r; e != 0; e >>= 1) {
        if (e & 1) {
          f = tb_mulmod(f, base);
        }
        base = tb_mulmod(base, base);
      }
      share = tb_mulmod(share, f);
    }
    tb.terms[r] = share;
  }
}

__global__ void __launch_bounds__(128) pdc_tb_finish_kernel(const pdc_dev_tb* __restrict__ tbs,
                                                             const pdc_dev_cb* __restrict__ cbs, int ntb)
{
  const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (t >= ntb) {
    return;
  }
  const pdc_dev_tb tb = tbs[t];
  const int        C  = (int)tb.nof_cbs;
  int              nok = 0;
  for (int c = lane; c < C; c += 32) {
    nok += tb_cb_ok(cbs[tb.first_cb + c]) ? 1 : 0;
  }
  nok = __reduce_add_sync(0xffffffffu, nok);
  const bool all_ok = (nok == C);
  uint32_t   rem    = 0;
  if (all_ok && C > 1) {
    for (int c = lane; c < C; c += 32) {
      rem ^= tb.terms[c];
    }
    rem = __reduce_xor_sync(0xffffffffu, rem);
  }
  if (lane == 0) {
    pdc_tb_result res;
    res.tb_crc_ok  = (all_ok && rem == 0u) ? 1 : 0;
    res.written    = all_ok ? 1 : 0;
    res.nof_cbs_ok = (uint16_t)nok;
    *tb.res        = res;
  }
}

extern "C" cudaError_t pdc_launch_tb(const pdc_dev_tb* d_tbs, const pdc_dev_cb* d_cbs, const uint32_t* crc_pow, int ntb,
                                     int max_cbs, cudaStream_t stream)
{
  // grid.y is limited to 65535 transport blocks per launch
  for (int t0 = 0; t0 < ntb; t0 += 65535) {
    const int nt = min(65535, ntb - t0);
    pdc_tb_assemble_kernel<<<dim3((unsigned)max_cbs, (unsigned)nt), 128, 0, stream>>>(d_tbs + t0, d_cbs, crc_pow);
  }
  pdc_tb_finish_kernel<<<(ntb + 3) / 4, 128, 0, stream>>>(d_tbs, d_cbs, ntb);
  return cudaGetLastError();
}
