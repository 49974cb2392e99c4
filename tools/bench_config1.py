#!/usr/bin/env python3
"""BASELINE configs[0]: ldpc_decoder_benchmark (BG1, Z = 384, rate 1/3: 25344 de-matched soft bits per code block, 6 iterations,
no CRC -> every block runs all 6 iterations; ldpc_decoder_benchmark.cpp:143-147,193-196).

Input: the benchmark's own pattern, `(rgen() & 1) * 20 - 10` with rgen = std::mt19937(0): code block 0 of the batch is the literal
block the benchmark decodes for BG1 / Z = 384 when it is run for that lifting size alone (the generator's first 24 Z outputs go to
the rate-0.917 block, the next 66 Z to this one); the other code blocks of the batch continue the same stream.

Two figures, as SURVEY 8d asks: the batch on one GPU (pdc_decode_device, LLRs resident) and ONE decode() call at a time through
the reference's ldpc_decoder interface (the CUDA adapter: latency, PCIe round trip included), next to the compiled reference
decoders on the host. run() returns a dict; as a script it prints one JSON line."""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
HARNESS = os.path.join(ROOT, "tests", "adapters", "_build", "libadapter_harness.so")


def mt19937_llrs(n, skip=0):
    """(rgen() & 1) * 20 - 10 for the outputs skip .. skip + n of std::mt19937(0)."""
    bg = np.random.MT19937()
    bg._legacy_seeding(0)  # init_genrand(0): the seeding of std::mt19937(0)
    raw = bg.random_raw(skip + n)[skip:]
    return ((raw & 1).astype(np.int16) * 20 - 10).astype(np.int8)


def decode_latency(impl, bg, Z, llrs, iters, reps):
    """One ldpc_decoder::decode per measurement through tests/adapters (impl: 0 generic, 1 avx2, 2 avx512, 3 cuda).
    -> (microseconds per call [reps], packed message bits) or None when the harness is not built."""
    if not os.path.exists(HARNESS):
        return None
    lib = ctypes.CDLL(HARNESS)
    vp, u = ctypes.c_void_p, ctypes.c_uint
    lib.adt_ldpc_decode_latency.argtypes = [ctypes.c_int, ctypes.c_int, u, vp, u, u, u, vp, vp]
    us = np.zeros(reps, np.float64)
    K = (22 if bg == 1 else 10) * Z
    bits = np.zeros((K + 7) // 8, np.uint8)
    llrs = np.ascontiguousarray(llrs, np.int8)
    rc = lib.adt_ldpc_decode_latency(impl, bg, Z, llrs.ctypes.data, llrs.size, iters, reps, us.ctypes.data, bits.ctypes.data)
    if rc < 0:
        return None
    return us, bits


def run(cbs=148 * 8, steps=20, bg=1, Z=384, iters=6, cpu=True, peak_ops=None):
    import torch

    import srsran_project_b200 as pdc
    from srsran_project_b200 import binding as B
    from tools.bench_configs import lane_ops

    kb, ns = (22, 66) if bg == 1 else (10, 50)
    K, N = kb * Z, ns * Z
    n = cbs
    llrs = mt19937_llrs(n * N, skip=(kb + 2) * Z)
    stride = (K + 31) // 32 * 4
    d = B.make_descs(n)
    d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = bg, Z, 1, 0, B.PDC_CRC_NONE, iters
    d["flags"], d["E"], d["harq_slot"] = B.PDC_F_DECODE_ONLY, N, -1
    d["llr_offset"] = np.arange(n, dtype=np.uint64) * N
    d["out_offset"] = np.arange(n, dtype=np.uint64) * stride
    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=n, max_batch_llrs=n * N + 64, harq_slots=0, nof_tickets=1)
    d_llrs = torch.from_numpy(llrs).cuda()
    d_bits = torch.zeros(n * stride, dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(n * 8, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        ctx.decode_device(d, d_llrs.data_ptr(), None, d_bits.data_ptr(), d_res.data_ptr(), st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        ctx.decode_device(d, d_llrs.data_ptr(), None, d_bits.data_ptr(), d_res.data_ptr(), st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    res = d_res.cpu().numpy().view(B.RESULT_DTYPE)
    gpu_bits = d_bits.cpu().numpy()
    L = int(res["nof_layers"][0])
    ops = n * lane_ops(bg, Z, iters, L, iters)
    out = {"config": f"config 1: ldpc_decoder_benchmark BG{bg} Z={Z} rate 1/3, {iters} iterations, no CRC, std::mt19937(0) +-10 LLRs",
           "code_blocks": n, "value": n * K / (ms * 1e-3) / 1e9, "unit": "Gbit/s", "gpu_us_per_cb": 1e3 * ms / n, "ms_per_step": ms,
           "nof_layers": L, "kernel": f"pdc_cb_kernel_h2g<{bg}>",
           "roofline": {"bound": "alu", "achieved": ops / (ms * 1e-3) / 1e12, "unit": "Tops/s (int8 lane-ops)",
                        "peak": None if peak_ops is None else peak_ops / 1e12,
                        "frac": None if peak_ops is None else ops / (ms * 1e-3) / peak_ops}}
    ctx.close()
    if cpu:
        from oracle import pyoracle as O

        m = min(n, 256)
        rd = np.zeros(m, O.REF_CB_DTYPE)
        rd["bg"], rd["Z"], rd["Qm"], rd["max_iter"], rd["E"], rd["decode_only"], rd["nof_crc_bits"] = bg, Z, 1, iters, N, 1, 16
        rd["llr_offset"] = np.arange(m, dtype=np.uint64) * N
        rd["out_offset"] = np.arange(m, dtype=np.uint64) * stride
        bits = np.zeros(m * stride, np.uint8)
        dec, dem = O.ref_best_flavours()
        nth = len(os.sched_getaffinity(0))
        sec, _ = O.ref_chain_run(dec, dem, nth, rd, llrs, np.zeros(1, np.int8), bits, 4)
        out["cpu_gbps_info"] = 4 * m * K / sec / 1e9
        out["cpu_threads"] = nth
        out["bit_exact"] = bool((bits == gpu_bits[:m * stride]).all())
        if not out["bit_exact"]:
            raise RuntimeError("config 1: the batch differs from the compiled reference decoder")
    # batch = 1: one decode() call per measurement through the reference's operator interface
    lat = decode_latency(3, bg, Z, llrs[:N], iters, 200)
    if lat is not None:
        us, lbits = lat
        us = us[20:]
        out["batch1"] = {"api": "srsran::ldpc_decoder::decode (CUDA adapter, one code block per call, host buffers)",
                         "us_p50": float(np.percentile(us, 50)), "us_p99": float(np.percentile(us, 99)),
                         "gbps_info": K / (np.percentile(us, 50) * 1e-6) / 1e9,
                         "bit_exact": bool((lbits == gpu_bits[:lbits.size]).all())}
        if cpu:
            from oracle import pyoracle as O

            dec, _ = O.ref_best_flavours()
            r = decode_latency(dec, bg, Z, llrs[:N], iters, 60)
            if r is not None:
                out["batch1"]["cpu_us_p50"] = float(np.percentile(r[0][5:], 50))
                out["batch1"]["cpu_us_p99"] = float(np.percentile(r[0][5:], 99))
                out["batch1"]["cpu_impl"] = {0: "generic", 1: "avx2", 2: "avx512"}[dec]
                out["batch1"]["bit_exact"] = out["batch1"]["bit_exact"] and bool((r[1] == lbits).all())
        if not out["batch1"]["bit_exact"]:
            raise RuntimeError("config 1: the single decode() call differs from the batch / the reference decoder")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cbs", type=int, default=148 * 8)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--bg", type=int, default=1)
    ap.add_argument("--Z", type=int, default=384)
    ap.add_argument("--iters", type=int, default=6)
    ap.add_argument("--cpu", action="store_true")
    args = ap.parse_args()
    print(json.dumps(run(args.cbs, args.steps, args.bg, args.Z, args.iters, args.cpu)))


if __name__ == "__main__":
    main()
