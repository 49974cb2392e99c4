#!/usr/bin/env python3
"""BASELINE configs[0]: ldpc_decoder_benchmark (BG1, Z = 384, rate 1/3: 25344 de-matched soft bits per code block, 6 iterations,
no CRC -> every block runs all 6 iterations; ldpc_decoder_benchmark.cpp:143-147,193-196) as a batch on one GPU, next to the
compiled reference decoder on the host threads. Prints one JSON line. Not the round's headline bench (that is bench.py)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cbs", type=int, default=148 * 8)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--bg", type=int, default=1)
    ap.add_argument("--Z", type=int, default=384)
    ap.add_argument("--iters", type=int, default=6)
    ap.add_argument("--cpu", action="store_true")
    args = ap.parse_args()
    import torch

    import srsran_project_b200 as pdc
    from srsran_project_b200 import binding as B

    kb, ns = (22, 66) if args.bg == 1 else (10, 50)
    K, N = kb * args.Z, ns * args.Z
    n = args.cbs
    rng = np.random.default_rng(0)
    llrs = (rng.integers(0, 2, n * N, dtype=np.int8) * 20 - 10).astype(np.int8)
    stride = (K + 31) // 32 * 4
    d = B.make_descs(n)
    d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = args.bg, args.Z, 1, 0, B.PDC_CRC_NONE, args.iters
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
    for _ in range(args.steps):
        ctx.decode_device(d, d_llrs.data_ptr(), None, d_bits.data_ptr(), d_res.data_ptr(), st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    out = {"workload": f"ldpc_decoder_benchmark BG{args.bg} Z={args.Z} rate 1/3, {args.iters} iterations, no CRC, {n} CBs per step",
           "gpu_gbps_info": n * K / (ms * 1e-3) / 1e9, "gpu_us_per_cb": 1e3 * ms / n, "ms_per_step": ms}
    res = d_res.cpu().numpy().view(B.RESULT_DTYPE)
    out["nof_layers"] = int(res["nof_layers"][0])
    if args.cpu:
        from oracle import pyoracle as O

        m = min(n, 256)
        rd = np.zeros(m, O.REF_CB_DTYPE)
        rd["bg"], rd["Z"], rd["Qm"], rd["max_iter"], rd["E"], rd["decode_only"], rd["nof_crc_bits"] = args.bg, args.Z, 1, args.iters, N, 1, 16
        rd["llr_offset"] = np.arange(m, dtype=np.uint64) * N
        rd["out_offset"] = np.arange(m, dtype=np.uint64) * stride
        bits = np.zeros(m * stride, np.uint8)
        dec, dem = O.ref_best_flavours()
        nth = len(os.sched_getaffinity(0))
        sec, _ = O.ref_chain_run(dec, dem, nth, rd, llrs, np.zeros(1, np.int8), bits, 4)
        out["cpu_gbps_info"] = 4 * m * K / sec / 1e9
        out["cpu_threads"] = nth
        out["bits_equal"] = bool((bits == d_bits.cpu().numpy()[:m * stride]).all())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
