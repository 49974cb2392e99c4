#!/usr/bin/env python3
"""The BASELINE.json configurations that are NOT the headline bench line (bench.py is config 2 / 5), each as one batch on one
GPU next to the compiled reference chain on all host threads, with a bit-exact comparison of the whole batch:

  config 1  ldpc_decoder_benchmark: BG1 Z=384 rate 1/3, 6 iterations, no CRC (decode-only)          -> tools/bench_config1.py
  config 3  BG2 small-TB sweep: all 51 lifting sizes x 8 code blocks, mixed Qm / rate / rv in one shuffled batch
  config 4  HARQ: config-2 code blocks at low SNR, rv0 (new data) then rv2 (combining) over a device-resident soft buffer

Prints one JSON line per configuration. Timing: CUDA events around `steps` repetitions of pdc_decode_device (LLRs resident);
the CPU figure is oracle/_ref (unmodified reference) through pusch_codeblock_decoder on every host thread.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def message_with_crc(S, rng, K, F, crc_id):
    ncrc = 16 if crc_id == 1 else 24
    payload = rng.integers(0, 2, K - F - ncrc, dtype=np.uint8)
    c = S.crc(crc_id, payload)
    crc_bits = np.array([(c >> (ncrc - 1 - i)) & 1 for i in range(ncrc)], np.uint8)
    return np.concatenate([payload, crc_bits, np.zeros(F, np.uint8)])


BG_DEG = {1: [19, 19, 19, 19, 3, 8, 9, 7, 10, 9, 7, 8, 7, 6, 7, 7, 6, 6, 6, 6, 6, 6, 5, 5, 6, 5, 5, 4, 5, 5, 5, 5, 5, 5, 5, 5, 5, 4, 5,
              5, 4, 5, 4, 5, 5, 4],
          2: [8, 10, 8, 10, 4, 6, 6, 6, 4, 5, 5, 5, 4, 5, 5, 4, 5, 5, 4, 4, 4, 4, 3, 4, 4, 3, 5, 3, 4, 3, 5, 3, 4, 4, 4, 4, 4, 3, 4,
              4, 4, 4]}


def lane_ops(bg, Z, iterations, layers, max_iter=6):
    """ops(CB) = I * Z * (8 * sum_{m<L} deg(m) + 2 L) int8 lane-operations."""
    cs = np.concatenate([[0], np.cumsum(BG_DEG[bg])])
    it = iterations if iterations > 0 else max_iter
    return float(it) * Z * (8.0 * cs[layers] + 2.0 * layers)


def run_batch(name, cases, txs, steps, cpu=True, peak_ops=None, stale=None):
    """cases: list of dict(bg, Z, Qm, F, Nref, crc_id, E); txs: list over transmissions of (rv list, llr list).
    peak_ops: the roofline denominator (two-lane issue peak x 2, ops/s) measured by the caller, or None.
    stale: (position, value) of one soft bit that every HARQ slot holds before the first transmission (None: empty slots)."""
    import torch

    import srsran_project_b200 as pdc
    from oracle import pyoracle as O
    from srsran_project_b200 import binding as B

    n = len(cases)
    kb = lambda bg: 22 if bg == 1 else 10
    ns = lambda bg: 66 if bg == 1 else 50
    out_off = np.zeros(n, np.uint64)
    soft_off = np.zeros(n, np.uint64)
    oo = so = 0
    for i, c in enumerate(cases):
        out_off[i], soft_off[i] = oo, so
        oo += (kb(c["bg"]) * c["Z"] + 31) // 32 * 4
        so += (ns(c["bg"]) * c["Z"] + 15) // 16 * 16
    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=n, max_batch_llrs=64 << 20, harq_slots=n, nof_tickets=1)
    d_bits = torch.zeros(oo, dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(n * 8, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    soft_cpu = np.zeros(so, np.int8)
    if stale is not None:
        for i, c in enumerate(cases):
            buf = np.zeros(ns(c["bg"]) * c["Z"], np.int8)
            buf[stale[0]] = stale[1]
            soft_cpu[int(soft_off[i]) + stale[0]] = stale[1]
            ctx.harq_write(i, buf)
    out = {"config": name, "code_blocks": n, "transmissions": []}
    for t, (rvs, llr_list) in enumerate(txs):
        llr_off = np.zeros(n, np.uint64)
        lo = 0
        for i in range(n):
            llr_off[i] = lo
            lo += (llr_list[i].size + 15) // 16 * 16
        llrs = np.zeros(lo, np.int8)
        for i in range(n):
            llrs[int(llr_off[i]):int(llr_off[i]) + llr_list[i].size] = llr_list[i]
        d = B.make_descs(n)
        rd = np.zeros(n, O.REF_CB_DTYPE)
        for i, c in enumerate(cases):
            d[i]["bg"], d[i]["Z"], d[i]["Qm"], d[i]["rv"], d[i]["crc_type"], d[i]["max_iter"] = c["bg"], c["Z"], c["Qm"], rvs[i], c["crc_id"], 6
            d[i]["flags"] = (B.PDC_F_NEW_DATA if t == 0 else 0) | B.PDC_F_EARLY_STOP
            d[i]["E"], d[i]["Nref"], d[i]["nof_filler"], d[i]["harq_slot"] = llr_list[i].size, c["Nref"], c["F"], i
            d[i]["llr_offset"], d[i]["out_offset"] = llr_off[i], out_off[i]
            r = rd[i]
            r["bg"], r["Z"], r["Qm"], r["rv"], r["crc_id"], r["max_iter"] = c["bg"], c["Z"], c["Qm"], rvs[i], c["crc_id"], 6
            r["new_data"], r["early_stop"], r["E"], r["Nref"], r["nof_filler"] = int(t == 0), 1, llr_list[i].size, c["Nref"], c["F"]
            r["nof_crc_bits"] = 16 if c["crc_id"] == 1 else 24
            r["llr_offset"], r["soft_offset"], r["out_offset"] = llr_off[i], soft_off[i], out_off[i]
        d_llrs = torch.from_numpy(llrs).cuda()
        # Timed repetitions restart from the same HARQ state (the arena content at this transmission's start). That matters
        # even for new data: with rv != 0 and k0 beyond the systematic part the reference does not clear [n_sys, k0) and
        # combines the wrapped LLRs with whatever is there (allot_llrs :148-158) - repeated runs would accumulate.
        snap = [ctx.harq_read(i, ns(c["bg"]) * c["Z"]) for i, c in enumerate(cases)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms = host_us = 0.0
        for rep in range(steps + 1):
            for i in range(n):
                if snap[i].any():
                    ctx.harq_write(i, snap[i])
                else:
                    ctx.harq_free(i)  # zero-filled and known to the runtime as empty (limited-buffer first transmissions)
            torch.cuda.synchronize()
            e0.record()
            t_host = time.perf_counter()
            ctx.decode_device(d, d_llrs.data_ptr(), None, d_bits.data_ptr(), d_res.data_ptr(), st)
            t_host = time.perf_counter() - t_host
            e1.record()
            torch.cuda.synchronize()
            if rep > 0:
                ms += e0.elapsed_time(e1)
                host_us += 1e6 * t_host
        ms /= steps
        host_us /= steps
        res = d_res.cpu().numpy().view(B.RESULT_DTYPE)
        bits = d_bits.cpu().numpy()
        info = sum(kb(c["bg"]) * c["Z"] - c["F"] - (16 if c["crc_id"] == 1 else 24) for c in cases)
        ops = sum(lane_ops(c["bg"], c["Z"], int(res["iterations"][i]), int(res["nof_layers"][i])) for i, c in enumerate(cases))
        rec = {"tx": t, "gpu_ms": ms, "host_call_us": host_us, "gpu_info_gbps": info / (ms * 1e-3) / 1e9, "value": info / (ms * 1e-3) / 1e9,
               "roofline": {"bound": "alu", "achieved": ops / (ms * 1e-3) / 1e12, "unit": "Tops/s (int8 lane-ops)",
                            "peak": None if peak_ops is None else peak_ops / 1e12,
                            "frac": None if peak_ops is None else ops / (ms * 1e-3) / peak_ops},
               "avg_iterations": float(np.where(res["iterations"] > 0, res["iterations"], 6).mean()),
               "crc_pass_fraction": float((res["crc_pass"] == 1).mean()),
               "layers_min_max": [int(res["nof_layers"].min()), int(res["nof_layers"].max())]}
        if cpu:
            nth = len(os.sched_getaffinity(0))
            dec, dem = O.ref_best_flavours()
            cbits = np.zeros(oo, np.uint8)
            soft_before = soft_cpu.copy()
            sec, it = O.ref_chain_run(dec, dem, nth, rd, llrs, soft_cpu, cbits, 1)
            reps = max(1, int(2.0 / max(sec, 1e-4)))
            scratch = soft_before.copy()
            if t == 0:
                # timing only (repetitions accumulate in the scratch buffer exactly as the reference would)
                sec, _ = O.ref_chain_run(dec, dem, nth, rd, llrs, scratch, np.zeros(oo, np.uint8), reps)
                sec /= reps
            rec["cpu_ms"], rec["cpu_threads"], rec["cpu_info_gbps"] = 1e3 * sec, nth, info / sec / 1e9
            same_it = bool((it == res["iterations"]).all())
            same_bits = all((bits[int(out_off[i]):int(out_off[i]) + (kb(c["bg"]) * c["Z"] + 7) // 8] ==
                             cbits[int(out_off[i]):int(out_off[i]) + (kb(c["bg"]) * c["Z"] + 7) // 8]).all()
                            for i, c in enumerate(cases))
            same_soft = all((ctx.harq_read(i, ns(c["bg"]) * c["Z"]) ==
                             soft_cpu[int(soft_off[i]):int(soft_off[i]) + ns(c["bg"]) * c["Z"]]).all()
                            for i, c in enumerate(cases))
            rec["bit_exact"] = {"iterations": same_it, "bits": bool(same_bits), "soft_buffers": bool(same_soft)}
            if not (same_it and same_bits and same_soft):
                for i, c in enumerate(cases):
                    N = ns(c["bg"]) * c["Z"]
                    gs = ctx.harq_read(i, N)
                    cs = soft_cpu[int(soft_off[i]):int(soft_off[i]) + N]
                    if it[i] != res["iterations"][i] or (gs != cs).any():
                        diff = np.nonzero(gs != cs)[0]
                        print("DIFF", i, c, "E", llr_list[i].size, "rv", rvs[i], "it cpu/gpu", int(it[i]), int(res["iterations"][i]),
                              "L", int(res["nof_layers"][i]), "soft diffs", diff.size, diff[:6], gs[diff[:6]], cs[diff[:6]],
                              file=sys.stderr)
                raise RuntimeError(f"{name} tx {t}: GPU differs from the reference: {rec['bit_exact']}")
        out["transmissions"].append(rec)
    ctx.close()
    return out


def config3(S, steps, cpu=True, peak_ops=None):
    from oracle.pyoracle import LIFTING_SIZES

    rng = np.random.default_rng(3)
    cases, rvs, llrs = [], [], []
    for zi, Z in enumerate(LIFTING_SIZES):
        K, N = 10 * Z, 50 * Z
        for k in range(8):
            Qm = (2, 4, 6, 8)[(zi + k) % 4]
            rate = (120, 308, 602)[(zi + 2 * k) % 3] / 1024.0
            rv = (0, 1, 2, 3)[k % 4] if k >= 4 else 0  # a self-decodable majority, as a scheduler would send
            F = int(rng.integers(0, max(1, Z // 2))) if K - Z // 2 > 40 else 0
            if K - F - 16 < 8:
                continue
            E = max(Qm, int((K - F) / rate) // Qm * Qm)
            msg = message_with_crc(S, rng, K, F, 1)
            cb = S.encode(2, Z, msg)
            tx = S.rate_match(2, Z, cb, rv, Qm, 0, F, E)
            x = 8.0 * (1 - 2.0 * tx) + rng.normal(0, 6.0 if rate < 0.4 else 3.5, E)
            cases.append(dict(bg=2, Z=Z, Qm=Qm, F=F, Nref=0, crc_id=1))
            rvs.append(rv)
            llrs.append(np.clip(np.round(x), -120, 120).astype(np.int8))
    order = rng.permutation(len(cases))
    cases = [cases[i] for i in order]
    rvs = [rvs[i] for i in order]
    llrs = [llrs[i] for i in order]
    return run_batch("config 3: BG2 sweep, all 51 lifting sizes x 8 code blocks, mixed Qm / rate / rv, shuffled", cases,
                     [(rvs, llrs)], steps, cpu=cpu, peak_ops=peak_ops)


def config4(S, steps, sigma1, sigma2, n_tbs=8, cpu=True, peak_ops=None):
    """n_tbs config-2 transport blocks (152 code blocks each): rv0 at a noise level where nothing decodes (six iterations
    everywhere: the worst case), then rv2 combined into the device-resident soft buffers."""
    c = S.CONFIG2
    seg = S.segmentation(c["bg"], c["tbs"])
    Es = S.rm_lengths(seg["C"], c["G"], c["nof_layers"], c["Qm"])
    cases = [dict(bg=1, Z=seg["Z"], Qm=8, F=seg["F"], Nref=0, crc_id=3) for _ in range(seg["C"] * n_tbs)]
    txs = [([], []), ([], [])]
    for k in range(n_tbs):
        rng = np.random.default_rng(4 + 100 * k)
        tb = rng.integers(0, 2, c["tbs"], dtype=np.uint8)
        for j, (rv, sigma, seed) in enumerate(((0, sigma1, 41), (2, sigma2, 42))):
            _, bits, _ = S.tb_encode(c["bg"], c["tbs"], tb, rv, c["Qm"], 0, c["nof_layers"], c["G"])
            x = S.awgn_llrs(bits, 8.0, sigma, seed + 1000 * k)
            p = 0
            for r in range(seg["C"]):
                txs[j][1].append(x[p:p + Es[r]])
                txs[j][0].append(rv)
                p += Es[r]
    return run_batch(f"config 4: HARQ rv0 (sigma {sigma1}) + rv2 (sigma {sigma2}) over the device soft buffer, {n_tbs} TBs x 152 "
                     f"code blocks BG1 Z=384", cases, txs, steps, cpu=cpu, peak_ops=peak_ops)


def config2_lbrm(S, steps, sigma=2.2, n_tbs=8, cpu=True, peak_ops=None):
    """BASELINE config 2' (the live stack's limited-buffer rate matching, Nref = 12611 < N = 25344,
    pusch_processor_impl.cpp:327-334): n_tbs config-2 transport blocks, first transmissions into released HARQ arena slots, and
    the same batch with Nref = 0 beside it."""
    c = S.CONFIG2
    seg = S.segmentation(c["bg"], c["tbs"])
    Es = S.rm_lengths(seg["C"], c["G"], c["nof_layers"], c["Qm"])
    out = None
    for nref in (12611, 0):
        cases = [dict(bg=1, Z=seg["Z"], Qm=8, F=seg["F"], Nref=nref, crc_id=3) for _ in range(seg["C"] * n_tbs)]
        txs = [([], [])]
        for k in range(n_tbs):
            rng = np.random.default_rng(5 + 100 * k)
            tb = rng.integers(0, 2, c["tbs"], dtype=np.uint8)
            _, bits, _ = S.tb_encode(c["bg"], c["tbs"], tb, 0, c["Qm"], nref, c["nof_layers"], c["G"])
            x = S.awgn_llrs(bits, 8.0, sigma, 51 + 1000 * k)
            p = 0
            for r in range(seg["C"]):
                txs[0][1].append(x[p:p + Es[r]])
                txs[0][0].append(0)
                p += Es[r]
        r = run_batch(f"config 2': limited-buffer rate matching (Nref = 12611), {n_tbs} TBs x 152 code blocks BG1 Z=384, first "
                      f"transmissions into released arena slots", cases, txs, steps, cpu=cpu and nref != 0, peak_ops=peak_ops)
        if nref != 0:
            out, txs_lbrm = r, txs
        else:
            out["same_batch_with_Nref_0"] = {k: r["transmissions"][0][k] for k in ("gpu_ms", "value", "avg_iterations")}
    # The same limited-buffer batch into slots that hold ONE stale soft bit in the stretch the reference leaves untouched
    # ([end, N - (Ncb - end)), allot_llrs :198-200): the decoder sees it, 35 layers take part, the general kernel runs.
    cases = [dict(bg=1, Z=seg["Z"], Qm=8, F=seg["F"], Nref=12611, crc_id=3) for _ in range(seg["C"] * n_tbs)]
    r = run_batch("config 2' with one stale soft bit at position 20000 of every slot", cases, txs_lbrm, max(2, steps // 2), cpu=cpu,
                  peak_ops=peak_ops, stale=(20000, 5))
    t = r["transmissions"][0]
    out["stale_slots"] = {k: t[k] for k in ("gpu_ms", "value", "avg_iterations", "layers_min_max", "bit_exact") if k in t}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--sigma1", type=float, default=3.0)
    ap.add_argument("--sigma2", type=float, default=6.0)
    ap.add_argument("--tbs", type=int, default=8, help="config 4: transport blocks in the batch")
    ap.add_argument("--only", type=int, default=0, help="2, 3 or 4: run that configuration alone")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    from tools import stimulus as S

    if args.only in (0, 2):
        print(json.dumps(config2_lbrm(S, args.steps, cpu=not args.no_cpu)), flush=True)
    if args.only in (0, 3):
        print(json.dumps(config3(S, args.steps, cpu=not args.no_cpu)), flush=True)
    if args.only in (0, 4):
        print(json.dumps(config4(S, args.steps, args.sigma1, args.sigma2, args.tbs, cpu=not args.no_cpu)), flush=True)


if __name__ == "__main__":
    main()
