"""LLR producers at the size of one BASELINE config-2 slot: n_tbs transport blocks x 13 OFDM symbols x 13104
256QAM symbols per demodulate_soft call. Prints one JSON object:

  demod            device-resident pdc_demodulate_device (soft demodulation + descrambling), G LLR/s and the HBM roofline
                   (algorithmic bytes: 8 B symbol + 4 B noise variance in, 8 B of LLRs out, per 256QAM symbol)
  symbols_to_tbs   equaliser outputs in HBM -> demodulate -> decode + CRC -> transport-block assembly, decoded info Gbit/s
  cpu              the compiled reference's demodulation_mapper_impl + pseudo_random_generator_impl on the host cores
                   (one call per OFDM symbol, calls spread over threads), only if oracle/_ref is present

Every run checks the first and the last segment of the LLRs against the CPU oracle and the payloads against the transmitted
transport blocks; a mismatch raises instead of printing. Usage: python tools/bench_demod.py [--tbs 64] [--steps 20]."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def build(n_tbs, sigma2=0.0016, seed=21, distinct=4):
    """`distinct` different transport blocks tiled to n_tbs (host preparation time; the device sees n_tbs full inputs)."""
    from tools.demod_stimulus import config2_transmission

    base = config2_transmission(min(distinct, n_tbs), sigma2, seed)
    if n_tbs <= distinct:
        return base
    from srsran_project_b200 import binding as B

    reps = -(-n_tbs // distinct)
    nsym = base["symbols"].shape[0] // distinct
    w = dict(base)
    w["symbols"] = np.tile(base["symbols"], (reps, 1))[:n_tbs * nsym]
    w["noise"] = np.tile(base["noise"], reps)[:n_tbs * nsym]
    nseg, C = base["segs"].size // distinct, base["C"]
    segs, descs, tbd = np.tile(base["segs"], reps)[:n_tbs * nseg], np.tile(base["descs"], reps)[:n_tbs * C], np.tile(base["tbd"], reps)[:n_tbs]
    for t in range(n_tbs):
        q = t // distinct
        segs["sym_offset"][t * nseg:(t + 1) * nseg] += q * distinct * nsym
        segs["llr_offset"][t * nseg:(t + 1) * nseg] += q * distinct * nsym * 8
        descs["llr_offset"][t * C:(t + 1) * C] += q * distinct * base["G"]
        descs["soft_offset"][t * C:(t + 1) * C] = (t * C + np.arange(C)) * 25344
        descs["out_offset"][t * C:(t + 1) * C] = (t * C + np.arange(C)) * base["stride"]
        tbd["first_cb"][t], tbd["tb_offset"][t] = t * C, t * base["tb_stride"]
    w["segs"], w["descs"], w["tbd"] = segs, descs, tbd
    w["tbs"] = [base["tbs"][t % distinct] for t in range(n_tbs)]
    return w


def run(ctx, n_tbs=64, steps=20, warmup=3, cpu=True, hbm_peak_gbs=None):
    import torch

    from srsran_project_b200 import binding as B

    pyoracle = None
    if cpu:  # the oracle is only ever the checker / the baseline of the CPU leg
        from oracle import pyoracle

    w = build(n_tbs)
    nsym = w["noise"].size
    # Two input sets (2 x 131 MB at 64 TBs) + 87 MB of LLRs written per step: every step streams from HBM (L2 = 126 MB).
    d_sym = [torch.from_numpy(w["symbols"]).cuda(), torch.from_numpy(w["symbols"]).cuda()]
    d_nv = [torch.from_numpy(w["noise"]).cuda(), torch.from_numpy(w["noise"]).cuda()]
    n_cbs = w["descs"].size
    d_llrs = torch.zeros(nsym * 8, dtype=torch.int8, device="cuda")
    d_soft = torch.zeros(n_cbs * 25344, dtype=torch.int8, device="cuda")
    d_bits = torch.zeros(n_cbs * w["stride"], dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(n_cbs * 8, dtype=torch.uint8, device="cuda")
    d_tb = torch.zeros(n_tbs * w["tb_stride"], dtype=torch.uint8, device="cuda")
    d_tbres = torch.zeros(n_tbs * 4, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def demod(k):
        ctx.demodulate_device(w["segs"], d_sym[k & 1].data_ptr(), d_nv[k & 1].data_ptr(), d_llrs.data_ptr(), stream)

    def chain(k):
        demod(k)
        ctx.decode_device_tbs(w["descs"], w["tbd"], d_llrs.data_ptr(), d_soft.data_ptr(), d_bits.data_ptr(), d_res.data_ptr(),
                              d_tb.data_ptr(), d_tbres.data_ptr(), stream)

    def timed(fn):
        for k in range(warmup):
            fn(k)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(steps):
            fn(k)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    t_heat = time.perf_counter()
    while time.perf_counter() - t_heat < 0.2:
        demod(0)
        torch.cuda.synchronize()
    ms_demod = timed(demod)
    # Parity against the CPU oracle (first and last call) belongs to the CPU-baseline leg: the oracle is only ever the checker.
    if cpu:
        orc = pyoracle.DemodOracle()
        llrs = d_llrs.cpu().numpy()
        for sg in (w["segs"][0], w["segs"][-1]):
            a, n = int(sg["sym_offset"]), int(sg["nof_symbols"])
            want = orc.descramble(orc.demodulate_soft(w["symbols"][a:a + n], w["noise"][a:a + n], 8), int(sg["c_init"]),
                                  int(sg["scramble_offset"]))
            o = int(sg["llr_offset"])
            if not np.array_equal(llrs[o:o + want.size], want):
                raise RuntimeError("device LLRs differ from the oracle")
    ms_chain = timed(chain)
    tres = d_tbres.cpu().numpy().view(B.TB_RESULT_DTYPE)
    out = d_tb.cpu().numpy().reshape(n_tbs, -1)[:, :w["tbs_bytes"]]
    if not (tres["tb_crc_ok"] == 1).all() or any(not np.array_equal(out[t], w["tbs"][t]) for t in range(n_tbs)):
        raise RuntimeError("transport blocks decoded from device-born LLRs differ from the transmitted ones")
    it = d_res.cpu().numpy().view(B.RESULT_DTYPE)["iterations"]
    alg_bytes = nsym * (8 + 4 + 8)
    gbs = alg_bytes / (ms_demod * 1e-3) / 1e9
    res = {
        "workload": f"{n_tbs} TBs x 13 calls x 13104 256QAM symbols (config 2), descrambling on",
        "demod": {"ms_per_step": ms_demod, "gllr_per_s": nsym * 8 / (ms_demod * 1e-3) / 1e9,
                  "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak_gbs, "unit": "GB/s",
                               "frac": (gbs / hbm_peak_gbs) if hbm_peak_gbs else None,
                               "algorithmic_bytes_per_step": alg_bytes}},
        "symbols_to_tbs": {"ms_per_step": ms_chain, "gbit_per_s": n_tbs * 1277992 / (ms_chain * 1e-3) / 1e9,
                           "avg_iterations": float(it.mean()), "kernels": "pdc_demod_kernel + code-block kernel + 2 assembly kernels"},
        "parity": ("LLRs of the first and last call bit-exact vs oracle; " if cpu else "") +
                  "all payloads decoded from the device-born LLRs equal to the transmitted TBs",
    }
    res["ulsch_demux"] = run_demux(ctx, n_tbs, w["G"], d_llrs, timed, cpu, hbm_peak_gbs)
    if cpu and pyoracle.ref_available():
        from concurrent.futures import ThreadPoolExecutor

        ref = pyoracle.DemodReference()
        nthreads = os.cpu_count() or 1
        segs = w["segs"][:min(w["segs"].size, 8 * nthreads)]

        def one(sg):
            a, n = int(sg["sym_offset"]), int(sg["nof_symbols"])
            return ref.descramble(ref.demodulate_soft(w["symbols"][a:a + n], w["noise"][a:a + n], 8), int(sg["c_init"]),
                                  int(sg["scramble_offset"]))

        one(segs[0])
        t0 = time.perf_counter()
        reps = 0
        with ThreadPoolExecutor(nthreads) as ex:
            while time.perf_counter() - t0 < 3.0:
                list(ex.map(one, segs))
                reps += 1
        dt = time.perf_counter() - t0
        res["cpu"] = {"kind": "reference", "cores": nthreads, "gllr_per_s": reps * segs.size * 13104 * 8 / dt / 1e9,
                      "sample": f"{segs.size} calls x {reps} repetitions, {dt:.1f} s wall, demodulation_mapper_impl (AVX512) + "
                                "pseudo_random_generator_impl apply_xor"}
    return res


def run_demux(ctx, n_tbs, G, d_llrs, timed, cpu, hbm_peak_gbs):
    """UL-SCH demultiplexer at the same size: every codeword of the slot (273 PRB, 4 layers, 256QAM, DM-RS in symbol 2 with no
    data REs) carries a 2-bit HARQ-ACK on reserved REs, CSI Part 1 and CSI Part 2; the input is the demodulator's output in HBM."""
    import torch

    from srsran_project_b200 import binding as B

    nbpr = 32
    descs = np.zeros(n_tbs, B.ULSCH_DEMUX_DTYPE)
    descs["Qm"], descs["nof_layers"], descs["nof_prb"], descs["nof_symbols"] = 8, 4, 273, 14
    descs["dmrs_type"], descs["nof_cdm_groups_without_data"], descs["dmrs_symbol_mask"] = 1, 2, 1 << 2
    descs["nof_harq_ack_bits"], descs["nof_enc_harq_ack_bits"], descs["nof_harq_ack_rvd"] = 2, 40 * nbpr, 60 * nbpr
    descs["nof_csi_part1_bits"], descs["nof_enc_csi_part1_bits"] = 30, 150 * nbpr
    descs["nof_csi_part2_bits"], descs["nof_enc_csi_part2_bits"] = 50, 260 * nbpr
    descs["c_init"] = 0x5a5a + np.arange(n_tbs)
    lens = B.ulsch_demux_lengths(descs[:1])
    assert lens[0] == G
    stride = sum((n + 15) // 16 * 16 for n in lens[1:])
    for t in range(n_tbs):
        o = t * stride
        descs[t]["in_offset"] = t * G
        for f, n in zip(("sch_offset", "harq_ack_offset", "csi_part1_offset", "csi_part2_offset"), lens[1:]):
            descs[t][f] = o
            o += (n + 15) // 16 * 16
    d_out = torch.zeros(n_tbs * stride, dtype=torch.int8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    ms = timed(lambda k: ctx.ulsch_demultiplex_device(descs, d_llrs.data_ptr(), d_out.data_ptr(), stream))
    alg = n_tbs * (lens[0] + sum(lens[1:]))
    gbs = alg / (ms * 1e-3) / 1e9
    out = {"workload": f"{n_tbs} codewords x 13 OFDM symbols x 3276 REs x 32 soft bits; HARQ-ACK 2 bits on 40 of 60 reserved REs, "
                       "CSI Part 1 on 150 REs, CSI Part 2 on 260 REs",
           "ms_per_step": ms, "gllr_per_s": n_tbs * lens[0] / (ms * 1e-3) / 1e9,
           "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak_gbs, "unit": "GB/s",
                        "frac": (gbs / hbm_peak_gbs) if hbm_peak_gbs else None, "algorithmic_bytes_per_step": alg}}
    if cpu:  # the restatement is only ever the checker: last codeword of the slot, all four streams
        from oracle import demux_oracle as D

        t = n_tbs - 1
        cfg = {"qm": 8, "nof_layers": 4, "nof_prb": 273, "start_symbol": 0, "nof_symbols": 14, "dmrs_type": 1,
               "dmrs_symbol_mask": 1 << 2, "nof_cdm_groups_without_data": 2, "nof_harq_ack_bits": 2,
               "nof_enc_harq_ack_bits": 40 * nbpr, "nof_harq_ack_rvd": 60 * nbpr, "nof_csi_part1_bits": 30,
               "nof_enc_csi_part1_bits": 150 * nbpr, "nof_csi_part2_bits": 50, "nof_enc_csi_part2_bits": 260 * nbpr,
               "csi2_mode": 0, "c_init": int(descs[t]["c_init"])}
        want = D.demultiplex(cfg, d_llrs[t * G:(t + 1) * G].cpu().numpy())
        got = d_out.cpu().numpy()
        for f, b in zip(("sch_offset", "harq_ack_offset", "csi_part1_offset", "csi_part2_offset"), want):
            if not np.array_equal(got[int(descs[t][f]):int(descs[t][f]) + b.size], b):
                raise RuntimeError("device UL-SCH demultiplexer differs from the restatement (" + f + ")")
        out["parity"] = "four streams of the last codeword bit-exact vs the restatement"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tbs", type=int, default=64)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import srsran_project_b200 as pdc

    peak = None
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p)).get("hbm_gbs")
    n_cbs = args.tbs * 152
    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=n_cbs, max_batch_llrs=1 << 20, harq_slots=0, nof_tickets=2)
    print(json.dumps(run(ctx, args.tbs, args.steps, args.warmup, not args.no_cpu, peak)))
    ctx.close()


if __name__ == "__main__":
    main()
