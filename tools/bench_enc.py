"""PDSCH encoder at the size of one BASELINE config-2 slot mirrored to the downlink: n_tbs transport blocks of
1 277 968 bits (152 code blocks, BG1, Z = 384), 256QAM, 4 layers, 170 352 channel symbols. Prints one JSON object:

  encode   device-resident pdc_pdsch_encode_device, encoded information Gbit/s and the HBM roofline (algorithmic bytes per
           transport block: tbs / 8 read, nof_ch_symbols * Qm bytes written — the reference's codeword is one bit per byte)
  cpu      the compiled reference's pdsch_encoder_impl (ldpc_encoder_avx2) on the host cores, one transport block per thread

The first and the last codeword of the batch are compared with the TX-chain restatement; a mismatch raises."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CASE = dict(bg=1, tbs=1277968, Qm=8, L=4, nsym=170352, rv=0)


def run(ctx, n_tbs=64, steps=20, warmup=3, cpu=True, hbm_peak_gbs=None):
    import torch

    from srsran_project_b200 import binding as B

    pyoracle = None
    if cpu:  # the compiled reference is only ever the baseline of the CPU leg
        from oracle import pyoracle

    c = CASE
    nb, G = c["tbs"] // 8, c["nsym"] * c["Qm"]
    data = np.random.default_rng(77).integers(0, 256, n_tbs * nb, dtype=np.uint8)
    tbs = np.zeros(n_tbs, B.PDSCH_TB_DTYPE)
    tbs["bg"], tbs["Qm"], tbs["rv"], tbs["nof_layers"] = c["bg"], c["Qm"], c["rv"], c["L"]
    tbs["tbs_bits"], tbs["nof_ch_symbols"] = c["tbs"], c["nsym"]
    tbs["tb_offset"], tbs["cw_offset"] = np.arange(n_tbs) * nb, np.arange(n_tbs) * G
    d_tb = torch.from_numpy(data).cuda()
    d_cw = [torch.zeros(n_tbs * G, dtype=torch.uint8, device="cuda") for _ in range(2)]  # 2 x 87 MB: outputs leave L2
    stream = torch.cuda.current_stream().cuda_stream

    def step(k):
        ctx.pdsch_encode_device(tbs, d_tb.data_ptr(), d_cw[k & 1].data_ptr(), stream)

    t_heat = time.perf_counter()
    while time.perf_counter() - t_heat < 0.2:
        step(0)
        torch.cuda.synchronize()
    for k in range(warmup):
        step(k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(steps):
        step(k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    cw = d_cw[(steps - 1) & 1].cpu().numpy()
    from tools import stimulus as S

    for t in (0, n_tbs - 1):  # the TX-chain restatement of the stimulus generator (tools/, pinned against the reference in tests/)
        _, want, _ = S.tb_encode(c["bg"], c["tbs"], np.unpackbits(data[t * nb:(t + 1) * nb]), c["rv"], c["Qm"], 0, c["L"], G)
        if not np.array_equal(cw[t * G:(t + 1) * G], want):
            raise RuntimeError("device codeword differs from the TX-chain restatement")
    alg = n_tbs * (nb + G)
    gbs = alg / (ms * 1e-3) / 1e9
    res = {"workload": f"{n_tbs} TBs x {c['tbs']} bits (152 code blocks, BG1 Z=384), 256QAM, 4 layers, rv0",
           "encode": {"ms_per_step": ms, "gbit_per_s": n_tbs * c["tbs"] / (ms * 1e-3) / 1e9,
                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak_gbs, "unit": "GB/s",
                                   "frac": (gbs / hbm_peak_gbs) if hbm_peak_gbs else None, "algorithmic_bytes_per_step": alg},
                      "kernels": "pdc_enc_crc_kernel + pdc_enc_cb_kernel"},
           "parity": "first and last codeword bit-exact vs the TX-chain restatement"}
    if cpu and pyoracle.ref_available():
        from concurrent.futures import ThreadPoolExecutor

        nthreads = os.cpu_count() or 1
        refs = [pyoracle.PdschEncReference() for _ in range(nthreads)]

        def one(i):
            t = i % n_tbs
            return refs[i % nthreads].encode(data[t * nb:(t + 1) * nb], c["tbs"], c["bg"], c["rv"], c["Qm"], 0, c["L"], c["nsym"])

        one(0)
        t0 = time.perf_counter()
        done = 0
        with ThreadPoolExecutor(nthreads) as ex:
            while time.perf_counter() - t0 < 3.0:
                list(ex.map(one, range(nthreads)))
                done += nthreads
        dt = time.perf_counter() - t0
        res["cpu"] = {"kind": "reference", "cores": nthreads, "gbit_per_s": done * c["tbs"] / dt / 1e9,
                      "sample": f"{done} transport blocks, {dt:.1f} s wall, pdsch_encoder_impl (ldpc_encoder_avx2)"}
    return res


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
    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=256, max_batch_llrs=1 << 20, harq_slots=0, nof_tickets=2)
    print(json.dumps(run(ctx, args.tbs, args.steps, args.warmup, not args.no_cpu, peak)))
    ctx.close()


if __name__ == "__main__":
    main()
