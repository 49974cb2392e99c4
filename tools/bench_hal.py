#!/usr/bin/env python3
"""Throughput of the drop-in boundary (bench.py leg `e2e_hal`): T host threads, each running the UNMODIFIED reference
pusch_decoder_hw_impl (compiled from /root/reference into tests/adapters/_build/libadapter_harness.so) over the CUDA
hal::hw_accelerator_pusch_dec, on BASELINE config-2 transport blocks, the way pusch_decoder_hwacc_benchmark.cpp:430-495 drives a
decoder (new_data + one on_new_softbits + on_end_softbits per transport block). Everything between the reference's decoder
interface and the decoded transport block is inside the timed region: the reference's segmenter, soft-bit buffering, host CRC
checks and TB assembly, the adapter's copies into page-locked staging, H2D, the kernels, D2H.

Every decoded transport block is compared with the transmitted one. Prints one JSON line when run as a script."""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
HARNESS = os.path.join(ROOT, "tests", "adapters", "_build", "libadapter_harness.so")


def available():
    return os.path.exists(HARNESS)


def _lib():
    lib = ctypes.CDLL(HARNESS)
    u, vp = ctypes.c_uint, ctypes.c_void_p
    lib.adt_bench_hal.argtypes = [u, u, u, vp, u, u, vp, u, ctypes.c_int, u, u, u, vp, u, u, u, vp, vp, vp]
    lib.adt_bench_hal.restype = ctypes.c_int
    return lib


def run_raw(llrs, expected, tbs_bits, bg, Qm, nof_layers, max_iter, threads, tbs_per_thread, warmup=2, devices=(0,),
            max_parts=8, close_after_us=150, replay=None, retx=None):
    """llrs: [n_distinct, nof_llrs] int8, expected: [n_distinct, tbs_bits / 8] uint8. replay: None = the CUDA accelerator;
    [n_distinct, C, ceil(K / 8)] uint8 = code-block messages for the harness's replay accelerator (answers at once: measures the
    host cost of the reference's pusch_decoder_hw_impl alone). retx: None, or (llrs2 [n_distinct, nof_llrs] int8, rv2): a transport
    block whose first transmission fails its CRC is sent again as redundancy version rv2 and combined in the accelerator's soft
    buffer (HARQ). -> dict(seconds, ok, attempted, batches, code_blocks, retransmitted)."""
    rp = None
    if replay is not None:
        replay = np.ascontiguousarray(replay, np.uint8)
        rp = replay.ctypes.data
    llrs = np.ascontiguousarray(llrs, np.int8)
    expected = np.ascontiguousarray(expected, np.uint8)
    dev = np.asarray(devices, np.int32)
    sec = ctypes.c_double()
    stats = np.zeros(4, np.uint64)
    lib = _lib()
    lib.adt_bench_hal_set_retx.argtypes = [ctypes.c_void_p, ctypes.c_uint]
    lib.adt_bench_hal_retransmitted.restype = ctypes.c_uint64
    llrs2 = None
    if retx is not None:
        llrs2 = np.ascontiguousarray(retx[0], np.int8)
        assert llrs2.shape == llrs.shape
        lib.adt_bench_hal_set_retx(llrs2.ctypes.data, int(retx[1]))
    else:
        lib.adt_bench_hal_set_retx(None, 2)
    rc = lib.adt_bench_hal(threads, tbs_per_thread, warmup, llrs.ctypes.data, llrs.shape[1], llrs.shape[0],
                              expected.ctypes.data, tbs_bits, bg, Qm, nof_layers, max_iter, dev.ctypes.data, dev.size, max_parts,
                              close_after_us, rp, ctypes.byref(sec), stats.ctypes.data)
    lib.adt_bench_hal_set_retx(None, 2)
    if rc != 0:
        raise RuntimeError(f"adt_bench_hal failed: {rc} (-1: no CUDA device, -2: soft buffer pool exhausted)")
    return dict(seconds=sec.value, ok=int(stats[0]), attempted=int(stats[1]), batches=int(stats[2]), code_blocks=int(stats[3]),
                retransmitted=int(lib.adt_bench_hal_retransmitted()))


def config2_inputs(n_distinct, sigma=2.2, seed0=100):
    from tools import stimulus as S

    c = S.CONFIG2
    llrs = np.zeros((n_distinct, c["G"]), np.int8)
    tbs = np.zeros((n_distinct, c["tbs"] // 8), np.uint8)
    msgs = None
    for t in range(n_distinct):
        rng = np.random.default_rng(seed0 + t)
        tb = rng.integers(0, 2, c["tbs"], dtype=np.uint8)
        _, bits, m = S.tb_encode(c["bg"], c["tbs"], tb, c["rv"], c["Qm"], 0, c["nof_layers"], c["G"])
        llrs[t] = S.awgn_llrs(bits, 8.0, sigma, 7919 * (seed0 + t) + 1)
        tbs[t] = np.packbits(tb)
        if msgs is None:
            msgs = np.zeros((n_distinct, m.shape[0], (m.shape[1] + 7) // 8), np.uint8)
        msgs[t] = np.packbits(m, axis=1)
    return c, llrs, tbs, msgs


def run(threads, tbs_per_thread=24, n_distinct=16, devices=(0,), max_parts=8, close_after_us=150, inputs=None):
    """The bench leg: config-2 transport blocks. -> dict for the JSON line."""
    c, llrs, tbs, msgs = inputs if inputs is not None else config2_inputs(n_distinct)
    r = run_raw(llrs, tbs, c["tbs"], c["bg"], c["Qm"], c["nof_layers"], 6, threads, tbs_per_thread, devices=devices,
                max_parts=max_parts, close_after_us=close_after_us)
    if r["ok"] != r["attempted"]:
        raise RuntimeError(f"e2e_hal: {r['attempted'] - r['ok']} of {r['attempted']} transport blocks differ from the transmitted ones")
    nb = max(r["batches"], 1)
    # The same with limited-buffer rate matching as the live stack always configures it (Nref = 12611 for this transport block:
    # pusch_processor_impl.cpp:327-334, SURVEY 8 config 2'); a first transmission's LLRs are the same with and without the limit.
    os.environ["ADT_NREF"] = "12611"
    try:
        rl = run_raw(llrs, tbs, c["tbs"], c["bg"], c["Qm"], c["nof_layers"], 6, threads, tbs_per_thread, devices=devices,
                     max_parts=max_parts, close_after_us=close_after_us)
    finally:
        del os.environ["ADT_NREF"]
    if rl["ok"] != rl["attempted"]:
        raise RuntimeError(f"e2e_hal (Nref = 12611): {rl['attempted'] - rl['ok']} of {rl['attempted']} transport blocks differ")
    # The same threads over an accelerator that answers at once: what the reference's own host code (soft-bit buffering,
    # segmentation, per-code-block CRC checks, bit-granular TB assembly, TB CRC) allows through this boundary on this host.
    h = run_raw(llrs, tbs, c["tbs"], c["bg"], c["Qm"], c["nof_layers"], 6, threads, tbs_per_thread, replay=msgs)
    if h["ok"] != h["attempted"]:
        raise RuntimeError("e2e_hal: the replay accelerator run did not reproduce the transport blocks")
    ceiling = h["attempted"] * c["tbs"] / h["seconds"] / 1e9
    return {"value": r["attempted"] * c["tbs"] / r["seconds"] / 1e9, "unit": "Gbit/s", "host_threads": threads,
            "value_limited_buffer_Nref_12611": rl["attempted"] * c["tbs"] / rl["seconds"] / 1e9,
            "host_ceiling": ceiling, "frac_of_host_ceiling": r["attempted"] * c["tbs"] / r["seconds"] / 1e9 / ceiling,
            "host_ceiling_source": "same threads and transport blocks through the reference's pusch_decoder_hw_impl over a replay "
                                   "accelerator that answers at once (tests/adapters/harness.cpp): the host cost of the boundary",
            "transport_blocks": r["attempted"], "seconds": r["seconds"], "gpu_submissions": r["batches"],
            "code_blocks_per_submission": r["code_blocks"] / nb, "devices": list(devices),
            "h2d_bytes_per_tb": int(llrs.shape[1]), "d2h_bytes_per_tb": int(152 * (1056 + 8)),
            "api": "reference pusch_decoder_hw_impl (unmodified, compiled from the reference) -> hal::hw_accelerator_pusch_dec "
                   "(CUDA, external HARQ, cross-instance batching); every transport block compared with the transmitted one"}


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
    ap.add_argument("--tbs-per-thread", type=int, default=24)
    ap.add_argument("--max-parts", type=int, default=8)
    ap.add_argument("--close-after-us", type=int, default=150)
    ap.add_argument("--devices", default="0")
    a = ap.parse_args()
    print(json.dumps(run(a.threads, a.tbs_per_thread, devices=tuple(int(x) for x in a.devices.split(",")),
                         max_parts=a.max_parts, close_after_us=a.close_after_us)))
