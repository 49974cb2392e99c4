"""Generates tests/golden/pusch_cb_chain_v1.npz from the COMPILED, UNMODIFIED reference (oracle/_ref/libsrsran_ref.so,
built by oracle/build_ref.sh from /root/reference): the reference's own golden vector tarballs are absent from the
checkout, so the pins are the reference's outputs on seeded inputs.

For every case the file stores the inputs (configuration, initial soft-buffer content, the LLRs of each transmission)
and, for BOTH reference flavours (ldpc_decoder_avx512 + ldpc_rate_dematcher_avx512_impl, ldpc_decoder_generic +
ldpc_rate_dematcher_impl) what pusch_codeblock_decoder (pusch_codeblock_decoder.cpp:26-74) / ldpc_decoder::decode
(ldpc_decoder_impl.cpp:57-148) returned: the whole soft buffer after every transmission, the K output bits (also of
failing blocks), and the iteration count (-1 = std::nullopt).

Run in the build container only (needs /root/reference once, to compile oracle/_ref):
    python -m tools.gen_golden
The tests (tests/test_golden.py) read the .npz; they never need the reference.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import pyoracle as O  # noqa: E402  (generator = test infrastructure)
from tools import stimulus as S  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "pusch_cb_chain_v1.npz")
MAX_TX = 3

CASE_DTYPE = np.dtype([("bg", np.int32), ("Z", np.int32), ("Qm", np.int32), ("E", np.int32), ("F", np.int32),
                       ("Nref", np.int32), ("crc_id", np.int32), ("max_iter", np.int32), ("early_stop", np.int32),
                       ("decode_only", np.int32), ("force_decoding", np.int32), ("ntx", np.int32),
                       ("rv", np.int32, MAX_TX), ("llr_off", np.int64, MAX_TX), ("soft0_off", np.int64),
                       ("soft_off", np.int64, MAX_TX), ("bits_off", np.int64, MAX_TX)])


def message_with_crc(rng, K, F, crc_id):
    ncrc = 16 if crc_id == O.CRC16 else 24
    payload = rng.integers(0, 2, K - F - ncrc, dtype=np.uint8)
    c = S.crc(crc_id, payload)
    crc_bits = np.array([(c >> (ncrc - 1 - i)) & 1 for i in range(ncrc)], np.uint8)
    return np.concatenate([payload, crc_bits, np.zeros(F, np.uint8)])


def awgn(rng, bits, sigma, amp=8.0):
    x = amp * (1.0 - 2.0 * bits.astype(np.float64)) + rng.normal(0.0, sigma, bits.size)
    return np.clip(np.round(x), -120, 120).astype(np.int8)


def chain_cases(rng):
    """(config dict, [llrs per tx], soft0). One small and one large Z per lifting-size set index, both base graphs,
    every Qm, fillers, LBRM, the four rv, CRC16/24A/24B, passing and failing blocks, E < n_info and E >> Ncb."""
    sets = (2, 3, 5, 7, 9, 11, 13, 15)
    plan = []
    for si, a in enumerate(sets):
        small = a * 2 ** (si % 3)
        large = max(z for z in O.LIFTING_SIZES if z % a == 0 and (z // a) & (z // a - 1) == 0)
        plan += [(small, si), (large, si + 8)]
    plan += [(384, 16), (384, 17), (384, 18), (352, 19), (32, 20), (96, 21)]
    cases = []
    for bg in (1, 2):
        kb, ns = O.kb_of(bg), O.nshort_of(bg)
        for Z, k in plan:
            K, N = kb * Z, ns * Z
            Qm = (1, 2, 4, 6, 8)[k % 5]
            F = 0 if k % 3 == 0 else int(rng.integers(1, min(2 * Z, K - 30)))
            crc_id = O.CRC16 if K - F < 200 else (O.CRC24A, O.CRC24B, O.CRC16)[k % 3]
            Nref = 0 if k % 4 else int(rng.integers((kb + 5) * Z, N))
            if k in (16, 17, 18):  # BASELINE config 2 / 2' / 4 shapes (SURVEY.md section 8 size table)
                Qm, F, crc_id = 8, 16, O.CRC24B
                Nref = 12611 if (k == 17 and bg == 1) else 0
                E = 8960 if bg == 1 else 4000
                rvs, sig = ((0,), (2.2,)) if k != 18 else ((0, 2), (3.2, 3.0))
            elif k == 19:  # E below the systematic part (write-set quirk) and then heavy repetition
                E = (K // 2) // Qm * Qm
                rvs, sig = (0, 1), (4.0, 4.0)
            elif k == 20:  # many laps of the circular buffer
                E = (3 * N + 7 * Qm) // Qm * Qm
                rvs, sig = (3, 0), (12.0, 12.0)
            else:
                rate = (0.92, 0.6, 0.34, 0.2)[k % 4]
                E = max(Qm, int(K / rate) // Qm * Qm)
                rvs = [(0, 2, 3), (0, 3), (0, 1), (0,)][k % 4]
                base = 9.0 if bg == 1 else 7.0
                sig = [base * (0.45 if rate > 0.9 else 1.0)] * len(rvs)
            msg = message_with_crc(rng, K, F, crc_id)
            cb = S.encode(bg, Z, msg)
            llrs = [awgn(rng, S.rate_match(bg, Z, cb, rv, Qm, Nref, F, E), s) for rv, s in zip(rvs, sig)]
            soft0 = rng.integers(-120, 121, N).astype(np.int8)
            cfg = dict(bg=bg, Z=Z, Qm=Qm, E=E, F=F, Nref=Nref, crc_id=crc_id, max_iter=6, early_stop=int(k % 7 != 6),
                       decode_only=0, force_decoding=0, rvs=rvs)
            cases.append((cfg, llrs, soft0))
    return cases


def decode_only_cases(rng):
    """ldpc_decoder::decode alone, including the behavioural known answers of ldpc_enc_dec_test.cpp:293-364."""
    cases = []
    for bg, Z in ((1, 384), (2, 384), (1, 2), (2, 3), (1, 44), (2, 208), (1, 160), (2, 30)):
        kb, ns = O.kb_of(bg), O.nshort_of(bg)
        K, N = kb * Z, ns * Z
        msg = message_with_crc(rng, K, 0, O.CRC16)
        cb = S.encode(bg, Z, msg)
        # noiseless +-10 at three shortened lengths: one iteration
        for n in ((kb + 4) * Z, (kb + (ns - kb) // 2) * Z, N):
            llrs = (10 - 20 * cb[:n].astype(np.int16)).astype(np.int8)
            cases.append((dict(bg=bg, Z=Z, Qm=1, E=n, F=0, Nref=0, crc_id=O.CRC16, max_iter=6, early_stop=1,
                               decode_only=1, force_decoding=0, rvs=(0,)), [llrs], np.zeros(0, np.int8)))
        # rate 1/3 random +-10 (the ldpc_decoder_benchmark stimulus shape), no CRC: runs max_iter
        llrs = (rng.integers(0, 2, N) * 20 - 10).astype(np.int8)
        cases.append((dict(bg=bg, Z=Z, Qm=1, E=N, F=0, Nref=0, crc_id=O.CRC_NONE, max_iter=6, early_stop=0,
                           decode_only=1, force_decoding=0, rvs=(0,)), [llrs], np.zeros(0, np.int8)))
        # noisy, full length, CRC drives the early stop
        llrs = awgn(rng, cb, 7.0 if bg == 2 else 9.0)
        cases.append((dict(bg=bg, Z=Z, Qm=1, E=N, F=0, Nref=0, crc_id=O.CRC16, max_iter=6, early_stop=1,
                           decode_only=1, force_decoding=0, rvs=(0,)), [llrs], np.zeros(0, np.int8)))
        # all-zero and sparse inputs: nullopt; with force_decoding (inverted flag) and no CRC the output is all ones
        sparse = np.zeros(N, np.int8)
        sparse[rng.integers(K + 4 * Z, N, 5)] = 1
        for llrs in (np.zeros(N, np.int8), sparse):
            for force in (0, 1):
                for crc_id in (O.CRC_NONE, O.CRC16):
                    cases.append((dict(bg=bg, Z=Z, Qm=1, E=N, F=0, Nref=0, crc_id=crc_id, max_iter=2,
                                       early_stop=int(crc_id != 0), decode_only=1, force_decoding=force, rvs=(0,)),
                                  [llrs], np.zeros(0, np.int8)))
    return cases


_refs = {}


def run_reference(flavour, cfg, llrs, soft0):
    """-> ([soft after tx], [packed bits after tx], [iterations after tx])"""
    ref = _refs.setdefault(flavour, O.Reference(flavour))
    bg, Z = cfg["bg"], cfg["Z"]
    K, N = O.kb_of(bg) * Z, O.nshort_of(bg) * Z
    softs, bits, iters = [], [], []
    if cfg["decode_only"]:
        it, out = ref.decode(bg, Z, llrs[0], cfg["F"], cfg["max_iter"], cfg["crc_id"], bool(cfg["force_decoding"]))
        return [np.zeros(0, np.int8)], [out], [it]
    dec_fl = dem_fl = flavour
    soft = soft0.copy()
    for tx, (rv, x) in enumerate(zip(cfg["rvs"], llrs)):
        d = np.zeros(1, O.REF_CB_DTYPE)
        d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_id"], d["max_iter"] = bg, Z, cfg["Qm"], rv, cfg["crc_id"], cfg["max_iter"]
        d["new_data"], d["early_stop"], d["E"], d["Nref"], d["nof_filler"] = int(tx == 0), cfg["early_stop"], cfg["E"], \
            cfg["Nref"], cfg["F"]
        d["nof_crc_bits"] = 16 if cfg["crc_id"] == O.CRC16 else 24
        out = np.zeros((K + 7) // 8, np.uint8)
        _, it = O.ref_chain_run(dec_fl, dem_fl, 1, d, np.ascontiguousarray(x), soft, out)
        softs.append(soft.copy()), bits.append(out), iters.append(int(it[0]))
    return softs, bits, iters


def main():
    if not O.ref_available():
        raise SystemExit("oracle/_ref/libsrsran_ref.so missing: run oracle/build_ref.sh (needs /root/reference)")
    rng = np.random.default_rng(20250101)
    cases = chain_cases(rng) + decode_only_cases(rng)
    meta = np.zeros(len(cases), CASE_DTYPE)
    llr_blob, soft0_blob = [], []
    out = {fl: dict(soft=[], bits=[], iters=np.full((len(cases), MAX_TX), -99, np.int32))
           for fl in (O.FLAVOUR_AVX512, O.FLAVOUR_GENERIC)}
    lo = so = 0
    soft_pos = bits_pos = 0
    for i, (cfg, llrs, soft0) in enumerate(cases):
        m = meta[i]
        for k in ("bg", "Z", "Qm", "E", "F", "Nref", "crc_id", "max_iter", "early_stop", "decode_only", "force_decoding"):
            m[k] = cfg[k]
        m["ntx"] = len(llrs)
        m["soft0_off"] = so
        soft0_blob.append(soft0)
        so += soft0.size
        K = O.kb_of(cfg["bg"]) * cfg["Z"]
        for tx, x in enumerate(llrs):
            m["rv"][tx] = cfg["rvs"][tx]
            m["llr_off"][tx] = lo
            llr_blob.append(x)
            lo += x.size
            m["soft_off"][tx] = soft_pos
            m["bits_off"][tx] = bits_pos
            soft_pos += 0 if cfg["decode_only"] else soft0.size
            bits_pos += (K + 7) // 8
        for fl in out:
            softs, bits, iters = run_reference(fl, cfg, llrs, soft0)
            out[fl]["soft"] += softs
            out[fl]["bits"] += bits
            out[fl]["iters"][i, :len(iters)] = iters
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(
        OUT, cases=meta, llrs=np.concatenate(llr_blob), soft0=np.concatenate(soft0_blob),
        soft_avx512=np.concatenate(out[O.FLAVOUR_AVX512]["soft"]), bits_avx512=np.concatenate(out[O.FLAVOUR_AVX512]["bits"]),
        iters_avx512=out[O.FLAVOUR_AVX512]["iters"],
        soft_generic=np.concatenate(out[O.FLAVOUR_GENERIC]["soft"]), bits_generic=np.concatenate(out[O.FLAVOUR_GENERIC]["bits"]),
        iters_generic=out[O.FLAVOUR_GENERIC]["iters"])
    n_pass = int((out[O.FLAVOUR_AVX512]["iters"] > 0).sum())
    print(f"wrote {OUT}: {len(cases)} cases, {lo} LLRs, {os.path.getsize(OUT)} bytes; avx512 passing transmissions {n_pass}")


if __name__ == "__main__":
    main()
