"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs. Bit-exact."""
import numpy as np
import pytest

from oracle.pyoracle import (CRC16, CRC24A, CRC24B, CRC_NONE, FLAVOUR_AVX512, FLAVOUR_GENERIC, LIFTING_SIZES, kb_of,
                             nshort_of)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def acc():
    import srsran_project_b200 as pdc

    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=2048, max_batch_llrs=96 << 20, harq_slots=64, nof_tickets=2)
    yield ctx
    ctx.close()


def awgn_llrs(rng, bits, sigma, amp=8.0):
    x = amp * (1.0 - 2.0 * bits.astype(np.float64)) + rng.normal(0.0, sigma, bits.size)
    return np.clip(np.round(x), -120, 120).astype(np.int8)


def message_with_crc(oracle, rng, K, F, crc_id):
    ncrc = 16 if crc_id == CRC16 else 24
    payload = rng.integers(0, 2, K - F - ncrc, dtype=np.uint8)
    c = oracle.crc(crc_id, np.packbits(payload), payload.size)
    crc_bits = np.array([(c >> (ncrc - 1 - i)) & 1 for i in range(ncrc)], np.uint8)
    return np.concatenate([payload, crc_bits, np.zeros(F, np.uint8)])


def out_stride(K):
    return ((K + 31) // 32) * 4


def run_decode_only_batch(acc, oracle, cases, flavour):
    """cases: list of dict(bg, Z, llrs, F, max_iter, crc_id, force). Compares every CB with the oracle."""
    from srsran_project_b200 import binding as B

    n = len(cases)
    descs = B.make_descs(n)
    llr_off, out_off = 0, 0
    chunks = []
    for i, c in enumerate(cases):
        K = kb_of(c["bg"]) * c["Z"]
        d = descs[i]
        d["bg"], d["Z"], d["Qm"], d["rv"] = c["bg"], c["Z"], 1, 0
        d["crc_type"], d["max_iter"] = c["crc_id"], c["max_iter"]
        flags = B.PDC_F_DECODE_ONLY | (B.PDC_F_EARLY_STOP if c["crc_id"] != CRC_NONE else 0)
        flags |= B.PDC_F_GENERIC if flavour == FLAVOUR_GENERIC else 0
        flags |= B.PDC_F_FORCE_DECODING if c.get("force") else 0
        d["flags"], d["E"], d["nof_filler"] = flags, c["llrs"].size, c["F"]
        d["llr_offset"], d["out_offset"] = llr_off, out_off
        chunks.append(c["llrs"])
        llr_off += (c["llrs"].size + 15) // 16 * 16
        out_off += out_stride(K)
    llrs = np.zeros(llr_off, np.int8)
    for i, c in enumerate(cases):
        o = int(descs[i]["llr_offset"])
        llrs[o:o + c["llrs"].size] = c["llrs"]
    res, bits = acc.run(descs, llrs, out_off)
    for i, c in enumerate(cases):
        K = kb_of(c["bg"]) * c["Z"]
        it, want = oracle.decode(flavour, c["bg"], c["Z"], c["llrs"], c["F"], c["max_iter"], c["crc_id"],
                                 bool(c.get("force")))
        tag = (i, c["bg"], c["Z"], c["crc_id"], c["max_iter"])
        assert res[i]["iterations"] == it, tag
        assert res[i]["crc_pass"] == (1 if it > 0 else 0), tag
        if res[i]["status"] == 0:
            o = int(descs[i]["out_offset"])
            got = np.unpackbits(bits[o:o + (K + 7) // 8])[:K]
            assert (got == np.unpackbits(want)[:K]).all(), tag
    return res


@pytest.mark.parametrize("flavour", [FLAVOUR_AVX512, FLAVOUR_GENERIC])
@pytest.mark.parametrize("bg", [1, 2])
def test_decoder_all_lifting_sizes(acc, oracle, flavour, bg):
    rng = np.random.default_rng(100 + bg + flavour)
    kb, ns = kb_of(bg), nshort_of(bg)
    cases = []
    for Z in LIFTING_SIZES:
        K = kb * Z
        for trial in range(4):
            msg = rng.integers(0, 2, K, dtype=np.uint8)
            cb = oracle.encode(bg, Z, msg)
            n = ns * Z if trial == 0 else int(rng.integers(kb + 2, ns + 1)) * Z
            llrs = awgn_llrs(rng, cb[:n], (2.0, 5.0, 7.0, 9.0)[trial])
            if trial == 3:
                llrs[rng.integers(0, n, n // 3)] = 0
            if trial == 2:
                llrs[int(rng.integers(K, n)):] = 0
            for crc_id, max_iter in ((CRC_NONE, 6), (CRC16, 6), (CRC24B, 2), (CRC_NONE, 1)):
                cases.append(dict(bg=bg, Z=Z, llrs=llrs, F=0, max_iter=max_iter, crc_id=crc_id))
    run_decode_only_batch(acc, oracle, cases, flavour)


def test_decoder_early_stop_iteration_counts(acc, oracle):
    rng = np.random.default_rng(7)
    cases, seen = [], set()
    for bg, Z, crc_id in ((1, 384, CRC24B), (1, 96, CRC24A), (2, 56, CRC16), (2, 384, CRC24B), (1, 2, CRC16),
                          (2, 15, CRC16), (1, 208, CRC24B)):
        K = kb_of(bg) * Z
        F = min(16, Z)
        for sigma in (1.0, 4.0, 6.0, 7.0, 8.0, 10.0):
            msg = message_with_crc(oracle, rng, K, F, crc_id)
            cb = oracle.encode(bg, Z, msg)
            llrs = awgn_llrs(rng, cb, sigma)
            llrs[K - 2 * Z - F:K - 2 * Z] = 127
            cases.append(dict(bg=bg, Z=Z, llrs=llrs, F=F, max_iter=6, crc_id=crc_id))
    res = run_decode_only_batch(acc, oracle, cases, FLAVOUR_AVX512)
    seen = set(int(x) for x in res["iterations"])
    assert len(seen) >= 4, seen


def test_decoder_reference_known_answers(acc, oracle):
    # ldpc_enc_dec_test.cpp:340-364 — all-zero and almost-zero LLRs give nullopt and an all-ones message.
    cases = []
    rng = np.random.default_rng(3)
    for bg, Z in ((1, 8), (2, 384), (1, 384), (2, 3)):
        n = nshort_of(bg) * Z
        K = kb_of(bg) * Z
        sparse = np.zeros(n, np.int8)
        sparse[rng.integers(K + 4 * Z, n, 5)] = 1
        for llrs in (np.zeros(n, np.int8), sparse):
            for force in (True, False):
                for crc_id in (CRC_NONE, CRC16):
                    cases.append(dict(bg=bg, Z=Z, llrs=llrs, F=0, max_iter=2, crc_id=crc_id, force=force))
    run_decode_only_batch(acc, oracle, cases, FLAVOUR_AVX512)


def chain_case(rng, oracle, bg, Z, Qm, E, F, Nref, crc_id, sigmas, rvs, garbage=True):
    K, N = kb_of(bg) * Z, nshort_of(bg) * Z
    msg = message_with_crc(oracle, rng, K, F, crc_id)
    cb = oracle.encode(bg, Z, msg)
    txs = []
    for sigma, rv in zip(sigmas, rvs):
        tx_bits = oracle.rate_match(bg, Z, cb, rv, Qm, Nref, F, E)
        txs.append((rv, awgn_llrs(rng, tx_bits, sigma)))
    soft0 = rng.integers(-120, 121, N).astype(np.int8) if garbage else np.zeros(N, np.int8)
    return dict(bg=bg, Z=Z, Qm=Qm, E=E, F=F, Nref=Nref, crc_id=crc_id, txs=txs, soft0=soft0, msg=msg)


def run_chain(acc, oracle, cases, flavour, early_stop=True, use_arena=False, max_iter=6, arena_init="write"):
    """Runs every case's transmissions in sequence (one batch per transmission index) and compares with the oracle:
    soft buffer after every transmission, bits, iterations, CRC. arena_init (use_arena): "write" = the case's initial buffer is
    written into its slot, "free" = the slot is zero-filled first, "release" = pdc_harq_release, "keep" = the slot keeps what earlier
    transmissions left in it (the oracle starts from a read-back of it)."""
    from srsran_project_b200 import binding as B

    n = len(cases)
    ntx = max(len(c["txs"]) for c in cases)
    soft_off, out_off = [], []
    so = oo = 0
    for c in cases:
        soft_off.append(so), out_off.append(oo)
        so += (nshort_of(c["bg"]) * c["Z"] + 15) // 16 * 16
        oo += out_stride(kb_of(c["bg"]) * c["Z"])
    soft_gpu = np.zeros(so, np.int8)
    soft_cpu = [c["soft0"].copy() for c in cases]
    for i, c in enumerate(cases):
        soft_gpu[soft_off[i]:soft_off[i] + c["soft0"].size] = c["soft0"]
        if use_arena and arena_init == "write":
            acc.harq_write(i, c["soft0"])
        elif use_arena and arena_init == "free":
            acc.harq_free(i)
            soft_cpu[i][:] = 0
        elif use_arena and arena_init == "release":
            acc.harq_release(i)  # reads as zeros from now on, cleared lazily
            soft_cpu[i][:] = 0
        elif use_arena:
            soft_cpu[i][:] = acc.harq_read(i, c["soft0"].size)
    for tx in range(ntx):
        idx = [i for i, c in enumerate(cases) if tx < len(c["txs"])]
        descs = B.make_descs(len(idx))
        lo = 0
        for k, i in enumerate(idx):
            c = cases[i]
            rv, llrs = c["txs"][tx]
            d = descs[k]
            d["bg"], d["Z"], d["Qm"], d["rv"] = c["bg"], c["Z"], c["Qm"], rv
            d["crc_type"], d["max_iter"] = c["crc_id"], max_iter
            flags = (B.PDC_F_NEW_DATA if tx == 0 else 0) | (B.PDC_F_EARLY_STOP if early_stop else 0)
            flags |= B.PDC_F_GENERIC if flavour == FLAVOUR_GENERIC else 0
            d["flags"], d["E"], d["Nref"], d["nof_filler"] = flags, c["E"], c["Nref"], c["F"]
            d["harq_slot"] = i if use_arena else -1
            d["llr_offset"], d["soft_offset"], d["out_offset"] = lo, soft_off[i], out_off[i]
            lo += (c["E"] + 15) // 16 * 16
        llrs_all = np.zeros(lo, np.int8)
        for k, i in enumerate(idx):
            o = int(descs[k]["llr_offset"])
            llrs_all[o:o + cases[i]["E"]] = cases[i]["txs"][tx][1]
        res, bits = acc.run(descs, llrs_all, oo, None if use_arena else soft_gpu)
        for k, i in enumerate(idx):
            c = cases[i]
            K, N = kb_of(c["bg"]) * c["Z"], nshort_of(c["bg"]) * c["Z"]
            rv, llrs = c["txs"][tx]
            it, want = oracle.codeblock_decode(flavour, c["bg"], c["Z"], soft_cpu[i], llrs, tx == 0, rv, c["Qm"],
                                               c["Nref"], c["F"], c["crc_id"], max_iter, early_stop)
            tag = (i, tx, c["bg"], c["Z"], c["Qm"], c["E"], c["F"], c["Nref"])
            got_soft = acc.harq_read(i, N) if use_arena else soft_gpu[soft_off[i]:soft_off[i] + N]
            assert (got_soft == soft_cpu[i]).all(), ("soft", tag, int((got_soft != soft_cpu[i]).sum()))
            assert res[k]["iterations"] == it, ("iterations", tag, int(res[k]["iterations"]), it)
            got = np.unpackbits(bits[out_off[i]:out_off[i] + (K + 7) // 8])[:K]
            assert (got == np.unpackbits(want)[:K]).all(), ("bits", tag)
    return


@pytest.mark.parametrize("flavour", [FLAVOUR_AVX512, FLAVOUR_GENERIC])
def test_chain_all_lifting_sizes_mixed_batch(acc, oracle, flavour):
    # BASELINE config 3 in spirit: both base graphs, every Z, mixed Qm / rate / rv / fillers / LBRM in one batch.
    rng = np.random.default_rng(11 + flavour)
    cases = []
    for bg in (1, 2):
        kb, ns = kb_of(bg), nshort_of(bg)
        for zi, Z in enumerate(LIFTING_SIZES):
            K, N = kb * Z, ns * Z
            Qm = (2, 4, 6, 8, 1)[zi % 5]
            F = int(rng.integers(0, Z)) if zi % 2 else 0
            crc_id = CRC16 if K - F < 400 else (CRC24B if zi % 3 else CRC24A)
            if K - F <= 24:
                crc_id = CRC16
            if K - F <= 16:
                continue
            Nref = 0 if zi % 4 else int(rng.integers((kb + 5) * Z, N))
            rate = (0.9, 0.6, 0.35, 0.2)[zi % 4]
            E = max(Qm, int(K / rate) // Qm * Qm)
            rvs = [(0, 2, 3, 1), (0, 3), (0, 0, 2), (0,)][zi % 4]
            sig = [9.0 if bg == 1 else 7.0] * len(rvs)
            cases.append(chain_case(rng, oracle, bg, Z, Qm, E, F, Nref, crc_id, sig, rvs))
    order = rng.permutation(len(cases))
    run_chain(acc, oracle, [cases[i] for i in order], flavour)


def test_chain_harq_arena_rv0_rv2(acc, oracle):
    # BASELINE config 4: RV0 + RV2 over a persistent device-resident soft buffer, low SNR (max iterations, CRC fails on
    # the first transmission), garbage initial buffer content to pin the dematcher's write set.
    rng = np.random.default_rng(21)
    cases = []
    for k in range(8):
        E = (8960, 8992)[k % 2]
        cases.append(chain_case(rng, oracle, 1, 384, 8, E, 16, (0, 12611)[k % 2], CRC24B, (3.2, 3.0), (0, 2)))
    run_chain(acc, oracle, cases, FLAVOUR_AVX512, use_arena=True)


def test_chain_limited_buffer_first_transmissions_on_arena_slots(acc, oracle):
    """BASELINE config 2' (what the live stack asks for: Nref = 12611 < N, pusch_processor_impl.cpp:327-334). The reference leaves
    [end, N - (Ncb - end)) of the soft buffer as it was (allot_llrs :198-200), so the decoder may see stale soft bits there. The
    runtime keeps a bound on the non-zero extent of every arena slot: released (zeroed) slots take the four-row kernel, slots
    with stale content beyond the new transmission's end take the general one. Both must reproduce the oracle, which starts
    from the slot's real content: zeros, then whatever the previous transport block left."""
    rng = np.random.default_rng(77)

    def cases_for(Es):
        out = []
        for k, E in enumerate(Es):
            out.append(chain_case(rng, oracle, 1, 384, 8, E, 16, 12611, CRC24B, (2.2,), (0,), garbage=False))
        out.append(chain_case(rng, oracle, 2, 208, 4, 4 * 560, 8, 5700, CRC24A, (2.0,), (0,), garbage=False))
        out.append(chain_case(rng, oracle, 1, 96, 6, 6 * 380, 0, 3100, CRC24B, (2.2,), (0,), garbage=False))
        return out

    for flavour in (FLAVOUR_AVX512, FLAVOUR_GENERIC):
        # released slots: every untouched stretch holds zeros
        run_chain(acc, oracle, cases_for([8960, 8992, 8960, 8992, 8992, 8960]), flavour, use_arena=True, arena_init="free")
        # the same slots again, new data: where the new transmission is SHORTER than the previous one, [end_new, end_old) is stale
        run_chain(acc, oracle, cases_for([8992, 8960, 8960, 8992, 8960, 8992]), flavour, use_arena=True, arena_init="keep")
        # and once more: the stale stretches persist (or are covered by a longer transmission)
        run_chain(acc, oracle, cases_for([8960, 8960, 8992, 8960, 8992, 8992]), flavour, use_arena=True, arena_init="keep")
    # a retransmission raises the bound to the whole buffer; new data afterwards must still be right
    cases = [chain_case(rng, oracle, 1, 384, 8, 8960, 16, 12611, CRC24B, (3.4, 3.2), (0, 2), garbage=False) for _ in range(3)]
    run_chain(acc, oracle, cases, FLAVOUR_AVX512, use_arena=True, arena_init="free")
    run_chain(acc, oracle, cases_for([8960, 8992, 8960])[:3], FLAVOUR_AVX512, use_arena=True, arena_init="keep")
    # pdc_harq_release (what free_harq_context_entry maps to): the slots read as zeros without having been cleared; limited-buffer
    # first transmissions rewrite them, ...
    run_chain(acc, oracle, cases, FLAVOUR_AVX512, use_arena=True, arena_init="free")  # leaves rv0 + rv2 content everywhere
    run_chain(acc, oracle, cases_for([8960, 8992, 8960, 8992, 8992, 8960]), FLAVOUR_AVX512, use_arena=True, arena_init="release")
    # ... a low-rate limited-buffer first transmission (does not fit the staging tile, leaves a stretch untouched) and a
    # combining transmission into a released slot have the slot cleared on the stream first
    run_chain(acc, oracle, cases, FLAVOUR_AVX512, use_arena=True, arena_init="free")
    low = [chain_case(rng, oracle, 1, 384, 4, 15000, 16, 20000, CRC24B, (3.0,), (0,), garbage=False) for _ in range(2)]
    low.append(chain_case(rng, oracle, 1, 384, 8, 8960, 16, 12611, CRC24B, (3.0,), (2,), garbage=False))  # starts at rv 2
    run_chain(acc, oracle, low, FLAVOUR_AVX512, use_arena=True, arena_init="release")
    assert (acc.harq_read(5, 25344) != 0).any()  # slot 5 still holds the first batch's second case ...
    acc.harq_release(5)
    assert not acc.harq_read(5, 25344).any()     # ... and reads as zeros once released


def test_chain_limited_buffer_random_sizes_on_arena_slots(acc, oracle):
    """The same bookkeeping over random lifting sizes, base graphs, modulation orders, filler counts and buffer limits: three
    rounds of new data on the same 60 arena slots (zero-filled, then re-used twice with other lengths, so that some slots hold
    stale soft bits beyond the new transmission and some do not), high code rates so that the four-row kernels are candidates."""
    rng = np.random.default_rng(2024)
    even = [Z for Z in LIFTING_SIZES if Z % 2 == 0 and Z >= 16]
    geo = []
    for k in range(60):
        bg = 1 + (k % 2)
        Z = int(even[int(rng.integers(0, len(even)))])
        kb, ns = kb_of(bg), nshort_of(bg)
        F = int(rng.integers(0, Z // 4)) * (2 if k % 3 else 1)
        Qm = (2, 4, 6, 8)[k % 4]
        geo.append((bg, Z, kb * Z, ns * Z, F, Qm))
    for rnd, init in enumerate(("free", "keep", "keep")):
        cases = []
        for k, (bg, Z, K, N, F, Qm) in enumerate(geo):
            kb = kb_of(bg)
            n_sys = (kb - 2) * Z
            # between "just covers the systematic part" and "fills the staging tile of the four-row kernel" ((Kb + 2) Z)
            lo, hi = n_sys - F + Qm, (kb + 2) * Z - F
            E = int(rng.integers(lo, hi)) // Qm * Qm
            E = max(E, (n_sys - F + Qm - 1) // Qm * Qm)
            Nref = int(rng.integers((kb + 3) * Z, N)) if (k + rnd) % 4 else 0
            crc_id = CRC16 if K - F < 400 else (CRC24B if k % 3 else CRC24A)
            cases.append(chain_case(rng, oracle, bg, Z, Qm, E, F, Nref, crc_id, (2.0,), (0,), garbage=False))
        run_chain(acc, oracle, cases, FLAVOUR_AVX512 if rnd != 1 else FLAVOUR_GENERIC, use_arena=True, arena_init=init)


def test_chain_without_early_stop(acc, oracle):
    rng = np.random.default_rng(31)
    cases = [chain_case(rng, oracle, 1, 384, 8, 8960, 16, 0, CRC24B, (s,), (0,)) for s in (1.0, 2.0, 3.5)]
    cases += [chain_case(rng, oracle, 2, 96, 4, 4000, 8, 0, CRC16, (s,), (0,)) for s in (1.0, 6.0)]
    run_chain(acc, oracle, cases, FLAVOUR_AVX512, early_stop=False, max_iter=3)


def test_chain_heavy_repetition_and_tiny_inputs(acc, oracle):
    # E >> Ncb (many wraps of the circular buffer) and E smaller than the systematic part.
    rng = np.random.default_rng(41)
    cases = []
    for bg, Z, Qm, E in ((2, 2, 2, 6000), (1, 3, 1, 4001), (2, 8, 6, 30000), (1, 16, 8, 64), (2, 24, 2, 50),
                         (1, 384, 2, 60000), (2, 384, 4, 2048)):
        cases.append(chain_case(rng, oracle, bg, Z, Qm, E, min(4, Z - 1), 0, CRC16, (6.0, 6.0), (0, 1)))
    run_chain(acc, oracle, cases, FLAVOUR_AVX512)


# ---- packed half2 kernel (pdc_kernel_h2.cuh): high-rate code blocks, four core rows, Z a multiple of 32 -----------------
H2_SIZES = [Z for Z in LIFTING_SIZES if Z % 32 == 0]


def high_rate_E(bg, Z, F, Qm, frac):
    """Largest E (multiple of Qm) that keeps nof_layers == 4: nz = E + F <= (Kb + 2) Z, scaled by frac."""
    kb = kb_of(bg)
    lo = (kb - 2) * Z - F + 8  # at least a few parity bits
    hi = (kb + 2) * Z - F
    return max(Qm, int(lo + frac * (hi - lo)) // Qm * Qm)


@pytest.mark.parametrize("flavour", [FLAVOUR_AVX512, FLAVOUR_GENERIC])
def test_h2_chain_high_rate_all_sizes(acc, oracle, flavour):
    rng = np.random.default_rng(71 + flavour)
    cases = []
    for bg in (1, 2):
        for zi, Z in enumerate(H2_SIZES):
            K = kb_of(bg) * Z
            for trial in range(6):
                Qm = (8, 6, 4, 2, 1, 8)[trial]
                F = (0, 16, Z - 1, 5, 0, 2 * Z - 3)[trial]
                crc_id = (CRC24B, CRC24A, CRC16, CRC24B, CRC16, CRC24B)[(trial + zi) % 6]
                E = high_rate_E(bg, Z, F, Qm, (1.0, 0.9, 0.5, 1.0, 0.2, 0.97)[trial])
                # Noise from "decodes at once" to "never decodes"; amplitude up to the LLR range to exercise promotion.
                sigma = (1.0, 2.0, 2.6, 3.5, 8.0, 2.3)[(trial + zi) % 6]
                c = chain_case(rng, oracle, bg, Z, Qm, E, F, 0, crc_id, (sigma, sigma), (0, 0), garbage=(trial % 2 == 0))
                cases.append(c)
    order = rng.permutation(len(cases))
    run_chain(acc, oracle, [cases[i] for i in order], flavour)


def test_h2_chain_strong_llrs_and_no_early_stop(acc, oracle):
    # Large amplitudes drive soft bits into the +-127 "infinity" states (promotion, sticky infinities).
    rng = np.random.default_rng(83)
    cases = []
    for bg, Z in ((1, 384), (1, 352), (2, 384), (2, 64), (1, 32), (2, 32), (1, 96)):
        for amp, sigma in ((40.0, 30.0), (100.0, 60.0), (8.0, 2.2), (20.0, 14.0)):
            K = kb_of(bg) * Z
            F = 16 if Z > 16 else 0
            E = high_rate_E(bg, Z, F, 8, 1.0)
            msg = message_with_crc(oracle, rng, K, F, CRC24B)
            cb = oracle.encode(bg, Z, msg)
            txs = []
            for rv in (0, 0):
                bits = oracle.rate_match(bg, Z, cb, rv, 8, 0, F, E)
                txs.append((rv, awgn_llrs(rng, bits, sigma, amp)))
            cases.append(dict(bg=bg, Z=Z, Qm=8, E=E, F=F, Nref=0, crc_id=CRC24B, txs=txs,
                              soft0=np.zeros(nshort_of(bg) * Z, np.int8), msg=msg))
    run_chain(acc, oracle, cases, FLAVOUR_AVX512)
    run_chain(acc, oracle, cases, FLAVOUR_AVX512, early_stop=False, max_iter=4)
    run_chain(acc, oracle, cases, FLAVOUR_GENERIC, early_stop=False, max_iter=2)


def test_h2_decode_only_high_rate(acc, oracle):
    rng = np.random.default_rng(97)
    cases = []
    for bg in (1, 2):
        kb = kb_of(bg)
        for Z in H2_SIZES:
            K = kb * Z
            for trial in range(3):
                n = (kb + 2) * Z
                msg = message_with_crc(oracle, rng, K, 0, CRC24B if trial else CRC16)
                cb = oracle.encode(bg, Z, msg)
                llrs = awgn_llrs(rng, cb[:n], (2.0, 3.0, 6.0)[trial])
                for crc_id, max_iter in ((CRC_NONE, 6), (CRC24B if trial else CRC16, 6), (CRC24A, 1)):
                    cases.append(dict(bg=bg, Z=Z, llrs=llrs, F=0, max_iter=max_iter, crc_id=crc_id))
    run_decode_only_batch(acc, oracle, cases, FLAVOUR_AVX512)


def test_h2_unaligned_buffers_and_byte_paths(acc, oracle):
    """The packed kernel's vector paths need aligned LLRs / soft buffers and Kq, F multiples of 4; everything else takes the
    byte paths. Same results either way: odd LLR and soft-buffer offsets, every Qm, fillers of every residue."""
    from srsran_project_b200 import binding as B

    rng = np.random.default_rng(131)
    cases = []
    for bg, Z in ((1, 384), (2, 384), (1, 64), (2, 96)):
        for k, (Qm, F) in enumerate(((8, 16), (8, 17), (6, 12), (6, 3), (4, 8), (4, 6), (2, 4), (2, 1), (1, 0), (8, 0))):
            E = high_rate_E(bg, Z, F, Qm, (1.0, 0.93, 0.8)[k % 3])
            if k == 9:
                E = E // (4 * Qm) * (4 * Qm) + Qm  # Kq not a multiple of 4
            cases.append(chain_case(rng, oracle, bg, Z, Qm, E, F, 0, CRC24B, (2.4,), (0,)))
    for shift_llr, shift_soft in ((0, 0), (1, 0), (8, 4), (3, 7), (16, 16)):
        n = len(cases)
        descs = B.make_descs(n)
        lo = so = oo = 0
        soft_off, out_off = [], []
        for i, c in enumerate(cases):
            d = descs[i]
            d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = c["bg"], c["Z"], c["Qm"], 0, c["crc_id"], 6
            d["flags"], d["E"], d["nof_filler"], d["harq_slot"] = B.PDC_F_NEW_DATA | B.PDC_F_EARLY_STOP, c["E"], c["F"], -1
            lo = (lo + 15) // 16 * 16 + shift_llr
            so = (so + 15) // 16 * 16 + shift_soft
            d["llr_offset"], d["soft_offset"], d["out_offset"] = lo, so, oo
            soft_off.append(so), out_off.append(oo)
            lo += c["E"]
            so += nshort_of(c["bg"]) * c["Z"]
            oo += out_stride(kb_of(c["bg"]) * c["Z"])
        llrs = np.zeros(lo + 16, np.int8)
        soft = rng.integers(-120, 121, so + 16).astype(np.int8)
        soft_before = soft.copy()
        for i, c in enumerate(cases):
            o = int(descs[i]["llr_offset"])
            llrs[o:o + c["E"]] = c["txs"][0][1]
        res, bits = acc.run(descs, llrs, oo, soft)
        for i, c in enumerate(cases):
            K, N = kb_of(c["bg"]) * c["Z"], nshort_of(c["bg"]) * c["Z"]
            want_soft = soft_before[soft_off[i]:soft_off[i] + N].copy()
            it, want = oracle.codeblock_decode(FLAVOUR_AVX512, c["bg"], c["Z"], want_soft, c["txs"][0][1], True, 0, c["Qm"],
                                               0, c["F"], c["crc_id"], 6, True)
            tag = (shift_llr, shift_soft, i, c["bg"], c["Z"], c["Qm"], c["E"], c["F"])
            assert (soft[soft_off[i]:soft_off[i] + N] == want_soft).all(), ("soft", tag)
            assert res[i]["iterations"] == it, ("iterations", tag)
            got = np.unpackbits(bits[out_off[i]:out_off[i] + (K + 7) // 8])[:K]
            assert (got == np.unpackbits(want)[:K]).all(), ("bits", tag)
        # nothing outside the code blocks' own soft buffers was written
        mask = np.ones(soft.size, bool)
        for i, c in enumerate(cases):
            mask[soft_off[i]:soft_off[i] + nshort_of(c["bg"]) * c["Z"]] = False
        assert (soft[mask] == soft_before[mask]).all()
