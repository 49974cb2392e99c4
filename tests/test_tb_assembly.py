"""Transport-block assembly on the device: pdc_submit_tbs / pdc_fetch_tbs / pdc_decode_device_tbs against the
oracle's restatement of pusch_decoder_impl::join_and_notify + concatenate_codeblocks (pusch_decoder_impl.cpp:396-505) and
against the transmitted transport block (round trip through the TS 38.212 transmit chain of tools/stimulus).

CPU part: the oracle restatement itself (round trip, bit-granular code-block sizes, the untouched-buffer and checksum-mismatch
cases). GPU part: the CUDA path, bit-exact against the oracle applied to the same code-block bits.
"""
import numpy as np
import pytest

from oracle.pyoracle import CRC16, CRC24A, CRC24B

# (bg, tbs_bits): C = 1 with CRC16, C = 1 with CRC24A, C = 3 on BG2, C = 3 on BG1, C = 16 with an ODD
# number of data bits per code block (BG2: 16 * 3601 = 57592 + 24), C = 5 on BG1, C = 3 with one bit of zero padding after
# the TB CRC (20024 = 3 * 6675 - 1: not a TS 38.214 size, but the reference segments it: ldpc_segmenter_rx_impl.cpp:52-106).
TB_CASES = [(2, 1032), (1, 5120), (2, 7656), (1, 20520), (2, 57592), (1, 40976), (1, 20000)]


def tx_side(S, bg, tbs, seed, corrupt_tb_crc=False):
    """-> dict(seg, tb_bits, msgs [C, K] (payload | CB CRC | fillers = 0))"""
    rng = np.random.default_rng(seed)
    tb = rng.integers(0, 2, tbs, dtype=np.uint8)
    seg = S.segmentation(bg, tbs)
    C = seg["C"]
    G = C * 2 * ((seg["K"] * 3 // 2) // 2)  # about rate 2/3, QPSK, one layer, equal E for every code block
    _, _, msgs = S.tb_encode(bg, tbs, tb, 0, 2, 0, 1, G)
    if corrupt_tb_crc:
        # One payload bit of the first code block flipped, its CB CRC recomputed: every CB CRC passes, the TB CRC does not.
        assert C > 1
        D = seg["K"] - seg["F"] - 24
        msgs[0, 5] ^= 1
        c = S.crc(3, msgs[0, :D])
        msgs[0, D:D + 24] = [(c >> (23 - i)) & 1 for i in range(24)]
    return dict(seg=seg, tb=tb, msgs=msgs, G=G)


def pack_rows(msgs):
    K = msgs.shape[1]
    stride = (K + 31) // 32 * 4
    out = np.zeros((msgs.shape[0], stride), np.uint8)
    out[:, :(K + 7) // 8] = np.packbits(msgs, axis=1)
    return out


# ---- CPU: the oracle restatement -------------------------------------------------------------------------------------------
@pytest.mark.parametrize("bg,tbs", TB_CASES)
def test_oracle_tb_round_trip(oracle, bg, tbs):
    from tools import stimulus as S

    t = tx_side(S, bg, tbs, seed=tbs)
    seg = t["seg"]
    if seg["C"] > 1:
        D = seg["K"] - seg["F"] - 24
        assert (seg["C"] - 1) * D + 24 < tbs + 24 <= seg["C"] * D
    out = np.full(tbs // 8, 0xA5, np.uint8)
    ok, written = oracle.tb_assemble(seg["K"], seg["F"], pack_rows(t["msgs"]), np.ones(seg["C"], np.uint8), tbs, out)
    assert ok and written
    assert (out == np.packbits(t["tb"])).all()
    # one code block whose CRC failed: nothing is written, no TB CRC
    flags = np.ones(seg["C"], np.uint8)
    flags[-1] = 0
    out[:] = 0xA5
    ok, written = oracle.tb_assemble(seg["K"], seg["F"], pack_rows(t["msgs"]), flags, tbs, out)
    assert not ok and not written and (out == 0xA5).all()


def test_oracle_tb_checksum_mismatch(oracle):
    from tools import stimulus as S

    bg, tbs = 2, 57592
    t = tx_side(S, bg, tbs, seed=7, corrupt_tb_crc=True)
    seg = t["seg"]
    out = np.zeros(tbs // 8, np.uint8)
    ok, written = oracle.tb_assemble(seg["K"], seg["F"], pack_rows(t["msgs"]), np.ones(seg["C"], np.uint8), tbs, out)
    want = t["tb"].copy()
    want[5] ^= 1
    assert written and not ok and (out == np.packbits(want)).all()


# ---- GPU -------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def acc():
    import srsran_project_b200 as pdc

    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=256, max_batch_llrs=16 << 20, harq_slots=64, nof_tickets=2)
    yield ctx
    ctx.close()


def cb_llrs(S, seg, msgs, r, E, sigma, seed, rv=0):
    cw = S.encode(seg["bg"], seg["Z"], msgs[r])
    tx = S.rate_match(seg["bg"], seg["Z"], cw, rv, 2, 0, seg["F"], E)
    return S.awgn_llrs(tx, 8.0, sigma, seed)


def build_batch(S, B, tbs_list, sigma_of, use_slots, seed0=0, corrupt=()):
    """One submission holding every TB of tbs_list. sigma_of(tb index, cb index) -> noise. Returns what the checks need."""
    txs = [tx_side(S, bg, tbs, seed=seed0 + 31 * k, corrupt_tb_crc=(k in corrupt)) for k, (bg, tbs) in enumerate(tbs_list)]
    n = sum(t["seg"]["C"] for t in txs)
    descs, tbd = B.make_descs(n), np.zeros(len(txs), B.TB_DESC_DTYPE)
    chunks, i, lo, oo, to = [], 0, 0, 0, 0
    for k, t in enumerate(txs):
        seg = t["seg"]
        C, E = seg["C"], t["G"] // seg["C"]
        tbd[k]["first_cb"], tbd[k]["nof_cbs"], tbd[k]["tbs_bits"], tbd[k]["tb_offset"] = i, C, tbs_list[k][1], to
        to += (tbs_list[k][1] // 8 + 3) // 4 * 4
        crc_id = CRC24B if C > 1 else (CRC24A if tbs_list[k][1] > 3824 else CRC16)
        for r in range(C):
            d = descs[i]
            d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = seg["bg"], seg["Z"], 2, 0, crc_id, 6
            d["flags"] = B.PDC_F_NEW_DATA | B.PDC_F_EARLY_STOP
            d["E"], d["nof_filler"], d["harq_slot"] = E, seg["F"], (i if use_slots else -1)
            d["llr_offset"], d["out_offset"], d["soft_offset"] = lo, oo, i * 25344
            chunks.append(cb_llrs(S, seg, t["msgs"], r, E, sigma_of(k, r), seed=1000 * k + r + seed0))
            lo += (E + 15) // 16 * 16
            oo += (seg["K"] + 31) // 32 * 4
            i += 1
    llrs = np.zeros(lo, np.int8)
    for d, c in zip(descs, chunks):
        llrs[int(d["llr_offset"]):int(d["llr_offset"]) + c.size] = c
    return dict(txs=txs, descs=descs, tbd=tbd, llrs=llrs, bits_bytes=oo, tb_bytes=to, n=n)


def check_against_oracle(oracle, b, cb_res, cb_bits, tb_res, tb_bytes, fill):
    """cb_bits: per-CB packed message bits (from the classic path) the oracle assembles; compares TB results and payloads."""
    for k, t in enumerate(b["txs"]):
        seg, tb = t["seg"], b["tbd"][k]
        i0, C, tbs = int(tb["first_cb"]), int(tb["nof_cbs"]), int(tb["tbs_bits"])
        stride = (seg["K"] + 31) // 32 * 4
        rows = np.stack([cb_bits[int(b["descs"][i0 + r]["out_offset"]):][:stride] for r in range(C)])
        want = np.full(tbs // 8, fill, np.uint8)
        ok, written = oracle.tb_assemble(seg["K"], seg["F"], rows, cb_res["crc_pass"][i0:i0 + C], tbs, want)
        got = tb_bytes[int(tb["tb_offset"]):int(tb["tb_offset"]) + tbs // 8]
        assert bool(tb_res[k]["tb_crc_ok"]) == ok and bool(tb_res[k]["written"]) == written, (k, tb_res[k], ok, written)
        assert int(tb_res[k]["nof_cbs_ok"]) == int(cb_res["crc_pass"][i0:i0 + C].sum())
        assert (got == want).all(), ("payload", k, int((got != want).sum()))
    # the bytes between the payloads (4-byte alignment padding) are not touched
    used = np.zeros(tb_bytes.size, bool)
    for tb in b["tbd"]:
        used[int(tb["tb_offset"]):int(tb["tb_offset"]) + int(tb["tbs_bits"]) // 8] = True
    assert (tb_bytes[~used] == fill).all()


@pytest.mark.gpu
@pytest.mark.parametrize("use_slots", [False, True])
def test_cuda_tb_assembly_matches_oracle(acc, oracle, use_slots):
    """Seven TBs (C = 1 .. 16, both base graphs, byte- and bit-granular code-block sizes) in one submission; one code block of
    the C = 3 TB is hopeless, the C = 5 TB carries a wrong TB CRC."""
    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    sigma = lambda k, r: 40.0 if (k == 3 and r == 1) else 2.0
    b = build_batch(S, B, TB_CASES, sigma, use_slots, seed0=11, corrupt={5})
    # classic path first: per-CB bits and results (the oracle assembles from these)
    res, bits = acc.run(b["descs"], b["llrs"], b["bits_bytes"], None)
    assert res["crc_pass"].sum() == b["n"] - 1 and res["crc_pass"][int(b["tbd"][3]["first_cb"]) + 1] == 0
    tb_bytes = np.full(b["tb_bytes"], 0x5A, np.uint8)
    t = acc.submit_tbs(b["descs"], b["tbd"], b["llrs"])
    tb_res, cb_res = acc.fetch_tbs(t, b["n"], len(TB_CASES), tb_bytes)
    assert (cb_res == res).all()
    check_against_oracle(oracle, b, res, bits, tb_res, tb_bytes, 0x5A)
    # and against what was transmitted
    for k, tx in enumerate(b["txs"]):
        got = tb_bytes[int(b["tbd"][k]["tb_offset"]):][:TB_CASES[k][1] // 8]
        if k == 3:
            assert not tb_res[k]["written"] and not tb_res[k]["tb_crc_ok"] and (got == 0x5A).all()
        elif k == 5:
            want = tx["tb"].copy()
            want[5] ^= 1
            assert tb_res[k]["written"] and not tb_res[k]["tb_crc_ok"] and (got == np.packbits(want)).all()
        else:
            assert tb_res[k]["tb_crc_ok"] and (got == np.packbits(tx["tb"])).all()


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["decoded", "dematch_only"])
def test_cuda_tb_assembly_across_harq_transmissions(acc, oracle, mode):
    """First transmission: one code block of a 16-CB TB fails, nothing is delivered. Second transmission: only that code block
    is sent again (the others are flagged PDC_F_DECODED as pusch_decoder_hw_impl skips them, or PDC_F_DEMATCH_ONLY as
    pusch_decoder_impl combines them without decoding: :335-350); the TB is assembled from bits decoded in both."""
    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    case = [(2, 57592)]
    bad = 9
    b = build_batch(S, B, case, lambda k, r: 8.0 if r == bad else 2.0, True, seed0=5)  # fails alone, decodes once combined
    tb_bytes = np.full(b["tb_bytes"], 0x11, np.uint8)
    tb_res, cb_res = acc.fetch_tbs(acc.submit_tbs(b["descs"], b["tbd"], b["llrs"]), b["n"], 1, tb_bytes)
    assert not tb_res[0]["written"] and tb_res[0]["nof_cbs_ok"] == b["n"] - 1 and cb_res["crc_pass"][bad] == 0
    assert (tb_bytes == 0x11).all()
    # retransmission (rv0 again, clean): combine into the same HARQ slot
    seg, t = b["txs"][0]["seg"], b["txs"][0]
    d2 = b["descs"].copy()
    E = int(d2[bad]["E"])
    if mode == "decoded":
        d2["flags"] = B.PDC_F_DECODED
        d2["E"], d2["llr_offset"] = 0, 0
        d2[bad]["flags"], d2[bad]["E"] = B.PDC_F_EARLY_STOP, E
        llrs2 = cb_llrs(S, seg, t["msgs"], bad, E, 1.0, seed=99)
    else:
        d2["flags"] = B.PDC_F_DEMATCH_ONLY
        d2[bad]["flags"] = B.PDC_F_EARLY_STOP
        llrs2 = np.zeros(b["llrs"].size, np.int8)
        for r in range(b["n"]):
            o = int(d2[r]["llr_offset"])
            llrs2[o:o + E] = cb_llrs(S, seg, t["msgs"], r, E, 1.0, seed=200 + r)
    tb_res, cb_res = acc.fetch_tbs(acc.submit_tbs(d2, b["tbd"], llrs2), b["n"], 1, tb_bytes)
    assert tb_res[0]["written"] and tb_res[0]["tb_crc_ok"] and tb_res[0]["nof_cbs_ok"] == b["n"]
    assert cb_res["crc_pass"][bad] == 1
    assert (tb_bytes[:case[0][1] // 8] == np.packbits(t["tb"])).all()


@pytest.mark.gpu
def test_cuda_tb_assembly_device_pointers(acc, oracle):
    """pdc_decode_device_tbs: everything stays in HBM (LLRs in, TB payloads + results out)."""
    import torch

    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    b = build_batch(S, B, TB_CASES[:5], lambda k, r: 2.0, False, seed0=3)
    res, bits = acc.run(b["descs"], b["llrs"], b["bits_bytes"], None)
    dev = "cuda:0"
    d_llrs = torch.from_numpy(b["llrs"]).to(dev)
    d_soft = torch.zeros(b["n"] * 25344, dtype=torch.int8, device=dev)
    d_bits = torch.zeros(b["bits_bytes"], dtype=torch.uint8, device=dev)
    d_res = torch.zeros(b["n"] * 8, dtype=torch.uint8, device=dev)
    d_tb = torch.full((b["tb_bytes"],), 0x77, dtype=torch.uint8, device=dev)
    d_tbres = torch.zeros(len(b["tbd"]) * 4, dtype=torch.uint8, device=dev)
    acc.decode_device_tbs(b["descs"], b["tbd"], d_llrs.data_ptr(), d_soft.data_ptr(), d_bits.data_ptr(), d_res.data_ptr(),
                          d_tb.data_ptr(), d_tbres.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    tb_res = d_tbres.cpu().numpy().view(B.TB_RESULT_DTYPE)
    assert (d_res.cpu().numpy().view(B.RESULT_DTYPE) == res).all()
    check_against_oracle(oracle, b, res, bits, tb_res, d_tb.cpu().numpy(), 0x77)
    assert tb_res["tb_crc_ok"].all()


@pytest.mark.gpu
def test_cuda_tb_descriptor_contract(acc):
    """What plan_tbs rejects (PDC_ERR_INVALID_ARG, nothing launched)."""
    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    b = build_batch(S, B, [(2, 7656)], lambda k, r: 2.0, False)
    n = b["n"]
    for field, value in (("tbs_bits", 7664), ("nof_cbs", n + 1), ("nof_cbs", n - 1), ("first_cb", n), ("tb_offset", 2),
                         ("tbs_bits", 7657)):
        tbd = b["tbd"].copy()
        tbd[0][field] = value
        with pytest.raises(B.PdcError) as e:
            acc.submit_tbs(b["descs"], tbd, b["llrs"])
        assert e.value.status == B.PDC_ERR_INVALID_ARG, field
    d = b["descs"].copy()
    d[1]["flags"] = B.PDC_F_DECODED  # needs a HARQ slot
    with pytest.raises(B.PdcError):
        acc.submit_tbs(d, b["tbd"], b["llrs"])
    # a TB ticket is not a CB ticket
    t = acc.submit_tbs(b["descs"], b["tbd"], b["llrs"])
    with pytest.raises(B.PdcError) as e:
        acc.fetch(t, b["n"], np.zeros(b["bits_bytes"], np.uint8))
    assert e.value.status == B.PDC_ERR_BAD_TICKET
    tb_res, _ = acc.fetch_tbs(t, b["n"], 1, np.zeros(b["tb_bytes"], np.uint8))
    assert tb_res[0]["tb_crc_ok"]
