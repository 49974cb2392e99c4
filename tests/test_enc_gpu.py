"""PDSCH encoder on the device: pdc_pdsch_encode / pdc_pdsch_encode_device against the golden codewords recorded
from the compiled pdsch_encoder_impl, against the TX-chain restatement on a batch of transport blocks in one submission, and
as a loop: encode on the device -> BPSK-like LLRs -> the device decode chain gives the transport block back. Bit-exact."""
import hashlib
import os

import numpy as np
import pytest

from oracle import pyoracle
from tools.enc_cases import CASES

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pdsch_enc_v1.npz")


@pytest.fixture(scope="module")
def acc():
    import srsran_project_b200 as pdc

    a = pdc.PuschDecCuda(device=0, max_batch_cbs=512, max_batch_llrs=4 << 20, harq_slots=0, nof_tickets=2)
    yield a
    a.close()


def tb_bytes(case_index, tbs_bits):
    return np.random.default_rng(500 + case_index).integers(0, 256, tbs_bits // 8, dtype=np.uint8)


def make_batch(B, cases, first_index=0):
    tbs = np.zeros(len(cases), B.PDSCH_TB_DTYPE)
    to = co = 0
    chunks = []
    for i, (bg, nbits, Qm, L, nsym, rv, Nref) in enumerate(cases):
        d = tbs[i]
        d["bg"], d["Qm"], d["rv"], d["nof_layers"], d["tbs_bits"], d["nof_ch_symbols"], d["Nref"] = bg, Qm, rv, L, nbits, nsym, Nref
        d["tb_offset"], d["cw_offset"] = to, co
        chunks.append(tb_bytes(first_index + i, nbits))
        to += nbits // 8
        co += nsym * Qm
    return tbs, np.concatenate(chunks), co


def test_golden_codewords_one_by_one(acc):
    from srsran_project_b200 import binding as B

    g = np.load(GOLDEN)
    for k, case in enumerate(CASES):
        tbs, data, cw_len = make_batch(B, [case], k)
        cw = acc.pdsch_encode(tbs, data, cw_len)
        assert cw.max() <= 1
        got = np.frombuffer(hashlib.sha256(np.packbits(cw).tobytes()).digest(), np.uint8)
        assert np.array_equal(got, g["digests"][k]), case
        if f"cw{k}" in g.files:
            assert np.array_equal(cw, np.unpackbits(g[f"cw{k}"])[:cw.size]), case


def test_batch_of_transport_blocks_matches_restatement(acc):
    """Every case in ONE submission (code blocks of different base graphs, lifting sizes and lengths in one grid), plus every
    redundancy version of the mid-sized cases."""
    from srsran_project_b200 import binding as B

    orc = pyoracle.PdschEncOracle()
    cases = list(CASES[:-1])
    for (bg, nbits, Qm, L, nsym, rv, Nref) in CASES[5:12]:
        cases += [(bg, nbits, Qm, L, nsym, r, Nref) for r in range(4) if r != rv]
    tbs, data, cw_len = make_batch(B, cases)
    cw = acc.pdsch_encode(tbs, data, cw_len)
    for d, (bg, nbits, Qm, L, nsym, rv, Nref) in zip(tbs, cases):
        a, n = int(d["tb_offset"]), nbits // 8
        want = orc.encode(data[a:a + n], nbits, bg, rv, Qm, Nref, L, nsym)
        o = int(d["cw_offset"])
        got = cw[o:o + want.size]
        assert np.array_equal(got, want), ((bg, nbits, Qm, L, nsym, rv, Nref), np.nonzero(got != want)[0][:4])


def test_contract_violations(acc):
    from srsran_project_b200 import binding as B

    tbs, data, cw_len = make_batch(B, [(2, 96, 2, 1, 150, 0, 0)])
    for field, value in (("bg", 3), ("Qm", 3), ("rv", 4), ("nof_layers", 5), ("tbs_bits", 100), ("nof_ch_symbols", 0),
                         ("tbs_bits", 1277976)):
        bad = tbs.copy()
        bad[0][field] = value
        with pytest.raises(B.PdcError):
            acc.pdsch_encode(bad, np.zeros(200000, np.uint8), 400000)
    odd = tbs.copy()
    odd[0]["nof_layers"], odd[0]["nof_ch_symbols"] = 2, 151  # not a multiple of the number of layers
    with pytest.raises(B.PdcError):
        acc.pdsch_encode(odd, data, cw_len + 2)
    with pytest.raises(B.PdcError):  # codeword array too small
        acc.pdsch_encode(tbs, data, cw_len - 1)


def test_explicit_rate_matched_lengths_give_the_same_codewords(acc):
    """nof_ch_symbols == 0: the code-block lengths come as hal::hw_pdsch_encoder_configuration carries them (cw_length_a,
    cw_length_b, nof_short_segments); same codewords as the derived lengths, and the contract of that mode."""
    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    cases = list(CASES)
    tbs, data, cw_len = make_batch(B, cases)
    want = acc.pdsch_encode(tbs, data, cw_len)
    explicit = tbs.copy()
    for d, (bg, nbits, Qm, L, nsym, rv, Nref) in zip(explicit, cases):
        C = B.pdsch_segmentation(bg, nbits)[0]
        Es = S.rm_lengths(C, nsym * Qm, L, Qm)
        nshort = sum(1 for e in Es if e == Es[0])
        d["nof_layers"], d["nof_ch_symbols"] = 0, 0
        d["cw_length_a"], d["cw_length_b"], d["nof_short_segments"] = Es[0], (Es[-1] if nshort < C else 0), nshort
    got = acc.pdsch_encode(explicit, data, cw_len)
    assert np.array_equal(got, want)
    one = explicit[-2:-1].copy()  # a segmented transport block
    assert B.pdsch_segmentation(int(one[0]["bg"]), int(one[0]["tbs_bits"]))[0] > 1
    one[0]["tb_offset"], one[0]["cw_offset"] = 0, 0
    for field, value in (("cw_length_a", 0), ("cw_length_a", int(one[0]["cw_length_a"]) + 1), ("nof_short_segments", 0),
                         ("nof_short_segments", 153), ("reserved", 1)):
        bad = one.copy()
        bad[0][field] = value
        with pytest.raises(B.PdcError):
            acc.pdsch_encode(bad, np.zeros(200000, np.uint8), 4 << 20)
    mixed = tbs[:1].copy()  # both forms at once
    mixed[0]["cw_length_a"] = 8
    with pytest.raises(B.PdcError):
        acc.pdsch_encode(mixed, data, cw_len)


def test_encode_then_decode_on_the_device(acc):
    """Downlink mirror meets uplink chain: the device encodes a 152-code-block transport block, the codeword bits become LLRs
    (+-20, three in a thousand flipped to weak wrong values) on the device, and pdc_decode_device_tbs returns the transport block."""
    import torch

    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    case = CASES[-1]
    bg, nbits, Qm, L, nsym, rv, Nref = case
    tbs, data, cw_len = make_batch(B, [case], len(CASES) - 1)
    d_tb = torch.from_numpy(data).cuda()
    d_cw = torch.zeros(cw_len, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    acc.pdsch_encode_device(tbs, d_tb.data_ptr(), d_cw.data_ptr(), stream)
    llr = (20 - 40 * d_cw.to(torch.int16))
    noise = torch.randint(0, 1000, (cw_len,), device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    llr = torch.where(noise < 3, -(llr // 10), llr).to(torch.int8).contiguous()
    seg = S.segmentation(bg, nbits)
    Es = S.rm_lengths(seg["C"], cw_len, L, Qm)
    C, stride = seg["C"], (seg["K"] + 31) // 32 * 4
    descs, tbd = B.make_descs(C), np.zeros(1, B.TB_DESC_DTYPE)
    o = 0
    for q in range(C):
        d = descs[q]
        d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = bg, seg["Z"], Qm, rv, pyoracle.CRC24B, 6
        d["flags"], d["E"], d["nof_filler"] = B.PDC_F_NEW_DATA | B.PDC_F_EARLY_STOP, Es[q], seg["F"]
        d["llr_offset"], d["soft_offset"], d["out_offset"] = o, q * 25344, q * stride
        o += Es[q]
    tbd[0]["first_cb"], tbd[0]["nof_cbs"], tbd[0]["tbs_bits"] = 0, C, nbits
    d_soft = torch.zeros(C * 25344, dtype=torch.int8, device="cuda")
    d_bits = torch.zeros(C * stride, dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(C * 8, dtype=torch.uint8, device="cuda")
    d_out = torch.zeros((nbits // 8 + 3) // 4 * 4, dtype=torch.uint8, device="cuda")
    d_tbres = torch.zeros(4, dtype=torch.uint8, device="cuda")
    acc.decode_device_tbs(descs, tbd, llr.data_ptr(), d_soft.data_ptr(), d_bits.data_ptr(), d_res.data_ptr(),
                          d_out.data_ptr(), d_tbres.data_ptr(), stream)
    torch.cuda.synchronize()
    tres = d_tbres.cpu().numpy().view(B.TB_RESULT_DTYPE)
    assert tres[0]["tb_crc_ok"] == 1 and tres[0]["nof_cbs_ok"] == C
    assert np.array_equal(d_out.cpu().numpy()[:nbits // 8], data)
