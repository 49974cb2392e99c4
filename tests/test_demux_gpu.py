"""UL-SCH demultiplexer on the device: pdc_ulsch_demultiplex / pdc_ulsch_demultiplex_device against the golden
streams recorded from the compiled ulsch_demultiplex_impl, against the numpy restatement on a whole batch in one submission (word
and byte paths), and behind the device demodulator + descrambler. Bit-exact."""
import hashlib
import os

import numpy as np
import pytest

from oracle import demux_oracle as D
from tests import demux_cases as C

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ulsch_demux_v1.npz")


@pytest.fixture(scope="module")
def acc():
    import srsran_project_b200 as pdc

    a = pdc.PuschDecCuda(device=0, max_batch_cbs=64, max_batch_llrs=1 << 20, harq_slots=0, nof_tickets=2)
    yield a
    a.close()


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a, np.int8).tobytes()).digest(), np.uint8)


def layout(B, cfgs, seeds, align):
    """One descriptor per codeword; inputs and the four streams of each packed one after the other, every start rounded up
    to `align` bytes (1 exercises the byte path of the kernel, 4 the word path where nbpr allows)."""
    up = lambda v: (v + align - 1) // align * align  # noqa: E731
    descs = np.zeros(len(cfgs), B.ULSCH_DEMUX_DTYPE)
    cws, io, oo = [], 0, 1 if align == 1 else 0
    for d, cfg, seed in zip(descs, cfgs, seeds):
        C.fill_desc(d, cfg)
        lens = B.ulsch_demux_lengths(descs[len(cws):len(cws) + 1])
        cw = C.codeword(cfg, seed)
        d["in_offset"] = io
        cws.append((io, cw))
        io = up(io + cw.size)
        for name, n in zip(("sch_offset", "harq_ack_offset", "csi_part1_offset", "csi_part2_offset"), lens[1:]):
            d[name] = oo
            oo = up(oo + n)
    data = np.zeros(io + 4, np.int8)
    for o, cw in cws:
        data[o:o + cw.size] = cw
    return descs, data, oo + 4


def streams_of(B, d, out):
    lens = B.ulsch_demux_lengths(np.array([d]))
    return [out[int(d[f]):int(d[f]) + n] for f, n in zip(("sch_offset", "harq_ack_offset", "csi_part1_offset", "csi_part2_offset"),
                                                         lens[1:])]


def test_golden_streams_one_by_one(acc):
    from srsran_project_b200 import binding as B

    g = np.load(GOLDEN)
    for k, cfg in enumerate(C.cases()):
        descs, data, out_len = layout(B, [cfg], [100 + k], 4)
        out = acc.ulsch_demultiplex(descs, data, out_len)
        for q, a in enumerate(streams_of(B, descs[0], out)):
            assert a.size == g["lengths"][k, q]
            assert np.array_equal(digest(a), g["digests"][k, q]), (k, q, cfg)


@pytest.mark.parametrize("align", [1, 4])
def test_batch_in_one_submission_matches_restatement(acc, align):
    import torch

    from srsran_project_b200 import binding as B

    cfgs = C.cases()[:-2]
    rng = np.random.default_rng(5)
    cfgs += [C.make_case(rng, nof_prb=int(rng.integers(1, 40))) for _ in range(20)]
    descs, data, out_len = layout(B, cfgs, range(300, 300 + len(cfgs)), align)
    d_in = torch.from_numpy(data).cuda()
    d_out = torch.full((out_len,), 77, dtype=torch.int8, device="cuda")
    acc.ulsch_demultiplex_device(descs, d_in.data_ptr(), d_out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out = d_out.cpu().numpy()
    written = np.zeros(out_len, bool)
    for k, (d, cfg) in enumerate(zip(descs, cfgs)):
        want = D.demultiplex(cfg, C.codeword(cfg, 300 + k))
        for q, (a, b) in enumerate(zip(streams_of(B, d, out), want)):
            assert np.array_equal(a, b), (k, q, cfg, np.nonzero(a != b)[0][:4] if a.size == b.size else (a.size, b.size))
        for f, b in zip(("sch_offset", "harq_ack_offset", "csi_part1_offset", "csi_part2_offset"), want):
            written[int(d[f]):int(d[f]) + b.size] = True
    assert (out[~written] == 77).all()  # nothing outside the streams is touched


def test_behind_the_device_demodulator(acc):
    """symbols -> pdc_demodulate_device (+ descrambling) -> pdc_ulsch_demultiplex_device on one stream: the UL-SCH stream that
    reaches the code-block chain equals the restatement applied to the demodulator's output."""
    import torch

    from srsran_project_b200 import binding as B
    from tools import demod_stimulus as S

    rng = np.random.default_rng(9)
    cfg = C.make_case(rng, qm=6, layers=2, nof_prb=24, harq_bits=2, csi1_bits=30, csi2_bits=9, dmrs=[2, 11], csi2_mode=1)
    n_soft = C.nof_softbits(cfg)
    nsym = n_soft // cfg["qm"]
    sym, nv, _ = S.equalised(nsym, cfg["qm"], seed=31)
    seg = np.zeros(1, B.DEMOD_SEG_DTYPE)
    seg[0]["nof_symbols"], seg[0]["mod"], seg[0]["flags"] = nsym, cfg["qm"], B.PDC_DEMOD_F_DESCRAMBLE
    seg[0]["c_init"] = cfg["c_init"]
    d_sym = torch.from_numpy(np.ascontiguousarray(sym).view(np.float32)).cuda()
    d_nv = torch.from_numpy(nv).cuda()
    d_llr = torch.zeros(n_soft + 4, dtype=torch.int8, device="cuda")
    descs = np.zeros(1, B.ULSCH_DEMUX_DTYPE)
    C.fill_desc(descs[0], cfg)
    lens = B.ulsch_demux_lengths(descs)
    o = 0
    for f, n in zip(("sch_offset", "harq_ack_offset", "csi_part1_offset", "csi_part2_offset"), lens[1:]):
        descs[0][f] = o
        o += (n + 3) // 4 * 4
    d_out = torch.zeros(o + 4, dtype=torch.int8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    acc.demodulate_device(seg, d_sym.data_ptr(), d_nv.data_ptr(), d_llr.data_ptr(), stream)
    acc.ulsch_demultiplex_device(descs, d_llr.data_ptr(), d_out.data_ptr(), stream)
    torch.cuda.synchronize()
    llr = d_llr.cpu().numpy()[:n_soft]
    want = D.demultiplex(cfg, llr)
    for a, b in zip(streams_of(B, descs[0], d_out.cpu().numpy()), want):
        assert np.array_equal(a, b)
    assert want[1].size and want[2].size and want[3].size


def test_contract_violations(acc):
    from srsran_project_b200 import binding as B

    cfg = C.cases()[12]
    descs, data, out_len = layout(B, [cfg], [1], 4)
    with pytest.raises(B.PdcError):  # input array shorter than the allocation
        acc.ulsch_demultiplex(descs, data[:C.nof_softbits(cfg) - 1], out_len)
    with pytest.raises(B.PdcError):  # output array shorter than the streams
        acc.ulsch_demultiplex(descs, data, int(descs[0]["sch_offset"]) + 10)
    bad = descs.copy()
    bad[0]["nof_enc_csi_part1_bits"] = 1 << 27  # does not fit the allocation: the reference asserts at the end of the codeword
    bad[0]["nof_csi_part1_bits"] = 5
    with pytest.raises(B.PdcError):
        acc.ulsch_demultiplex(bad, data, out_len)


def test_pusch_with_uci_symbols_to_transport_block(acc):
    """The whole receive side of a PUSCH that carries UCI, on the device: a transport block is encoded for the UL-SCH REs that the
    multiplexing leaves, laid into the codeword around random HARQ-ACK / CSI bits (the REs a 2-bit HARQ-ACK punctures carry
    noise for the UL-SCH), scrambled and modulated; then pdc_demodulate_device (+ descrambling) -> pdc_ulsch_demultiplex_device
    -> pdc_decode_device_tbs on one stream return the transport block, and the UCI streams equal the restatement's."""
    import torch

    from srsran_project_b200 import binding as B
    from tools import demod_stimulus as DS
    from tools import stimulus as T

    rng = np.random.default_rng(21)
    qm, layers = 6, 2
    cfg = C.make_case(rng, qm=qm, layers=layers, nof_prb=40, harq_bits=2, csi1_bits=30, csi2_bits=9, dmrs=[2, 11], csi2_mode=0)
    # A 2-bit HARQ-ACK occupies a few dozen REs: all of them puncture the first code block, which must stay decodable.
    cfg["nof_enc_harq_ack_bits"], cfg["nof_harq_ack_rvd"] = 40 * qm * layers, 60 * qm * layers
    n_soft = C.nof_softbits(cfg)
    # Where every soft bit of the codeword goes: push 1-based positions through the restatement (0 = punctured in UL-SCH).
    where = D.demultiplex(cfg, np.arange(1, n_soft + 1), dtype=np.int64)
    sch_pos = np.abs(where[0])
    G = int(sch_pos.size)
    assert G % (qm * layers) == 0
    tbs_bits = int(G * 0.6) // 8 * 8
    tb = rng.integers(0, 2, tbs_bits, dtype=np.uint8)
    seg, sch_bits, _ = T.tb_encode(1, tbs_bits, tb, 0, qm, 0, layers, G)
    tx = rng.integers(0, 2, n_soft, dtype=np.uint8)  # UCI positions and punctured REs: anything
    keep = sch_pos > 0
    tx[sch_pos[keep] - 1] = sch_bits[keep]
    z = DS.modulate(tx ^ DS.gold_sequence(n_soft, cfg["c_init"]), qm)
    sigma2 = 0.004
    z = z + (rng.normal(0, np.sqrt(sigma2 / 2), z.size) + 1j * rng.normal(0, np.sqrt(sigma2 / 2), z.size)).astype(np.complex64)
    sym = np.stack([z.real, z.imag], axis=1).astype(np.float32)
    nv = np.full(z.size, sigma2, np.float32)

    dseg = np.zeros(1, B.DEMOD_SEG_DTYPE)
    dseg[0]["nof_symbols"], dseg[0]["mod"], dseg[0]["flags"], dseg[0]["c_init"] = z.size, qm, B.PDC_DEMOD_F_DESCRAMBLE, cfg["c_init"]
    ud = np.zeros(1, B.ULSCH_DEMUX_DTYPE)
    C.fill_desc(ud[0], cfg)
    lens = B.ulsch_demux_lengths(ud)
    assert lens[0] == n_soft and lens[1] == G
    offs, o = [], 0
    for n in lens[1:]:
        offs.append(o)
        o += (n + 15) // 16 * 16
    ud[0]["sch_offset"], ud[0]["harq_ack_offset"], ud[0]["csi_part1_offset"], ud[0]["csi_part2_offset"] = offs
    Cn, K = seg["C"], seg["K"]
    Es = T.rm_lengths(Cn, G, layers, qm)
    stride = (K + 31) // 32 * 4
    descs, tbd = B.make_descs(Cn), np.zeros(1, B.TB_DESC_DTYPE)
    e0 = 0
    for q in range(Cn):
        d = descs[q]
        d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = 1, seg["Z"], qm, 0, B.PDC_CRC24B if Cn > 1 else B.PDC_CRC24A, 10
        d["flags"], d["E"], d["nof_filler"] = B.PDC_F_NEW_DATA | B.PDC_F_EARLY_STOP, Es[q], seg["F"]
        d["llr_offset"], d["soft_offset"], d["out_offset"] = offs[0] + e0, q * 25344, q * stride
        e0 += Es[q]
    tbd[0]["first_cb"], tbd[0]["nof_cbs"], tbd[0]["tbs_bits"] = 0, Cn, tbs_bits

    d_sym = torch.from_numpy(sym.reshape(-1)).cuda()
    d_nv = torch.from_numpy(nv).cuda()
    d_cw = torch.zeros(n_soft + 16, dtype=torch.int8, device="cuda")
    d_streams = torch.zeros(o + 16, dtype=torch.int8, device="cuda")
    d_soft = torch.zeros(Cn * 25344, dtype=torch.int8, device="cuda")
    d_bits = torch.zeros(Cn * stride, dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(Cn * 8, dtype=torch.uint8, device="cuda")
    d_tb = torch.zeros((tbs_bits // 8 + 3) // 4 * 4, dtype=torch.uint8, device="cuda")
    d_tbres = torch.zeros(4, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    acc.demodulate_device(dseg, d_sym.data_ptr(), d_nv.data_ptr(), d_cw.data_ptr(), stream)
    acc.ulsch_demultiplex_device(ud, d_cw.data_ptr(), d_streams.data_ptr(), stream)
    acc.decode_device_tbs(descs, tbd, d_streams.data_ptr(), d_soft.data_ptr(), d_bits.data_ptr(), d_res.data_ptr(),
                          d_tb.data_ptr(), d_tbres.data_ptr(), stream)
    torch.cuda.synchronize()
    want = D.demultiplex(cfg, d_cw.cpu().numpy()[:n_soft])
    got = d_streams.cpu().numpy()
    for off, w in zip(offs, want):
        assert np.array_equal(got[off:off + w.size], w)
    hard = (got[:G] < 0).astype(np.uint8)
    assert (hard[keep] != sch_bits[keep]).mean() < 0.02, "the demodulated UL-SCH stream is not the transmitted codeword"
    cres = d_res.cpu().numpy().view(B.RESULT_DTYPE)
    tres = d_tbres.cpu().numpy().view(B.TB_RESULT_DTYPE)
    assert tres[0]["tb_crc_ok"] == 1 and tres[0]["nof_cbs_ok"] == Cn, (cres, tres)
    assert np.array_equal(d_tb.cpu().numpy()[:tbs_bits // 8], np.packbits(tb))
    assert (want[0] == 0).sum() >= 2 * qm * layers  # the punctured REs reached the decoder as erasures


def test_host_buffer_entry_with_several_codewords(acc):
    """pdc_ulsch_demultiplex (host buffers) with several codewords in one call, byte-aligned layout: every stream comes back,
    bytes between the streams are left alone."""
    from srsran_project_b200 import binding as B

    cfgs = C.cases()[3:9] + C.cases()[30:33]
    descs, data, out_len = layout(B, cfgs, range(700, 700 + len(cfgs)), 1)
    out = np.full(out_len, 55, np.int8)
    acc._check(acc.lib.pdc_ulsch_demultiplex(acc.h, descs.ctypes.data, descs.size, data.ctypes.data, data.size, out.ctypes.data,
                                             out.size))
    touched = np.zeros(out_len, bool)
    for k, (d, cfg) in enumerate(zip(descs, cfgs)):
        want = D.demultiplex(cfg, C.codeword(cfg, 700 + k))
        for f, b in zip(("sch_offset", "harq_ack_offset", "csi_part1_offset", "csi_part2_offset"), want):
            o = int(d[f])
            assert np.array_equal(out[o:o + b.size], b), (k, f, cfg)
            touched[o:o + b.size] = True
    assert (out[~touched] == 55).all()
