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
