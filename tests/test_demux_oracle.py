"""UL-SCH demultiplexer, CPU side: the numpy restatement (oracle/demux_oracle.py) against the golden streams
recorded from the compiled, unmodified ulsch_demultiplex_impl (tests/golden/ulsch_demux_v1.npz, tools/gen_golden_demux.py), against
that reference itself when oracle/_ref is present, and the library's host-only planner (pdc_ulsch_demux_lengths) against both."""
import hashlib
import os

import numpy as np
import pytest

from oracle import demux_oracle as D
from tests import demux_cases as C

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ulsch_demux_v1.npz")


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a, np.int8).tobytes()).digest(), np.uint8)


def test_restatement_matches_golden_streams():
    g = np.load(GOLDEN)
    cs = C.cases()
    assert g["digests"].shape[0] == len(cs)
    for k, cfg in enumerate(cs):
        streams = D.demultiplex(cfg, C.codeword(cfg, 100 + k))
        for q, a in enumerate(streams):
            assert a.size == g["lengths"][k, q], (k, q)
            assert np.array_equal(digest(a), g["digests"][k, q]), (k, q, cfg)
            if f"s{k}_{q}" in g.files:
                assert np.array_equal(a, g[f"s{k}_{q}"])


def test_restatement_matches_compiled_reference():
    try:
        ref = D.UlschDemuxReference()
    except FileNotFoundError:
        pytest.skip("oracle/_ref not built (no reference tree on this machine)")
    cs = C.cases()
    rng = np.random.default_rng(77)
    extra = [C.make_case(rng, nof_prb=int(rng.integers(1, 30))) for _ in range(25)]  # not in the golden file
    for k, cfg in enumerate(cs[:20] + extra):
        cw = C.codeword(cfg, 900 + k)
        want = ref.demultiplex(cfg, cw, block=int(rng.integers(0, 3)) * 311)
        got = D.demultiplex(cfg, cw)
        for q, (a, b) in enumerate(zip(got, want)):
            assert np.array_equal(a, b), (k, q, cfg)


def test_golden_covers_the_paths():
    """The fixture is only worth something if it walks every branch: placeholders of 1 and 2 bits for each UCI kind and each
    modulation, reserved REs with and without HARQ-ACK, HARQ-ACK of more than two bits, both CSI Part 2 modes, and a case where
    the two modes differ."""
    cs = C.cases()
    for qm in (1, 2, 4, 6, 8):
        for bits in (1, 2):
            assert any(c["qm"] == qm and c["nof_harq_ack_bits"] == bits for c in cs)
            assert any(c["qm"] == qm and c["nof_csi_part1_bits"] == bits for c in cs)
            assert any(c["qm"] == qm and c["nof_csi_part2_bits"] == bits for c in cs)
    assert any(c["nof_harq_ack_rvd"] > 0 and c["nof_harq_ack_bits"] == 0 for c in cs)
    assert any(c["nof_harq_ack_bits"] > 2 for c in cs)
    assert any(c["nof_enc_csi_part2_bits"] and c["csi2_mode"] == 0 for c in cs)
    assert any(c["nof_enc_csi_part2_bits"] and c["csi2_mode"] == 1 for c in cs)
    differ = 0
    for k, c in enumerate(cs):
        if c["nof_enc_csi_part2_bits"]:
            cw = C.codeword(c, 100 + k)
            a, b = D.demultiplex(dict(c, csi2_mode=0), cw), D.demultiplex(dict(c, csi2_mode=1), cw)
            differ += any(not np.array_equal(x, y) for x, y in zip(a, b))
    assert differ >= 3


def test_library_planner_lengths_and_contract():
    from srsran_project_b200 import binding as B

    g = np.load(GOLDEN)
    cs = C.cases()
    for k, cfg in enumerate(cs):
        d = np.zeros(1, B.ULSCH_DEMUX_DTYPE)
        C.fill_desc(d[0], cfg)
        lens = B.ulsch_demux_lengths(d)
        assert lens[0] == C.nof_softbits(cfg)
        assert list(lens[1:]) == [int(v) for v in g["lengths"][k]], (k, cfg)
    base = np.zeros(1, B.ULSCH_DEMUX_DTYPE)
    C.fill_desc(base[0], cs[10])
    for field, value in (("Qm", 3), ("nof_layers", 0), ("nof_layers", 5), ("nof_prb", 0), ("nof_prb", 276), ("nof_symbols", 0),
                         ("nof_symbols", 15), ("dmrs_type", 3), ("nof_cdm_groups_without_data", 4), ("dmrs_symbol_mask", 0),
                         ("dmrs_symbol_mask", 0x3fff), ("flags", 2), ("reserved", 1), ("reserved2", 1),
                         ("nof_enc_harq_ack_bits", 1 << 28)):
        bad = base.copy()
        bad[0][field] = value
        with pytest.raises(B.PdcError):
            B.ulsch_demux_lengths(bad)
    # encoded bits that are not whole REs, UCI without message bits, reserved REs with a long HARQ-ACK
    nbpr = int(base[0]["Qm"]) * int(base[0]["nof_layers"])
    for fields in ({"nof_harq_ack_bits": 4, "nof_enc_harq_ack_bits": nbpr * 5 + 1, "nof_harq_ack_rvd": 0},
                   {"nof_harq_ack_bits": 0, "nof_enc_harq_ack_bits": nbpr, "nof_harq_ack_rvd": 0},
                   {"nof_harq_ack_bits": 4, "nof_enc_harq_ack_bits": nbpr * 5, "nof_harq_ack_rvd": nbpr * 5}):
        bad = base.copy()
        for f, v in fields.items():
            bad[0][f] = v
        with pytest.raises(B.PdcError):
            B.ulsch_demux_lengths(bad)


def test_library_planner_fuzz_against_restatement():
    """2000 random configurations, many of them with UCI that does NOT fit: the host planner either returns the stream lengths the
    restatement delivers, or rejects exactly the configurations for which the reference would assert at on_end_codeword (a UCI
    stream not delivered in full)."""
    from srsran_project_b200 import binding as B

    rng = np.random.default_rng(4242)
    accepted = rejected = 0
    for k in range(2000):
        cfg = C.make_case(rng, nof_prb=int(rng.integers(1, 6)), nsym=int(rng.integers(1, 15)), start=0)
        if cfg["start_symbol"] + cfg["nof_symbols"] > 14:
            continue
        if rng.integers(0, 3) == 0:  # stretch one encoded length so that it may not fit
            f = ("nof_enc_harq_ack_bits", "nof_enc_csi_part1_bits", "nof_enc_csi_part2_bits")[int(rng.integers(0, 3))]
            if cfg[f]:
                cfg[f] *= int(rng.integers(2, 12))
                if f == "nof_enc_harq_ack_bits" and cfg["nof_harq_ack_bits"] <= 2:
                    cfg["nof_harq_ack_rvd"] = max(cfg["nof_harq_ack_rvd"], cfg[f])
        # the restatement needs a symbol without DM-RS after the first DM-RS symbol, like the reference
        m = cfg["dmrs_symbol_mask"]
        first = next(l for l in range(14) if (m >> l) & 1)
        if all((m >> l) & 1 for l in range(first, 14)):
            continue
        streams = D.demultiplex(cfg, np.zeros(C.nof_softbits(cfg), np.int8))
        complete = (streams[1].size == cfg["nof_enc_harq_ack_bits"] and streams[2].size == cfg["nof_enc_csi_part1_bits"]
                    and streams[3].size == cfg["nof_enc_csi_part2_bits"])
        d = np.zeros(1, B.ULSCH_DEMUX_DTYPE)
        C.fill_desc(d[0], cfg)
        if complete:
            lens = B.ulsch_demux_lengths(d)
            assert list(lens) == [C.nof_softbits(cfg)] + [s.size for s in streams], (k, cfg)
            accepted += 1
        else:
            with pytest.raises(B.PdcError):
                B.ulsch_demux_lengths(d)
            rejected += 1
    assert accepted > 800 and rejected > 100, (accepted, rejected)
