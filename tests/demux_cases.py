"""Configurations for the UL-SCH demultiplexer tests: the parameter space of the reference's own vector test
(tests/unittests/phy/upper/channel_processors/pusch/ulsch_demultiplex_test_data.h: every modulation, 1-4 layers, both DM-RS types,
1-4 DM-RS symbols, 0 / 1 / 2 / many HARQ-ACK bits with and without reserved REs) widened with CSI Part 1 and Part 2 of 0 / 1 / 2 /
many bits, both ways of learning the CSI Part 2 size, short allocations and the largest one. Seeded; no reference needed."""
import numpy as np

DMRS_MASKS = ([2], [2, 11], [2, 7, 11], [2, 5, 8, 11], [0], [3, 4], [2, 3, 10, 11])


def _mask(symbols):
    m = 0
    for l in symbols:
        m |= 1 << l
    return m


def make_case(rng, qm=None, layers=None, nof_prb=None, harq_bits=None, csi1_bits=None, csi2_bits=None, start=0, nsym=14,
              dmrs=None, csi2_mode=None):
    qm = int(rng.choice([1, 2, 4, 6, 8])) if qm is None else qm
    layers = int(rng.integers(1, 5)) if layers is None else layers
    nof_prb = int(rng.integers(1, 53)) if nof_prb is None else nof_prb
    dmrs_type = int(rng.integers(1, 3))
    cdm = int(rng.integers(1, 3 if dmrs_type == 1 else 4))
    symbols = DMRS_MASKS[int(rng.integers(0, len(DMRS_MASKS)))] if dmrs is None else dmrs
    symbols = [l for l in symbols if start <= l < start + nsym] or [start]
    nbpr = qm * layers
    harq_bits = int(rng.choice([0, 1, 2, 3, 10, 40])) if harq_bits is None else harq_bits
    csi1_bits = int(rng.choice([0, 0, 1, 2, 7, 30])) if csi1_bits is None else csi1_bits
    csi2_bits = (int(rng.choice([0, 1, 2, 9, 50])) if csi1_bits else 0) if csi2_bits is None else csi2_bits
    M = 12 * nof_prb  # REs of a symbol without DM-RS
    free_syms = max(1, nsym - len(symbols) - 2)
    cap = M * free_syms  # a loose bound on the REs that can carry UCI

    def enc(bits, share):
        if bits == 0:
            return 0
        return nbpr * int(rng.integers(1, max(2, int(cap * share))))

    cfg = {"qm": qm, "nof_layers": layers, "nof_prb": nof_prb, "start_symbol": start, "nof_symbols": nsym,
           "dmrs_type": dmrs_type, "dmrs_symbol_mask": _mask(symbols), "nof_cdm_groups_without_data": cdm,
           "nof_harq_ack_bits": harq_bits, "nof_enc_harq_ack_bits": enc(harq_bits, 0.25),
           "nof_csi_part1_bits": csi1_bits, "nof_enc_csi_part1_bits": enc(csi1_bits, 0.2),
           "nof_csi_part2_bits": csi2_bits, "nof_enc_csi_part2_bits": enc(csi2_bits, 0.2),
           "csi2_mode": int(rng.integers(0, 2)) if csi2_mode is None else csi2_mode,
           "c_init": int(rng.integers(1, 1 << 31))}
    # Reserved REs exist when HARQ-ACK has at most two bits (TS 38.212 6.2.7 step 1); sometimes more of them than needed.
    if harq_bits <= 2 and (harq_bits > 0 or rng.integers(0, 4) == 0):
        cfg["nof_harq_ack_rvd"] = max(cfg["nof_enc_harq_ack_bits"], nbpr * int(rng.integers(1, max(2, cap // 4))))
    else:
        cfg["nof_harq_ack_rvd"] = 0
    return cfg


def cases():
    rng = np.random.default_rng(2024)
    out = []
    for qm in (1, 2, 4, 6, 8):  # the placeholder paths of every modulation, every UCI kind
        for bits in (1, 2):
            out.append(make_case(rng, qm=qm, layers=int(rng.integers(1, 3)), nof_prb=int(rng.integers(2, 12)), harq_bits=bits,
                                 csi1_bits=bits, csi2_bits=3 - bits))
    for _ in range(40):
        out.append(make_case(rng))
    out.append(make_case(rng, nof_prb=1, qm=2, layers=1, harq_bits=1, csi1_bits=0, csi2_bits=0))
    out.append(make_case(rng, start=2, nsym=6, dmrs=[2], nof_prb=20))
    out.append(make_case(rng, start=4, nsym=10, dmrs=[4, 9], nof_prb=7, harq_bits=2, csi1_bits=30, csi2_bits=50, csi2_mode=1))
    out.append(make_case(rng, nof_prb=275, qm=8, layers=4, harq_bits=10, csi1_bits=30, csi2_bits=50, dmrs=[2, 11]))
    out.append(make_case(rng, nof_prb=273, qm=8, layers=4, harq_bits=0, csi1_bits=0, csi2_bits=0, dmrs=[2]))
    return out


def nof_softbits(cfg):
    nbpr = cfg["qm"] * cfg["nof_layers"]
    per_prb_dmrs = cfg["nof_cdm_groups_without_data"] * (6 if cfg["dmrs_type"] == 1 else 4)
    n = 0
    for l in range(cfg["start_symbol"], cfg["start_symbol"] + cfg["nof_symbols"]):
        n += ((12 - per_prb_dmrs) if (cfg["dmrs_symbol_mask"] >> l) & 1 else 12) * cfg["nof_prb"] * nbpr
    return n


def codeword(cfg, seed):
    """Descrambled soft bits of the whole allocation: valid LLRs ([-120, 120] and the +-127 infinities,
    log_likelihood_ratio.h:60-66)."""
    n = nof_softbits(cfg)
    rng = np.random.default_rng(seed)
    v = rng.integers(-120, 121, n, dtype=np.int64)
    inf = rng.integers(0, 40, n)
    v = np.where(inf == 0, 127, np.where(inf == 1, -127, v))
    return v.astype(np.int8)


def fill_desc(d, cfg):
    """cfg -> one pdc_ulsch_demux_desc record (numpy, srsran_project_b200.binding.ULSCH_DEMUX_DTYPE); offsets left alone."""
    d["Qm"], d["nof_layers"], d["nof_prb"] = cfg["qm"], cfg["nof_layers"], cfg["nof_prb"]
    d["start_symbol_index"], d["nof_symbols"] = cfg["start_symbol"], cfg["nof_symbols"]
    d["dmrs_type"], d["dmrs_symbol_mask"] = cfg["dmrs_type"], cfg["dmrs_symbol_mask"]
    d["nof_cdm_groups_without_data"] = cfg["nof_cdm_groups_without_data"]
    d["nof_harq_ack_rvd"] = cfg["nof_harq_ack_rvd"]
    d["nof_harq_ack_bits"], d["nof_enc_harq_ack_bits"] = cfg["nof_harq_ack_bits"], cfg["nof_enc_harq_ack_bits"]
    d["nof_csi_part1_bits"], d["nof_enc_csi_part1_bits"] = cfg["nof_csi_part1_bits"], cfg["nof_enc_csi_part1_bits"]
    d["nof_csi_part2_bits"], d["nof_enc_csi_part2_bits"] = cfg["nof_csi_part2_bits"], cfg["nof_enc_csi_part2_bits"]
    d["c_init"] = cfg["c_init"]
    d["flags"] = 1 if (cfg["csi2_mode"] == 1 and cfg["nof_enc_csi_part2_bits"] != 0) else 0  # PDC_DEMUX_F_CSI2_AFTER_CSI1
