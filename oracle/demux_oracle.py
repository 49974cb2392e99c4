"""TEST INFRASTRUCTURE ONLY — not part of the product path (only tests/, smoke() and bench.py's cpu_baseline leg may import it).

numpy restatement of the reference's UL-SCH demultiplexer:
lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.cpp — RE set construction per OFDM symbol (:330-463), the
demultiplexing of one OFDM symbol (:465-590) and the 1- / 2-bit placeholder handling (:106-196). Pinned against the compiled
reference (oracle/_ref, `UlschDemuxReference`) by tests/test_demux_oracle.py and against tests/golden/ulsch_demux_v1.npz.

A configuration is a dict with the fields of ulsch_demultiplex::configuration (ulsch_demultiplex.h:39-66) as integers:
qm, nof_layers, nof_prb, start_symbol, nof_symbols, nof_harq_ack_rvd, dmrs_type (1 | 2), dmrs_symbol_mask (bit l = symbol l),
nof_cdm_groups_without_data, nof_harq_ack_bits, nof_enc_harq_ack_bits, nof_csi_part1_bits, nof_enc_csi_part1_bits,
nof_csi_part2_bits, nof_enc_csi_part2_bits, csi2_mode (0: CSI Part 2 known from the start, as the reference's unit test sets
it; 1: set when the CSI Part 1 stream ends, as pusch_processor_impl.cpp:73-98 does), c_init.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CFG_FIELDS = ("qm", "nof_layers", "nof_prb", "start_symbol", "nof_symbols", "nof_harq_ack_rvd", "dmrs_type",
              "dmrs_symbol_mask", "nof_cdm_groups_without_data", "nof_harq_ack_bits", "nof_enc_harq_ack_bits",
              "nof_csi_part1_bits", "nof_enc_csi_part1_bits", "nof_csi_part2_bits", "nof_enc_csi_part2_bits", "csi2_mode",
              "c_init")


def gold_bits(c_init, n):
    """TS 38.211 5.2.1 (pseudo_random_generator_impl.cpp): n bits of c() for c_init."""
    total = 1600 + n
    x1 = np.zeros(total + 31, np.uint8)
    x2 = np.zeros(total + 31, np.uint8)
    x1[0] = 1
    for i in range(31):
        x2[i] = (c_init >> i) & 1
    for i in range(total):
        x1[i + 31] = x1[i + 3] ^ x1[i]
        x2[i + 31] = x2[i + 3] ^ x2[i + 2] ^ x2[i + 1] ^ x2[i]
    return x1[1600:1600 + n] ^ x2[1600:1600 + n]


def _select(base, d, m):
    """re_set_select (:77-100): of the set positions of `base`, every d-th one, the first m of them."""
    idx = np.nonzero(base)[0][::d][:m]
    out = np.zeros(base.size, bool)
    out[idx] = True
    return out


def _step(remainder, avail):
    """d and the number of selected REs for `remainder` REs still to place among `avail` candidates (:361-366 etc.)."""
    if remainder < avail:
        return avail // remainder, remainder
    return 1, avail


def symbol_lengths(cfg):
    """Soft bits per OFDM symbol of the allocation."""
    nbpr = cfg["qm"] * cfg["nof_layers"]
    per_prb_dmrs = cfg["nof_cdm_groups_without_data"] * (6 if cfg["dmrs_type"] == 1 else 4)
    out = []
    for l in range(cfg["start_symbol"], cfg["start_symbol"] + cfg["nof_symbols"]):
        dm = (cfg["dmrs_symbol_mask"] >> l) & 1
        out.append(((12 - per_prb_dmrs) if dm else 12) * cfg["nof_prb"] * nbpr)
    return out


def _placeholder(re_data, mask, qm, nbits):
    """on_uci_placeholder_1bit / _2bit (:106-196) for the soft bits of one RE (every layer's symbol in turn)."""
    out = re_data.copy()
    if qm == 1:
        return out
    for s in range(re_data.size // qm):
        d, k = re_data[s * qm:(s + 1) * qm], mask[s * qm:(s + 1) * qm]
        o = out[s * qm:(s + 1) * qm]
        if nbits == 1:
            o[1] = -d[1] if (k[0] ^ k[1]) else d[1]
        for b in range(2, qm):
            o[b] = -d[b] if k[b] else d[b]
    return out


def demultiplex(cfg, llrs, dtype=np.int8):
    """Returns (sch, harq_ack, csi_part1, csi_part2) arrays for the descrambled codeword `llrs` (int8 soft bits; a wider dtype lets
    a test push position indices through to learn where each soft bit goes)."""
    mask_l = cfg["dmrs_symbol_mask"]
    first_dmrs = next(l for l in range(14) if (mask_l >> l) & 1)
    l1 = next(l for l in range(first_dmrs, 14) if not (mask_l >> l) & 1)  # :32-46
    l1_csi = next(l for l in range(14) if not (mask_l >> l) & 1)  # :48-55
    nbpr = cfg["qm"] * cfg["nof_layers"]
    lens = symbol_lengths(cfg)
    seq = gold_bits(cfg["c_init"], int(sum(lens)))
    llrs = np.asarray(llrs, dtype)
    assert llrs.size == sum(lens)
    m_rvd = m_harq = m_csi1 = m_csi2 = 0
    csi2_active = cfg["nof_enc_csi_part2_bits"] != 0 and cfg["csi2_mode"] == 0
    out = {"sch": [], "harq": [], "csi1": [], "csi2": []}
    off = 0
    for n_sym, l in zip(lens, range(cfg["start_symbol"], cfg["start_symbol"] + cfg["nof_symbols"])):
        M = n_sym // nbpr
        dm = bool((mask_l >> l) & 1)
        ulsch = np.ones(M, bool)
        uci = np.full(M, not dm)
        rvd = np.zeros(M, bool)
        harq = np.zeros(M, bool)
        csi1 = np.zeros(M, bool)
        csi2 = np.zeros(M, bool)
        M_uci = int(uci.sum())
        rem_rvd = (cfg["nof_harq_ack_rvd"] - m_rvd) // nbpr
        if l >= l1 and M_uci > 0 and rem_rvd > 0:  # step 1 (:355-373)
            d, m = _step(rem_rvd, M_uci)
            rvd = _select(ulsch, d, m)
            m_rvd += m * nbpr
        rem_harq = (cfg["nof_enc_harq_ack_bits"] - m_harq) // nbpr
        if l >= l1 and M_uci > 0 and cfg["nof_harq_ack_bits"] > 2 and rem_harq > 0:  # step 2 (:375-398)
            d, m = _step(rem_harq, M_uci)
            harq = _select(uci, d, m)
            ulsch &= ~harq
            uci &= ~harq
            M_uci = int(uci.sum())
            m_harq += m * nbpr
        rem_csi1 = (cfg["nof_enc_csi_part1_bits"] - m_csi1) // nbpr
        M_rvd = int(rvd.sum())
        if l >= l1_csi and (M_uci - M_rvd) > 0 and rem_csi1 > 0:  # step 3 (:400-425)
            d, m = _step(rem_csi1, M_uci - M_rvd)
            csi1 = _select(uci & ~rvd, d, m)
            ulsch &= ~csi1
            uci &= ~csi1
            m_csi1 += m * nbpr
        if csi2_active or (csi1.any() and m_csi1 == cfg["nof_enc_csi_part1_bits"] and cfg["csi2_mode"] == 1
                           and cfg["nof_enc_csi_part2_bits"] != 0):
            # configure_csi_part2_current_ofdm_symbol (:443-463). With the feedback (:241-251) it runs between the CSI Part 1 and
            # CSI Part 2 loops of the symbol in which CSI Part 1 ends: the sets it sees are the ones step 3 left.
            csi2_active = True
            Mu = int(uci.sum())
            rem = (cfg["nof_enc_csi_part2_bits"] - m_csi2) // nbpr
            if l >= l1_csi and Mu > 0 and rem > 0:
                d, m = _step(rem, Mu)
                csi2 = _select(uci, d, m)
                ulsch &= ~csi2
                uci &= ~csi2
                m_csi2 += m * nbpr
        if M_rvd > 0 and cfg["nof_harq_ack_bits"] <= 2 and rem_harq > 0:  # step 5 (:430-441)
            d, m = _step(rem_harq, M_rvd)
            harq = _select(rvd, d, m)
            m_harq += m * nbpr
        # demux_current_ofdm_symbol (:465-590)
        data = llrs[off:off + n_sym].copy()
        sq = seq[off:off + n_sym]
        for i in np.nonzero(harq)[0]:
            a = slice(i * nbpr, (i + 1) * nbpr)
            if cfg["nof_harq_ack_bits"] in (1, 2):
                out["harq"].append(_placeholder(data[a], sq[a], cfg["qm"], cfg["nof_harq_ack_bits"]))
                data[a] = 0
            else:
                out["harq"].append(data[a].copy())
        for name, sel, nbits in (("csi1", csi1, cfg["nof_csi_part1_bits"]), ("csi2", csi2, cfg["nof_csi_part2_bits"])):
            for i in np.nonzero(sel)[0]:
                a = slice(i * nbpr, (i + 1) * nbpr)
                if nbits in (1, 2):
                    out[name].append(_placeholder(data[a], sq[a], cfg["qm"], nbits))
                else:
                    out[name].append(data[a].copy())
        idx = np.nonzero(ulsch)[0]
        if idx.size:
            out["sch"].append(data.reshape(M, nbpr)[idx].reshape(-1))
        off += n_sym
    cat = lambda v: np.concatenate(v).astype(dtype) if v else np.zeros(0, dtype)  # noqa: E731
    return cat(out["sch"]), cat(out["harq"]), cat(out["csi1"]), cat(out["csi2"])


class UlschDemuxReference:
    """The compiled, unmodified ulsch_demultiplex_impl (oracle/_ref/libsrsran_ref.so, oracle/ref_shim_demux.cpp)."""

    def __init__(self):
        path = os.path.join(HERE, "_ref", "libsrsran_ref.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = ctypes.CDLL(path)
        if not hasattr(self.lib, "ref_ulsch_demultiplex"):
            raise FileNotFoundError("ref_ulsch_demultiplex not in " + path)

    def demultiplex(self, cfg, llrs, block=0):
        c = np.array([cfg[f] for f in CFG_FIELDS] + [block], np.uint32)
        llrs = np.ascontiguousarray(llrs, np.int8)
        outs = [np.zeros(llrs.size, np.int8) for _ in range(4)]
        lens = np.zeros(4, np.uint32)
        vp = ctypes.c_void_p
        rc = self.lib.ref_ulsch_demultiplex(vp(c.ctypes.data), vp(llrs.ctypes.data), ctypes.c_uint(llrs.size),
                                            *[vp(o.ctypes.data) for o in outs], ctypes.c_uint(llrs.size),
                                            vp(lens.ctypes.data))
        assert rc == 0
        return tuple(o[:int(n)] for o, n in zip(outs, lens))
