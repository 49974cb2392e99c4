"""TEST INFRASTRUCTURE ONLY — ctypes access to the CPU oracle (libpdc_oracle.so) and, when present, to the compiled
reference (oracle/_ref/libsrsran_ref.so).  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
import this module; nothing under srsran_project_b200/ does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libpdc_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libsrsran_ref.so")

FLAVOUR_GENERIC, FLAVOUR_AVX2, FLAVOUR_AVX512 = 0, 1, 2
CRC_NONE, CRC16, CRC24A, CRC24B = 0, 1, 2, 3

_c_i8p = ctypes.POINTER(ctypes.c_int8)
_c_u8p = ctypes.POINTER(ctypes.c_uint8)


def _i8(a):
    return a.ctypes.data_as(_c_i8p)


def _u8(a):
    return a.ctypes.data_as(_c_u8p)


def build_oracle():
    """Compiles the C restatement (seconds). Building the checker is not using it."""
    if not os.path.exists(ORACLE_SO) or any(
        os.path.getmtime(os.path.join(HERE, f)) > os.path.getmtime(ORACLE_SO)
        for f in ("pdc_oracle.c", "pdc_oracle_decode.c", "pdc_oracle_demod.c", "pdc_oracle.h")
    ):
        subprocess.check_call(["make", "-C", HERE, os.path.join(HERE, "libpdc_oracle.so")], stdout=subprocess.DEVNULL)


def kb_of(bg):
    return 22 if bg == 1 else 10


def nshort_of(bg):
    return 66 if bg == 1 else 50


class Oracle:
    """The plain-C restatement."""

    def __init__(self):
        build_oracle()
        self.lib = ctypes.CDLL(ORACLE_SO)
        self.lib.orc_crc_bits.restype = ctypes.c_uint32

    def crc(self, crc_id, packed, nbits):
        packed = np.ascontiguousarray(packed, dtype=np.uint8)
        return int(self.lib.orc_crc_bits(crc_id, _u8(packed), int(nbits)))

    def rate_dematch(self, flavour, buf, llrs, new_data, rv, Qm, Nref, nof_filler):
        assert buf.dtype == np.int8 and buf.flags.c_contiguous
        llrs = np.ascontiguousarray(llrs, dtype=np.int8)
        return self.lib.orc_rate_dematch(flavour, _i8(buf), buf.size, _i8(llrs), llrs.size, int(new_data), rv, Qm,
                                         Nref, nof_filler)

    def decode(self, flavour, bg, Z, llrs, nof_filler=0, max_iter=6, crc_id=CRC_NONE, force_decoding=False):
        llrs = np.ascontiguousarray(llrs, dtype=np.int8)
        K = kb_of(bg) * Z
        out = np.zeros((K + 7) // 8, np.uint8)
        it = self.lib.orc_ldpc_decode(flavour, bg, Z, nof_filler, max_iter, crc_id, int(force_decoding), _i8(llrs),
                                      llrs.size, _u8(out))
        return it, out

    def codeblock_decode(self, flavour, bg, Z, buf, llrs, new_data, rv, Qm, Nref, nof_filler, crc_id, max_iter=6,
                         early_stop=True, force_decoding=False):
        llrs = np.ascontiguousarray(llrs, dtype=np.int8)
        K = kb_of(bg) * Z
        out = np.zeros((K + 7) // 8, np.uint8)
        it = self.lib.orc_codeblock_decode(flavour, bg, Z, _i8(buf), _i8(llrs), llrs.size, int(new_data), rv, Qm,
                                           Nref, nof_filler, crc_id, max_iter, int(early_stop), int(force_decoding),
                                           _u8(out))
        return it, out

    def tb_assemble(self, K, nof_filler, cb_packed, cb_crc_ok, tbs_bits, tb_out):
        """pusch_decoder_impl::join_and_notify for one TB. cb_packed: [C, stride] uint8; tb_out (uint8, tbs_bits / 8) is
        written where the reference would write. -> (tb_crc_ok, written)"""
        cb_packed = np.ascontiguousarray(cb_packed, np.uint8)
        ok = np.ascontiguousarray(cb_crc_ok, np.uint8)
        assert tb_out.dtype == np.uint8 and tb_out.size == tbs_bits // 8 and tb_out.flags.c_contiguous
        w = ctypes.c_int(0)
        r = self.lib.orc_tb_assemble(cb_packed.shape[0], K, nof_filler, _u8(cb_packed), cb_packed.shape[1], _u8(ok),
                                     tbs_bits, _u8(tb_out), ctypes.byref(w))
        return bool(r), bool(w.value)

    # Stimulus (transmit side) lives in tools/stimulus; kept here as thin forwards so that tests read naturally.
    def encode(self, bg, Z, msg_bits):
        from tools import stimulus

        return stimulus.encode(bg, Z, msg_bits)

    def rate_match(self, bg, Z, cb_bits, rv, Qm, Nref, nof_filler, E):
        from tools import stimulus

        return stimulus.rate_match(bg, Z, cb_bits, rv, Qm, Nref, nof_filler, E)


def _demod_call(lib, prefix, symbols, noise_vars, mod):
    symbols = np.ascontiguousarray(symbols, np.float32).reshape(-1, 2)
    noise_vars = np.ascontiguousarray(noise_vars, np.float32)
    assert symbols.shape[0] == noise_vars.size
    out = np.zeros(noise_vars.size * (1 if mod == 0 else mod), np.int8)
    getattr(lib, prefix + "_demodulate_soft")(ctypes.c_void_p(out.ctypes.data), ctypes.c_void_p(symbols.ctypes.data),
                                              ctypes.c_void_p(noise_vars.ctypes.data), ctypes.c_uint(noise_vars.size),
                                              ctypes.c_int(mod))
    return out


def _descramble_call(lib, prefix, llrs, c_init, offset):
    llrs = np.ascontiguousarray(llrs, np.int8)
    out = np.zeros_like(llrs)
    getattr(lib, prefix + "_descramble")(ctypes.c_void_p(out.ctypes.data), ctypes.c_void_p(llrs.ctypes.data),
                                         ctypes.c_uint(llrs.size), ctypes.c_uint(c_init), ctypes.c_uint(offset))
    return out


def _prg_call(lib, prefix, n, c_init, offset):
    out = np.zeros(n, np.uint8)
    getattr(lib, prefix + "_prg_bits")(ctypes.c_void_p(out.ctypes.data), ctypes.c_uint(n), ctypes.c_uint(c_init),
                                       ctypes.c_uint(offset))
    return out


class DemodOracle:
    """LLR producers: the plain-C restatement in pdc_oracle_demod.c. mod: 0 pi/2-BPSK, 1 BPSK, 2, 4, 6, 8."""

    def __init__(self):
        build_oracle()
        self.lib = ctypes.CDLL(ORACLE_SO)

    def demodulate_soft(self, symbols, noise_vars, mod):
        return _demod_call(self.lib, "orc", symbols, noise_vars, mod)

    def descramble(self, llrs, c_init, offset=0):
        return _descramble_call(self.lib, "orc", llrs, c_init, offset)

    def prg_bits(self, n, c_init, offset=0):
        return _prg_call(self.lib, "orc", n, c_init, offset)

    def set_fused_mask(self, mask):
        ctypes.c_uint.in_dll(self.lib, "orc_demod_fused_mask").value = mask


class DemodReference:
    """The UNMODIFIED demodulation_mapper_impl / pseudo_random_generator_impl behind oracle/ref_shim_demod.cpp."""

    def __init__(self):
        self.lib = ctypes.CDLL(REF_SO)

    def demodulate_soft(self, symbols, noise_vars, mod):
        return _demod_call(self.lib, "ref", symbols, noise_vars, mod)

    def descramble(self, llrs, c_init, offset=0):
        return _descramble_call(self.lib, "ref", llrs, c_init, offset)

    def prg_bits(self, n, c_init, offset=0):
        return _prg_call(self.lib, "ref", n, c_init, offset)


class PdschEncReference:
    """The UNMODIFIED pdsch_encoder_impl (+ ldpc_segmenter_tx_impl, ldpc_encoder_avx2 | generic, ldpc_rate_matcher_impl) behind
    oracle/ref_shim_pdsch.cpp."""

    def __init__(self, generic=False):
        self.lib = ctypes.CDLL(REF_SO)
        self.generic = 1 if generic else 0

    def encode(self, tb_packed, tbs_bits, bg, rv, Qm, Nref, nof_layers, nof_ch_symbols):
        tb_packed = np.ascontiguousarray(tb_packed, np.uint8)
        cw = np.zeros(nof_ch_symbols * Qm, np.uint8)
        rc = self.lib.ref_pdsch_encode(ctypes.c_void_p(cw.ctypes.data), ctypes.c_void_p(tb_packed.ctypes.data),
                                       ctypes.c_uint(tbs_bits), ctypes.c_int(bg), ctypes.c_uint(rv), ctypes.c_uint(Qm),
                                       ctypes.c_uint(Nref), ctypes.c_uint(nof_layers), ctypes.c_uint(nof_ch_symbols),
                                       ctypes.c_int(self.generic))
        assert rc == 0
        return cw


class PdschEncOracle:
    """The TX-chain restatement (TB CRC, segmentation, CB CRC, LDPC encoding, rate matching, concatenation). It lives in
    tools/stimulus/pdc_stimulus.c because it doubles as the stimulus generator of the decoder tests; as the oracle of the PDSCH
    encoder it is pinned against the compiled pdsch_encoder_impl by tests/test_enc_oracle.py."""

    def encode(self, tb_packed, tbs_bits, bg, rv, Qm, Nref, nof_layers, nof_ch_symbols):
        from tools import stimulus as S

        tb_bits = np.unpackbits(np.ascontiguousarray(tb_packed, np.uint8))[:tbs_bits]
        _, bits, _ = S.tb_encode(bg, tbs_bits, tb_bits, rv, Qm, Nref, nof_layers, nof_ch_symbols * Qm)
        return bits


def ref_available():
    return os.path.exists(REF_SO)


class Reference:
    """The UNMODIFIED srsRAN classes behind oracle/ref_shim.cpp."""

    def __init__(self, flavour=FLAVOUR_AVX512):
        self.lib = ctypes.CDLL(REF_SO)
        self.lib.ref_create.restype = ctypes.c_void_p
        self.lib.ref_crc.restype = ctypes.c_uint32
        self.flavour = flavour
        self.h = ctypes.c_void_p(self.lib.ref_create(flavour))

    def __del__(self):
        try:
            self.lib.ref_destroy(self.h)
        except Exception:
            pass

    def crc(self, impl, crc_id, packed, nbits):
        packed = np.ascontiguousarray(packed, dtype=np.uint8)
        return int(self.lib.ref_crc(impl, crc_id, _u8(packed), int(nbits)))

    def rate_dematch(self, buf, llrs, new_data, bg, Z, rv, Qm, Nref, nof_filler):
        llrs = np.ascontiguousarray(llrs, dtype=np.int8)
        self.lib.ref_rate_dematch(self.h, _i8(buf), buf.size, _i8(llrs), llrs.size, int(new_data), bg, Z, rv, Qm, Nref,
                                  nof_filler)

    def decode(self, bg, Z, llrs, nof_filler=0, max_iter=6, crc_id=CRC_NONE, force_decoding=False, nof_crc_bits=None):
        llrs = np.ascontiguousarray(llrs, dtype=np.int8)
        K = kb_of(bg) * Z
        out = np.zeros((K + 7) // 8, np.uint8)
        if nof_crc_bits is None:
            nof_crc_bits = 16 if crc_id in (CRC_NONE, CRC16) else 24
        it = self.lib.ref_ldpc_decode(self.h, bg, Z, nof_filler, nof_crc_bits, max_iter, crc_id, int(force_decoding),
                                      _i8(llrs), llrs.size, _u8(out))
        return it, out

    def encode(self, bg, Z, msg_bits):
        packed = np.packbits(np.ascontiguousarray(msg_bits, dtype=np.uint8))
        cb = np.zeros(nshort_of(bg) * Z, np.uint8)
        self.lib.ref_ldpc_encode(self.h, bg, Z, _u8(packed), _u8(cb))
        return cb

    def encode_rm(self, bg, Z, msg_bits, rv, Qm, Nref, nof_filler, E):
        packed = np.packbits(np.ascontiguousarray(msg_bits, dtype=np.uint8))
        out = np.zeros(E, np.uint8)
        self.lib.ref_ldpc_encode_rm(self.h, bg, Z, _u8(packed), rv, Qm, Nref, nof_filler, E, _u8(out))
        return out

    def segment_rx(self, bg, rv, Qm, Nref, nof_layers, nof_llrs, tbs_bits):
        arrs = [np.zeros(152, np.uint32) for _ in range(6)]
        n = self.lib.ref_segment_rx(bg, rv, Qm, Nref, nof_layers, nof_llrs, tbs_bits,
                                    *[a.ctypes.data_as(ctypes.c_void_p) for a in arrs])
        keys = ("Z", "full_length", "rm_length", "nof_filler", "cw_offset", "nof_crc_bits")
        return [{k: int(a[i]) for k, a in zip(keys, arrs)} for i in range(n)]


REF_CB_DTYPE = np.dtype([("bg", np.int32), ("Z", np.int32), ("Qm", np.int32), ("rv", np.int32), ("crc_id", np.int32),
                         ("max_iter", np.int32), ("new_data", np.int32), ("early_stop", np.int32), ("E", np.uint32),
                         ("Nref", np.uint32), ("nof_filler", np.uint32), ("nof_crc_bits", np.uint32),
                         ("llr_offset", np.uint64), ("soft_offset", np.uint64), ("out_offset", np.uint64),
                         ("decode_only", np.int32), ("pad", np.int32)])


def ref_cpu_flags():
    """Bit 0 AVX2, bit 1 AVX512F+BW, bit 2 AVX512VBMI, bit 3 PCLMUL on THIS host (oracle/_ref must exist)."""
    lib = ctypes.CDLL(REF_SO)
    return int(lib.ref_cpu_flags())


def ref_best_flavours():
    """(decoder flavour, dematcher flavour) the reference's "auto" factories would pick on this host
    (channel_coding_factories.cpp:114-137, 284-320)."""
    f = ref_cpu_flags()
    dec = FLAVOUR_AVX512 if f & 2 else (FLAVOUR_AVX2 if f & 1 else FLAVOUR_GENERIC)
    dem = FLAVOUR_AVX512 if (f & 2 and f & 4) else (FLAVOUR_AVX2 if f & 1 else FLAVOUR_GENERIC)
    return dec, dem


def ref_chain_run(dec_flavour, dem_flavour, nthreads, descs, llrs, soft, bits, reps=1):
    """Runs the UNMODIFIED reference pusch_codeblock_decoder chain over descs (REF_CB_DTYPE) on nthreads host threads.
    Returns (seconds, iterations[int32])."""
    assert descs.dtype == REF_CB_DTYPE and descs.flags.c_contiguous
    lib = ctypes.CDLL(REF_SO)
    lib.ref_chain_run.restype = ctypes.c_double
    iters = np.zeros(descs.size, np.int32)
    vp = ctypes.c_void_p
    sec = lib.ref_chain_run(int(dec_flavour), int(dem_flavour), int(nthreads), int(descs.size), vp(descs.ctypes.data),
                            vp(llrs.ctypes.data), vp(soft.ctypes.data), vp(bits.ctypes.data), vp(iters.ctypes.data),
                            int(reps))
    return float(sec), iters


LIFTING_SIZES = sorted(a * 2 ** j for a in (2, 3, 5, 7, 9, 11, 13, 15) for j in range(8) if a * 2 ** j <= 384)
