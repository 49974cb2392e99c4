"""Stimulus generator for tests and bench.py (transmit side of TS 38.212 + an LLR-domain AWGN channel).

Input generation only: nothing here is on the product path and nothing here is the oracle. The C source is
tools/stimulus/pdc_stimulus.c; it is compiled on first use into tools/stimulus/libpdc_stimulus.so (gcc, seconds).
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libpdc_stimulus.so")
SRC = os.path.join(HERE, "pdc_stimulus.c")

_u8p = ctypes.POINTER(ctypes.c_uint8)
_i8p = ctypes.POINTER(ctypes.c_int8)


class Seg(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in ("bg", "Z", "C", "K", "Kp", "F", "tb_crc_bits", "cb_crc_bits", "B")]


def build():
    tables = os.path.join(HERE, "..", "..", "srsran_project_b200", "csrc", "ldpc_bg_tables.h")
    if (not os.path.exists(SO) or os.path.getmtime(SRC) > os.path.getmtime(SO)
            or os.path.getmtime(tables) > os.path.getmtime(SO)):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-std=c11", "-o", SO, SRC, "-lm"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(SO)
        _lib.stim_crc.restype = ctypes.c_uint32
        _lib.stim_awgn_llrs.argtypes = [_u8p, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_uint64, _i8p]
        _lib.stim_awgn_llrs.restype = None
    return _lib


def _u8(a):
    return a.ctypes.data_as(_u8p)


def kb_of(bg):
    return 22 if bg == 1 else 10


def nshort_of(bg):
    return 66 if bg == 1 else 50


def crc(crc_id, bits):
    bits = np.ascontiguousarray(bits, np.uint8)
    return int(lib().stim_crc(crc_id, _u8(bits), bits.size))


def encode(bg, Z, msg_bits):
    msg_bits = np.ascontiguousarray(msg_bits, np.uint8)
    cb = np.zeros(nshort_of(bg) * Z, np.uint8)
    assert lib().stim_ldpc_encode(bg, Z, _u8(msg_bits), _u8(cb)) == 0
    return cb


def rate_match(bg, Z, cb_bits, rv, Qm, Nref, nof_filler, E):
    cb_bits = np.ascontiguousarray(cb_bits, np.uint8)
    out = np.zeros(E, np.uint8)
    assert lib().stim_rate_match(bg, Z, _u8(cb_bits), rv, Qm, Nref, nof_filler, E, _u8(out)) == 0
    return out


def segmentation(bg, tbs):
    s = Seg()
    assert lib().stim_segmentation(bg, tbs, ctypes.byref(s)) == 0
    return {n: getattr(s, n) for n, _ in Seg._fields_}


def rm_lengths(C, G, nof_layers, Qm):
    return [int(lib().stim_rm_length(r, C, G, nof_layers, Qm)) for r in range(C)]


def tb_encode(bg, tbs, tb_bits, rv, Qm, Nref, nof_layers, G):
    """Returns (seg dict, rate-matched bits of all CBs back to back [G], messages [C, K])."""
    seg = segmentation(bg, tbs)
    tb_bits = np.ascontiguousarray(tb_bits, np.uint8)
    assert tb_bits.size == tbs
    out = np.zeros(G, np.uint8)
    msgs = np.zeros((seg["C"], seg["K"]), np.uint8)
    c = lib().stim_tb_encode(bg, tbs, _u8(tb_bits), rv, Qm, Nref, nof_layers, G, _u8(out), _u8(msgs))
    assert c == seg["C"]
    return seg, out, msgs


def awgn_llrs(bits, amp, sigma, seed):
    bits = np.ascontiguousarray(bits, np.uint8)
    out = np.zeros(bits.size, np.int8)
    lib().stim_awgn_llrs(_u8(bits), bits.size, float(amp), float(sigma), int(seed), out.ctypes.data_as(_i8p))
    return out


# BASELINE.json config 2: 100 MHz, 273 PRB, 14 symbols (1 DMRS symbol with 2 CDM groups without data), 4 layers, 256QAM,
# R = 948/1024, rv 0 (SURVEY.md section 8 size table).
CONFIG2 = dict(bg=1, tbs=1277992, Qm=8, nof_layers=4, G=1362816, rv=0)
