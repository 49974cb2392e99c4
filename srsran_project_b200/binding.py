"""ctypes binding of include/pusch_dec_cuda.h (test / benchmark glue; the product is the shared library)."""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

PDC_CRC_NONE, PDC_CRC16, PDC_CRC24A, PDC_CRC24B = 0, 1, 2, 3
PDC_F_NEW_DATA = 0x01
PDC_F_EARLY_STOP = 0x02
PDC_F_FORCE_DECODING = 0x04
PDC_F_GENERIC = 0x08
PDC_F_DECODE_ONLY = 0x10
PDC_F_DEMATCH_ONLY = 0x20
PDC_F_DECODED = 0x40

PDC_OK, PDC_ERR_INVALID_ARG, PDC_ERR_NO_DEVICE, PDC_ERR_CUDA, PDC_ERR_CAPACITY, PDC_ERR_NOT_READY, PDC_ERR_BAD_TICKET = range(7)
_STATUS_NAMES = ["PDC_OK", "PDC_ERR_INVALID_ARG", "PDC_ERR_NO_DEVICE", "PDC_ERR_CUDA", "PDC_ERR_CAPACITY",
                 "PDC_ERR_NOT_READY", "PDC_ERR_BAD_TICKET"]


class CbDesc(ctypes.Structure):
    """pdc_cb_desc."""
    _fields_ = [("bg", ctypes.c_uint8), ("Qm", ctypes.c_uint8), ("rv", ctypes.c_uint8), ("crc_type", ctypes.c_uint8),
                ("max_iter", ctypes.c_uint8), ("flags", ctypes.c_uint8), ("Z", ctypes.c_uint16),
                ("E", ctypes.c_uint32), ("Nref", ctypes.c_uint32), ("nof_filler", ctypes.c_uint32),
                ("harq_slot", ctypes.c_int32), ("llr_offset", ctypes.c_uint64), ("soft_offset", ctypes.c_uint64),
                ("out_offset", ctypes.c_uint64)]


class CbResult(ctypes.Structure):
    """pdc_cb_result."""
    _fields_ = [("iterations", ctypes.c_int32), ("crc_pass", ctypes.c_uint8), ("status", ctypes.c_uint8),
                ("nof_layers", ctypes.c_uint16)]


class PdcConfig(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int), ("max_batch_cbs", ctypes.c_uint32), ("max_batch_llrs", ctypes.c_uint64),
                ("harq_slots", ctypes.c_uint32), ("nof_tickets", ctypes.c_uint32)]


RESULT_DTYPE = np.dtype([("iterations", np.int32), ("crc_pass", np.uint8), ("status", np.uint8),
                         ("nof_layers", np.uint16)])
DESC_DTYPE = np.dtype([("bg", np.uint8), ("Qm", np.uint8), ("rv", np.uint8), ("crc_type", np.uint8),
                       ("max_iter", np.uint8), ("flags", np.uint8), ("Z", np.uint16), ("E", np.uint32),
                       ("Nref", np.uint32), ("nof_filler", np.uint32), ("harq_slot", np.int32),
                       ("llr_offset", np.uint64), ("soft_offset", np.uint64), ("out_offset", np.uint64)])
TB_DESC_DTYPE = np.dtype([("first_cb", np.uint32), ("nof_cbs", np.uint32), ("tbs_bits", np.uint32), ("reserved", np.uint32),
                          ("tb_offset", np.uint64)])  # pdc_tb_desc
TB_RESULT_DTYPE = np.dtype([("tb_crc_ok", np.uint8), ("written", np.uint8), ("nof_cbs_ok", np.uint16)])  # pdc_tb_result
DEMOD_SEG_DTYPE = np.dtype([("sym_offset", np.uint64), ("llr_offset", np.uint64), ("nof_symbols", np.uint32),
                            ("c_init", np.uint32), ("scramble_offset", np.uint32), ("mod", np.uint8), ("flags", np.uint8),
                            ("reserved", np.uint16)])  # pdc_demod_seg
PDSCH_TB_DTYPE = np.dtype([("bg", np.uint8), ("Qm", np.uint8), ("rv", np.uint8), ("nof_layers", np.uint8),
                           ("tbs_bits", np.uint32), ("nof_ch_symbols", np.uint32), ("Nref", np.uint32),
                           ("tb_offset", np.uint64), ("cw_offset", np.uint64), ("cw_length_a", np.uint32),
                           ("cw_length_b", np.uint32), ("nof_short_segments", np.uint32),
                           ("reserved", np.uint32)])  # pdc_pdsch_tb_desc
ULSCH_DEMUX_DTYPE = np.dtype([("Qm", np.uint8), ("nof_layers", np.uint8), ("start_symbol_index", np.uint8),
                              ("nof_symbols", np.uint8), ("dmrs_type", np.uint8), ("nof_cdm_groups_without_data", np.uint8),
                              ("dmrs_symbol_mask", np.uint16), ("nof_prb", np.uint16), ("flags", np.uint8), ("reserved", np.uint8),
                              ("nof_harq_ack_rvd", np.uint32), ("nof_harq_ack_bits", np.uint32),
                              ("nof_enc_harq_ack_bits", np.uint32), ("nof_csi_part1_bits", np.uint32),
                              ("nof_enc_csi_part1_bits", np.uint32), ("nof_csi_part2_bits", np.uint32),
                              ("nof_enc_csi_part2_bits", np.uint32), ("c_init", np.uint32), ("reserved2", np.uint32),
                              ("in_offset", np.uint64), ("sch_offset", np.uint64), ("harq_ack_offset", np.uint64),
                              ("csi_part1_offset", np.uint64), ("csi_part2_offset", np.uint64)])  # pdc_ulsch_demux_desc
PDC_DEMUX_F_CSI2_AFTER_CSI1 = 0x01
PDC_MOD_PI2_BPSK, PDC_MOD_BPSK, PDC_MOD_QPSK, PDC_MOD_QAM16, PDC_MOD_QAM64, PDC_MOD_QAM256 = 0, 1, 2, 4, 6, 8
PDC_DEMOD_F_DESCRAMBLE = 0x01
PDC_STATIC_MAX_CBS = 1024
assert DESC_DTYPE.itemsize == ctypes.sizeof(CbDesc) == 48
assert TB_DESC_DTYPE.itemsize == 24 and TB_RESULT_DTYPE.itemsize == 4
assert RESULT_DTYPE.itemsize == ctypes.sizeof(CbResult) == 8


class PdcError(RuntimeError):
    def __init__(self, status, msg=""):
        self.status = status
        name = _STATUS_NAMES[status] if 0 <= status < len(_STATUS_NAMES) else str(status)
        super().__init__(f"{name}: {msg}" if msg else name)


def library_path():
    """PDC_LIBRARY selects another build of the SAME library (kernel A/B runs on one GPU box); never a different backend."""
    return os.environ.get("PDC_LIBRARY") or os.path.join(HERE, "libpuschdec_cuda.so")


_lib = None


def load_library():
    """Loads libpuschdec_cuda.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise PdcError(PDC_ERR_NO_DEVICE, f"{path} not built: run `make -C srsran_project_b200/csrc`")
    lib = ctypes.CDLL(path)
    vp, u8p, i8p = ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint8), ctypes.POINTER(ctypes.c_int8)
    lib.pdc_create.argtypes = [ctypes.POINTER(PdcConfig), ctypes.POINTER(vp)]
    lib.pdc_destroy.argtypes = [vp]
    lib.pdc_destroy.restype = None
    lib.pdc_last_error.argtypes = [vp]
    lib.pdc_last_error.restype = ctypes.c_char_p
    lib.pdc_set_zero_copy.argtypes = [vp, ctypes.c_int]
    lib.pdc_host_alloc.argtypes = [ctypes.c_size_t]
    lib.pdc_host_alloc.restype = vp
    lib.pdc_host_free.argtypes = [vp]
    lib.pdc_host_free.restype = None
    lib.pdc_stage_open.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(vp)]
    lib.pdc_stage_close.argtypes = [vp]
    lib.pdc_stage_close.restype = None
    lib.pdc_stage_host.argtypes = [vp]
    lib.pdc_stage_host.restype = vp
    lib.pdc_stage_push.argtypes = [vp, ctypes.c_size_t, ctypes.c_size_t]
    lib.pdc_submit_staged.argtypes = [vp, vp, vp, ctypes.c_uint32, vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int64)]
    lib.pdc_submit.argtypes = [vp, vp, ctypes.c_uint32, vp, ctypes.c_size_t, vp, ctypes.c_size_t,
                               ctypes.POINTER(ctypes.c_int64)]
    lib.pdc_poll.argtypes = [vp, ctypes.c_int64]
    lib.pdc_fetch.argtypes = [vp, ctypes.c_int64, vp, ctypes.c_size_t, vp, vp, ctypes.c_size_t]
    lib.pdc_decode_device.argtypes = [vp, vp, ctypes.c_uint32, vp, vp, vp, vp, vp]
    lib.pdc_submit_tbs.argtypes = [vp, vp, ctypes.c_uint32, vp, ctypes.c_uint32, vp, ctypes.c_size_t, vp, ctypes.c_size_t,
                                   ctypes.POINTER(ctypes.c_int64)]
    lib.pdc_fetch_tbs.argtypes = [vp, ctypes.c_int64, vp, ctypes.c_size_t, vp, vp, vp, ctypes.c_size_t]
    lib.pdc_decode_device_tbs.argtypes = [vp, vp, ctypes.c_uint32, vp, ctypes.c_uint32, vp, vp, vp, vp, vp, vp, vp]
    lib.pdc_decode_device_tbs_static.argtypes = lib.pdc_decode_device_tbs.argtypes
    lib.pdc_demodulate_device.argtypes = [vp, vp, ctypes.c_uint32, vp, vp, vp, vp]
    lib.pdc_demodulate.argtypes = [vp, vp, ctypes.c_uint32, vp, vp, ctypes.c_size_t, vp, ctypes.c_size_t]
    lib.pdc_ulsch_demux_lengths.argtypes = [vp, vp]
    lib.pdc_ulsch_demultiplex_device.argtypes = [vp, vp, ctypes.c_uint32, vp, vp, vp]
    lib.pdc_ulsch_demultiplex.argtypes = [vp, vp, ctypes.c_uint32, vp, ctypes.c_size_t, vp, ctypes.c_size_t]
    lib.pdc_pdsch_encode.argtypes = [vp, vp, ctypes.c_uint32, vp, ctypes.c_size_t, vp, ctypes.c_size_t]
    lib.pdc_pdsch_encode_device.argtypes = [vp, vp, ctypes.c_uint32, vp, vp, vp]
    u32p = ctypes.POINTER(ctypes.c_uint32)
    lib.pdc_pdsch_segmentation.argtypes = [ctypes.c_int, ctypes.c_uint32, u32p, u32p, u32p, u32p]
    lib.pdc_harq_free.argtypes = [vp, ctypes.c_int32]
    lib.pdc_harq_release.argtypes = [vp, ctypes.c_int32]
    lib.pdc_harq_read.argtypes = [vp, ctypes.c_int32, vp, ctypes.c_size_t]
    lib.pdc_harq_write.argtypes = [vp, ctypes.c_int32, vp, ctypes.c_size_t]
    lib.pdc_kernel_launches.argtypes = [vp]
    lib.pdc_kernel_launches.restype = ctypes.c_uint64
    lib.pdc_probe_issue_rate.argtypes = [vp, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    lib.pdc_cb_message_bits.restype = ctypes.c_uint32
    lib.pdc_cb_buffer_len.restype = ctypes.c_uint32
    lib.pdc_cb_max_layers.argtypes = [vp]
    lib.pdc_cb_max_layers.restype = ctypes.c_uint32
    _lib = lib
    return lib


EXPORTED_SYMBOLS = ["pdc_create", "pdc_destroy", "pdc_last_error", "pdc_abi_version", "pdc_set_zero_copy", "pdc_host_alloc",
                    "pdc_host_free", "pdc_stage_open", "pdc_stage_close", "pdc_stage_host", "pdc_stage_push", "pdc_submit_staged", "pdc_submit", "pdc_poll",
                    "pdc_fetch", "pdc_decode_device", "pdc_submit_tbs", "pdc_fetch_tbs", "pdc_decode_device_tbs",
                    "pdc_decode_device_tbs_static",
                    "pdc_demodulate_device", "pdc_demodulate", "pdc_ulsch_demux_lengths",
                    "pdc_ulsch_demultiplex_device", "pdc_ulsch_demultiplex", "pdc_pdsch_encode", "pdc_pdsch_encode_device",
                    "pdc_pdsch_segmentation", "pdc_harq_free", "pdc_harq_release", "pdc_harq_read", "pdc_harq_write",
                    "pdc_kernel_launches", "pdc_probe_issue_rate", "pdc_cb_message_bits", "pdc_cb_buffer_len", "pdc_cb_max_layers"]


def pdsch_segmentation(bg, tbs_bits):
    """(C, Z, K', fillers) the device encoder will use for a transport block (pdc_pdsch_segmentation); host-only."""
    v = [ctypes.c_uint32() for _ in range(4)]
    st = load_library().pdc_pdsch_segmentation(bg, tbs_bits, *[ctypes.byref(x) for x in v])
    if st != 0:
        raise PdcError(st)
    return tuple(x.value for x in v)


def ulsch_demux_lengths(desc):
    """(input, UL-SCH, HARQ-ACK, CSI Part 1, CSI Part 2) soft bits for one pdc_ulsch_demux_desc record; host-only."""
    lens = np.zeros(5, np.uint32)
    st = load_library().pdc_ulsch_demux_lengths(desc.ctypes.data, lens.ctypes.data)
    if st != 0:
        raise PdcError(st)
    return tuple(int(v) for v in lens)


def make_descs(n):
    """Zero-initialised numpy array of pdc_cb_desc with harq_slot = -1."""
    d = np.zeros(n, DESC_DTYPE)
    d["harq_slot"] = -1
    return d


class PuschDecCuda:
    """One accelerator context (one GPU)."""

    def __init__(self, device=0, max_batch_cbs=1024, max_batch_llrs=64 << 20, harq_slots=0, nof_tickets=2):
        self.lib = load_library()
        cfg = PdcConfig(device, max_batch_cbs, max_batch_llrs, harq_slots, nof_tickets)
        h = ctypes.c_void_p()
        st = self.lib.pdc_create(ctypes.byref(cfg), ctypes.byref(h))
        if st != PDC_OK:
            raise PdcError(st, self.lib.pdc_last_error(None).decode())
        self.h = h
        self.device = device

    def set_zero_copy(self, enable=True):
        """pdc_set_zero_copy: page-locked LLR sources are DMA-read in place; keep them untouched until the ticket completes."""
        self._check(self.lib.pdc_set_zero_copy(self.h, 1 if enable else 0))

    def close(self):
        if getattr(self, "h", None):
            self.lib.pdc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st):
        if st != PDC_OK:
            raise PdcError(st, self.lib.pdc_last_error(self.h).decode())

    # --- host-buffer path ------------------------------------------------------------------------------------------
    def submit(self, descs, llrs, soft_in=None):
        assert descs.dtype == DESC_DTYPE and descs.flags.c_contiguous
        assert llrs.dtype == np.int8 and llrs.flags.c_contiguous
        t = ctypes.c_int64()
        sp, sn = (None, 0) if soft_in is None else (soft_in.ctypes.data, soft_in.nbytes)
        self._check(self.lib.pdc_submit(self.h, descs.ctypes.data, descs.size, llrs.ctypes.data, llrs.nbytes, sp, sn,
                                        ctypes.byref(t)))
        return t.value

    def poll(self, ticket):
        st = self.lib.pdc_poll(self.h, ticket)
        if st == PDC_ERR_NOT_READY:
            return False
        self._check(st)
        return True

    def fetch(self, ticket, n, bits, soft_out=None):
        res = np.zeros(n, RESULT_DTYPE)
        sp, sn = (None, 0) if soft_out is None else (soft_out.ctypes.data, soft_out.nbytes)
        self._check(self.lib.pdc_fetch(self.h, ticket, bits.ctypes.data, bits.nbytes, res.ctypes.data, sp, sn))
        return res

    def run(self, descs, llrs, bits_bytes, soft=None):
        """submit + fetch. ``soft`` (int8 array) is updated in place when given."""
        bits = np.zeros(bits_bytes, np.uint8)
        t = self.submit(descs, llrs, soft)
        res = self.fetch(t, descs.size, bits, soft)
        return res, bits

    # --- transport-block path ---------------------------------------------------------------------------------------
    def submit_tbs(self, descs, tbs, llrs, soft_in=None):
        assert descs.dtype == DESC_DTYPE and descs.flags.c_contiguous
        assert tbs.dtype == TB_DESC_DTYPE and tbs.flags.c_contiguous
        assert llrs.dtype == np.int8 and llrs.flags.c_contiguous
        t = ctypes.c_int64()
        sp, sn = (None, 0) if soft_in is None else (soft_in.ctypes.data, soft_in.nbytes)
        self._check(self.lib.pdc_submit_tbs(self.h, descs.ctypes.data, descs.size, tbs.ctypes.data, tbs.size,
                                            llrs.ctypes.data if llrs.size else None, llrs.nbytes, sp, sn, ctypes.byref(t)))
        return t.value

    def fetch_tbs(self, ticket, n, ntb, tb_bytes, soft_out=None):
        """-> (TB results, CB results); payloads of the written TBs land in ``tb_bytes`` (uint8 array)."""
        tb_res, cb_res = np.zeros(ntb, TB_RESULT_DTYPE), np.zeros(n, RESULT_DTYPE)
        sp, sn = (None, 0) if soft_out is None else (soft_out.ctypes.data, soft_out.nbytes)
        self._check(self.lib.pdc_fetch_tbs(self.h, ticket, tb_bytes.ctypes.data, tb_bytes.nbytes, tb_res.ctypes.data,
                                           cb_res.ctypes.data, sp, sn))
        return tb_res, cb_res

    def decode_device_tbs(self, descs, tbs, d_llrs_ptr, d_soft_ptr, d_bits_ptr, d_res_ptr, d_tb_bytes_ptr, d_tb_res_ptr,
                          stream_ptr=None):
        self._check(self.lib.pdc_decode_device_tbs(self.h, descs.ctypes.data, descs.size, tbs.ctypes.data, tbs.size,
                                                   d_llrs_ptr, d_soft_ptr, d_bits_ptr, d_res_ptr, d_tb_bytes_ptr,
                                                   d_tb_res_ptr, stream_ptr))

    def decode_device_tbs_static(self, descs, tbs, d_llrs_ptr, d_soft_ptr, d_bits_ptr, d_res_ptr, d_tb_bytes_ptr,
                                 d_tb_res_ptr, stream_ptr=None):
        """pdc_decode_device_tbs_static: the allocation-free variant (scratch in a static device array, <= 1024 CBs)."""
        self._check(self.lib.pdc_decode_device_tbs_static(self.h, descs.ctypes.data, descs.size, tbs.ctypes.data, tbs.size,
                                                          d_llrs_ptr, d_soft_ptr, d_bits_ptr, d_res_ptr, d_tb_bytes_ptr,
                                                          d_tb_res_ptr, stream_ptr))

    # --- device path -----------------------------------------------------------------------------------------------
    def decode_device(self, descs, d_llrs_ptr, d_soft_ptr, d_bits_ptr, d_res_ptr, stream_ptr=None):
        self._check(self.lib.pdc_decode_device(self.h, descs.ctypes.data, descs.size, d_llrs_ptr, d_soft_ptr,
                                               d_bits_ptr, d_res_ptr, stream_ptr))

    # --- LLR producers (soft demodulation + descrambling) ------------------------------------------------------------------
    def demodulate(self, segs, symbols, noise_vars, llr_bytes):
        """Host-buffer entry: symbols = float32 (n, 2) (re, im), noise_vars = float32 (n,). Returns the int8 LLR array."""
        symbols = np.ascontiguousarray(symbols, np.float32)
        noise_vars = np.ascontiguousarray(noise_vars, np.float32)
        llrs = np.zeros(llr_bytes, np.int8)
        self._check(self.lib.pdc_demodulate(self.h, segs.ctypes.data, segs.size, symbols.ctypes.data, noise_vars.ctypes.data,
                                            noise_vars.size, llrs.ctypes.data, llrs.size))
        return llrs

    def demodulate_device(self, segs, d_symbols_ptr, d_noise_ptr, d_llrs_ptr, stream_ptr=None):
        self._check(self.lib.pdc_demodulate_device(self.h, segs.ctypes.data, segs.size, d_symbols_ptr, d_noise_ptr,
                                                   d_llrs_ptr, stream_ptr))

    # --- UL-SCH demultiplexer -----------------------------------------------------------------------------------------------
    def ulsch_demultiplex(self, descs, llrs, out_len):
        """Host-buffer entry: llrs = descrambled codewords, returns the output buffer holding the four streams of each."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        out = np.zeros(out_len, np.int8)
        self._check(self.lib.pdc_ulsch_demultiplex(self.h, descs.ctypes.data, descs.size, llrs.ctypes.data, llrs.size,
                                                   out.ctypes.data, out.size))
        return out

    def ulsch_demultiplex_device(self, descs, d_in_ptr, d_out_ptr, stream_ptr=None):
        self._check(self.lib.pdc_ulsch_demultiplex_device(self.h, descs.ctypes.data, descs.size, d_in_ptr, d_out_ptr,
                                                          stream_ptr))

    # --- PDSCH transport-block encoder (downlink mirror) -------------------------------------------------------------------
    def pdsch_encode(self, tbs, tb_bytes, cw_len):
        """Host-buffer entry: tb_bytes = packed transport blocks, returns the codewords (one bit per byte)."""
        tb_bytes = np.ascontiguousarray(tb_bytes, np.uint8)
        cw = np.zeros(cw_len, np.uint8)
        self._check(self.lib.pdc_pdsch_encode(self.h, tbs.ctypes.data, tbs.size, tb_bytes.ctypes.data, tb_bytes.size,
                                              cw.ctypes.data, cw.size))
        return cw

    def pdsch_encode_device(self, tbs, d_tb_ptr, d_cw_ptr, stream_ptr=None):
        self._check(self.lib.pdc_pdsch_encode_device(self.h, tbs.ctypes.data, tbs.size, d_tb_ptr, d_cw_ptr, stream_ptr))

    # --- HARQ arena ------------------------------------------------------------------------------------------------
    def harq_free(self, slot):
        self._check(self.lib.pdc_harq_free(self.h, slot))

    def harq_release(self, slot):
        """The slot reads as zeros from now on (host-side bookkeeping, cleared lazily): pdc_harq_release."""
        self._check(self.lib.pdc_harq_release(self.h, slot))

    def harq_read(self, slot, n):
        out = np.zeros(n, np.int8)
        self._check(self.lib.pdc_harq_read(self.h, slot, out.ctypes.data, n))
        return out

    def harq_write(self, slot, data):
        data = np.ascontiguousarray(data, np.int8)
        self._check(self.lib.pdc_harq_write(self.h, slot, data.ctypes.data, data.size))

    def probe_issue_rate(self, mode):
        """Thread-level instructions per second of the probe `mode` (see pdc_probe_issue_rate)."""
        v = ctypes.c_double()
        self._check(self.lib.pdc_probe_issue_rate(self.h, mode, ctypes.byref(v)))
        return v.value

    def kernel_launches(self):
        return int(self.lib.pdc_kernel_launches(self.h))
