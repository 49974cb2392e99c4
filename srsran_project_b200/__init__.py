"""B200-native PUSCH code-block decode chain (rate dematch + HARQ combine -> layered min-sum LDPC -> CB CRC).

The product is the C-ABI shared library ``libpuschdec_cuda.so`` (include/pusch_dec_cuda.h) built from
``srsran_project_b200/csrc``; this package only holds the thin ctypes binding used by the tests and by bench.py, and
the host-side helpers (segmentation metadata) that mirror the reference's host logic.  There is no CPU fallback: if
the library is missing, or no CUDA device is present, creating a context raises.
"""
from .binding import (PDC_CRC16, PDC_CRC24A, PDC_CRC24B, PDC_CRC_NONE, PDC_F_DECODE_ONLY, PDC_F_DEMATCH_ONLY,  # noqa: F401
                      PDC_F_EARLY_STOP, PDC_F_FORCE_DECODING, PDC_F_GENERIC, PDC_F_NEW_DATA, CbDesc, CbResult,
                      PdcError, PuschDecCuda, library_path, load_library)
