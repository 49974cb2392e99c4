"""Drop-in boundary tests: the UNMODIFIED reference transport-block decoders (pusch_decoder_impl, pusch_decoder_hw_impl,
compiled from the reference sources into tests/adapters/_build/libadapter_harness.so) driving the CUDA adapters of
srsran_project_b200/adapters/, compared with the same decoders driving the reference's own AVX512 / generic operators; and the CUDA srsran::pusch_decoder
(code blocks and transport-block assembly + CRC24A on the device) in their place.

Mirrors tests/unittests/phy/upper/channel_processors/pusch/pusch_decoder_vectortest.cpp:286-403 (TB-level HARQ test: feed
the codeword in two chunks, check tb_crc_ok, the TB bytes, the per-CB statistics and the iteration bounds).
"""
import ctypes
import os

import numpy as np
import pytest

from tools import stimulus as S

HERE = os.path.dirname(os.path.abspath(__file__))
HARNESS = os.path.join(HERE, "adapters", "_build", "libadapter_harness.so")
ADAPTERS = os.path.join(HERE, "..", "srsran_project_b200", "libsrsran_cuda_adapters.so")

MODE_REF, MODE_CUDA_OPERATORS, MODE_CUDA_HAL, MODE_CUDA_TB, MODE_CUDA_HAL_METRICS = 0, 1, 2, 3, 4


def _lib():
    if not os.path.exists(HARNESS):
        pytest.skip("tests/adapters/_build not built (needs /root/reference at build time)")
    lib = ctypes.CDLL(HARNESS)
    lib.adt_create.restype = ctypes.c_void_p
    lib.adt_create.argtypes = [ctypes.c_int] * 5 + [ctypes.c_uint] * 2
    lib.adt_destroy.argtypes = [ctypes.c_void_p]
    lib.adt_decode_tb.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_int, ctypes.c_uint, ctypes.c_int] + \
        [ctypes.c_uint] * 4 + [ctypes.c_int, ctypes.c_uint, ctypes.c_void_p, ctypes.c_uint, ctypes.c_void_p,
                               ctypes.c_void_p]
    return lib


def best_ref_flavours():
    from oracle import pyoracle as O

    return O.ref_best_flavours()


class Decoder:
    def __init__(self, mode, cuda_generic=False, external_harq=True, nof_buffers=4, nof_codeblocks=4 * 152,
                 ref_generic=False):
        self.lib = _lib()
        dec, dem = (0, 0) if ref_generic else best_ref_flavours()
        self.h = self.lib.adt_create(mode, dec, dem, int(cuda_generic), int(external_harq), nof_buffers, nof_codeblocks)
        assert self.h

    def close(self):
        if self.h:
            self.lib.adt_destroy(self.h)
            self.h = None

    def decode(self, harq_id, new_data, rv, bg, Qm, nof_layers, Nref, max_iter, early_stop, tbs, llrs):
        llrs = np.ascontiguousarray(llrs, np.int8)
        tb = np.zeros(tbs // 8, np.uint8)
        stats = np.zeros(6, np.int64)
        rc = self.lib.adt_decode_tb(self.h, harq_id, int(new_data), rv, bg, Qm, nof_layers, Nref, max_iter,
                                    int(early_stop), tbs, llrs.ctypes.data, llrs.size, tb.ctypes.data, stats.ctypes.data)
        assert rc == 0, rc
        return tb, dict(tb_crc_ok=bool(stats[0]), nof_cbs=int(stats[1]), nof_obs=int(stats[2]), it_min=int(stats[3]),
                        it_max=int(stats[4]), it_mean_milli=int(stats[5]))


def make_tb(seed, bg, tbs, Qm, nof_layers, G, rvs, sigmas, Nref=0, amp=8.0):
    """Random TB -> reference TX chain restated in tools/stimulus -> one LLR vector per transmission."""
    rng = np.random.default_rng(seed)
    tb_bits = rng.integers(0, 2, tbs, dtype=np.uint8)
    txs = []
    for k, (rv, sigma) in enumerate(zip(rvs, sigmas)):
        _, bits, _ = S.tb_encode(bg, tbs, tb_bits, rv, Qm, Nref, nof_layers, G)
        txs.append(S.awgn_llrs(bits, amp, sigma, 1000 * seed + k))
    return np.packbits(tb_bits), txs


# (name, bg, tbs, Qm, layers, G, rvs, sigmas, Nref, max_iter, early_stop)
CASES = [
    ("cfg2_256qam_152cb", 1, 1277992, 8, 4, 1362816, (0,), (2.2,), 0, 6, True),
    ("cfg2_lbrm", 1, 1277992, 8, 4, 1362816, (0,), (2.0,), 12611, 6, True),
    ("bg1_two_cbs_harq", 1, 12296, 4, 2, 13600, (0, 2), (5.5, 5.5), 0, 6, True),
    ("bg1_single_cb_crc24a", 1, 6400, 6, 1, 9600, (0, 3), (4.5, 4.5), 0, 6, True),
    ("bg2_small_crc16", 2, 928, 2, 1, 7800, (0,), (7.0,), 0, 6, True),
    ("bg2_harq_fail_then_pass", 2, 3000, 2, 1, 4200, (0, 2, 3), (6.0, 6.0, 6.0), 0, 4, True),
    ("bg1_no_early_stop", 1, 30000, 8, 2, 35200, (0,), (2.0,), 0, 3, False),
    ("bg1_hopeless", 1, 20000, 4, 1, 23000, (0, 1), (30.0, 30.0), 0, 2, True),
]


def run_case(dec, case, harq_id=0):
    name, bg, tbs, Qm, layers, G, rvs, sigmas, Nref, max_iter, early = case
    want_tb, txs = make_tb(hash(name) % 1000, bg, tbs, Qm, layers, G, rvs, sigmas, Nref)
    out = []
    for k, (rv, llrs) in enumerate(zip(rvs, txs)):
        tb, st = dec.decode(harq_id, k == 0, rv, bg, Qm, layers, Nref, max_iter, early, tbs, llrs)
        out.append((tb, st))
        if st["tb_crc_ok"]:
            assert (tb == want_tb).all(), (name, k)
            break
    return out


def test_harness_reference_decoder_recovers_transport_blocks():
    """CPU-only sanity of the harness itself: the reference decoder over the reference operators (the oracle of the GPU
    tests below) decodes what the stimulus generator encoded."""
    dec = Decoder(MODE_REF)
    try:
        for case in CASES[2:5]:
            out = run_case(dec, case)
            assert out[-1][1]["tb_crc_ok"], case[0]
            assert out[-1][1]["nof_obs"] == out[-1][1]["nof_cbs"]
    finally:
        dec.close()


def test_adapter_library_exports():
    if not os.path.exists(ADAPTERS):
        pytest.skip("libsrsran_cuda_adapters.so not built (needs the reference headers at build time)")
    import subprocess

    syms = subprocess.run(["nm", "-D", "--defined-only", "-C", ADAPTERS], capture_output=True, text=True).stdout
    for s in ("srsran::create_ldpc_decoder_factory_cuda", "srsran::create_ldpc_rate_dematcher_factory_cuda",
              "srsran::hal::create_cuda_pusch_dec_acc_factory", "srsran::create_pusch_decoder_factory_cuda",
              "srsran::create_demodulation_mapper_factory_cuda", "srsran::create_pdsch_encoder_factory_cuda",
              "srsran::hal::create_cuda_pdsch_enc_acc_factory", "srsran::create_ulsch_demultiplex_factory_cuda"):
        assert s in syms, s


@pytest.mark.gpu
@pytest.mark.parametrize("mode,external", [(MODE_CUDA_HAL, True), (MODE_CUDA_HAL, False), (MODE_CUDA_OPERATORS, True),
                                           (MODE_CUDA_TB, True)])
@pytest.mark.parametrize("generic", [False, True])
def test_cuda_adapters_match_reference_decoder(mode, external, generic):
    """Same TB bytes, same tb_crc_ok, same per-CB iteration statistics as the reference decoder on every transmission."""
    ref = Decoder(MODE_REF, ref_generic=generic)
    dut = Decoder(mode, cuda_generic=generic, external_harq=external)
    try:
        cases = CASES if mode != MODE_CUDA_OPERATORS else CASES[2:]  # 152 synchronous single-CB calls per TB are slow: skip cfg2
        for i, case in enumerate(cases):
            a = run_case(ref, case, harq_id=i % 4)
            b = run_case(dut, case, harq_id=i % 4)
            assert len(a) == len(b), case[0]
            for k, ((tb_a, st_a), (tb_b, st_b)) in enumerate(zip(a, b)):
                assert st_a == st_b, (case[0], k, st_a, st_b)
                if st_a["tb_crc_ok"]:
                    assert (tb_a == tb_b).all(), (case[0], k)
    finally:
        ref.close()
        dut.close()


@pytest.mark.gpu
def test_hal_metric_decorator_reports_every_code_block():
    """The decorator that integration/srsran_cuda.patch adds to the reference (lib/phy/metrics, wired in upper_phy_factories.cpp)
    around the CUDA accelerator: same decoding results, one ldpc_decoder_metrics per decoded code block."""
    ref, dut = Decoder(MODE_REF), Decoder(MODE_CUDA_HAL_METRICS)
    try:
        nof_cbs = iters_ok = 0
        for i, case in enumerate(CASES[2:6]):
            a, b = run_case(ref, case, harq_id=i % 4), run_case(dut, case, harq_id=i % 4)
            assert [st for _, st in a] == [st for _, st in b], case[0]
            nof_cbs += sum(st["nof_obs"] for _, st in b)
        m = np.zeros(5, np.uint64)
        dut.lib.adt_metrics.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        dut.lib.adt_metrics(dut.h, m.ctypes.data)
        assert int(m[0]) == nof_cbs and int(m[3]) > 0 and int(m[4]) > 0 and int(m[1]) >= int(m[0]), (m, nof_cbs)
    finally:
        ref.close()
        dut.close()


@pytest.mark.gpu
def test_hal_accelerators_of_one_factory_share_gpu_submissions(monkeypatch):
    """Cross-instance batching (adapters/pusch_dec_batcher.h): six host threads, each with its own reference pusch_decoder_hw_impl,
    draw accelerators from one CUDA HAL factory and decode small transport blocks back to back. Every transport block comes out
    as transmitted, and the GPU saw fewer submissions than transport blocks (the objects' code blocks travelled together).
    The batcher is told to keep one submission in flight (batching.eager_inflight = 1: the setting for a loaded device; the
    default of 6 sends a lone transport block off at once while the device has room)."""
    monkeypatch.setenv("ADT_EAGER_INFLIGHT", "1")
    from tools import bench_hal

    if not bench_hal.available():
        pytest.skip("tests/adapters/_build not built (needs /root/reference at build time)")
    bg, tbs, Qm, layers, G = 1, 30000, 4, 2, 35200
    n_distinct = 5
    llrs, want = np.zeros((n_distinct, G), np.int8), np.zeros((n_distinct, tbs // 8), np.uint8)
    for t in range(n_distinct):
        want[t], (llrs[t],) = make_tb(300 + t, bg, tbs, Qm, layers, G, (0,), (2.5,))
    r = bench_hal.run_raw(llrs, want, tbs, bg, Qm, layers, 6, threads=6, tbs_per_thread=20, warmup=1, max_parts=8,
                          close_after_us=200)
    assert r["ok"] == r["attempted"] == 120, r
    assert r["batches"] < r["attempted"], r
    # one thread alone: nothing to coalesce with, every transport block is its own submission and still comes out right
    monkeypatch.delenv("ADT_EAGER_INFLIGHT")
    r1 = bench_hal.run_raw(llrs, want, tbs, bg, Qm, layers, 6, threads=1, tbs_per_thread=6, warmup=1)
    assert r1["ok"] == r1["attempted"] == 6 and r1["batches"] == 6, r1


@pytest.mark.gpu
def test_hal_harq_retransmissions_from_many_threads(monkeypatch):
    """HARQ through the drop-in boundary under concurrency: eight host threads, each with its own reference pusch_decoder_hw_impl
    over accelerators of one CUDA HAL factory (external HARQ: the soft bits live in the device arena, keyed by the absolute
    code-block identifier). First transmissions arrive at a noise level where most transport blocks fail; each failed one is
    sent again as redundancy version 2 and combined with what the accelerator kept. With the buffer limit the live stack
    configures (Nref < N: the arena bookkeeping of DESIGN.md 3.7 decides per code block which kernel it takes, slots are
    released only when a transport block passes). Every transport block must come out as transmitted in the end."""
    from tools import bench_hal
    from tools import stimulus as S

    if not bench_hal.available():
        pytest.skip("tests/adapters/_build not built (needs /root/reference at build time)")
    bg, tbs, Qm, layers, G = 1, 30000, 4, 2, 35200
    seg = S.segmentation(bg, tbs)
    nref = int(0.62 * 66 * seg["Z"])  # limits the circular buffer: rv 2 starts at floor(33 Ncb / N) Z
    monkeypatch.setenv("ADT_NREF", str(nref))
    n_distinct = 6
    tx1, tx2, want = np.zeros((n_distinct, G), np.int8), np.zeros((n_distinct, G), np.int8), np.zeros((n_distinct, tbs // 8), np.uint8)
    for t in range(n_distinct):
        want[t], (tx1[t], tx2[t]) = make_tb(500 + t, bg, tbs, Qm, layers, G, (0, 2), (5.5, 1.5), nref)
    r = bench_hal.run_raw(tx1, want, tbs, bg, Qm, layers, 6, threads=8, tbs_per_thread=12, warmup=1, retx=(tx2, 2))
    assert r["attempted"] == 96 and r["ok"] == r["attempted"], r
    assert r["retransmitted"] >= r["attempted"] // 2, r  # the stress is real: most first transmissions failed


@pytest.mark.gpu
def test_hal_factory_spreads_code_blocks_over_several_gpus():
    """cuda_hwacc_pusch_dec_factory_configuration::devices: code blocks are placed by absolute code-block identifier (GPU = id mod G)
    and every transport block still comes out as transmitted, first transmissions and HARQ-free repeats alike. Needs two GPUs."""
    import torch

    from tools import bench_hal

    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least two CUDA devices")
    if not bench_hal.available():
        pytest.skip("tests/adapters/_build not built (needs /root/reference at build time)")
    bg, tbs, Qm, layers, G = 1, 60000, 4, 2, 70400
    n_distinct = 4
    llrs, want = np.zeros((n_distinct, G), np.int8), np.zeros((n_distinct, tbs // 8), np.uint8)
    for t in range(n_distinct):
        want[t], (llrs[t],) = make_tb(400 + t, bg, tbs, Qm, layers, G, (0,), (2.5,))
    devices = tuple(range(min(torch.cuda.device_count(), 8)))
    r = bench_hal.run_raw(llrs, want, tbs, bg, Qm, layers, 6, threads=4, tbs_per_thread=12, warmup=1, devices=devices)
    assert r["ok"] == r["attempted"] == 48, r


@pytest.mark.gpu
def test_cuda_demodulation_mapper_adapter_reproduces_the_golden_calls():
    """srsran::demodulation_mapper implemented over pdc_demodulate (adapters/channel_modulation_factories_cuda.cpp), driven
    through the reference's interface, against the calls recorded from the compiled reference (tests/golden/demod_v1.npz)."""
    from tests.demod_cases import golden_calls

    lib = _lib()
    for name, mod, sym, nv, want in golden_calls():
        sym, nv = np.ascontiguousarray(sym, np.float32), np.ascontiguousarray(nv, np.float32)
        got = np.zeros(want.size, np.int8)
        rc = lib.adt_demodulate_soft(ctypes.c_void_p(got.ctypes.data), ctypes.c_void_p(sym.ctypes.data),
                                     ctypes.c_void_p(nv.ctypes.data), ctypes.c_uint(nv.size), ctypes.c_int(mod))
        assert rc == 0
        assert np.array_equal(got, want), name


@pytest.mark.gpu
def test_cuda_pdsch_encoder_adapter_reproduces_the_golden_codewords():
    """srsran::pdsch_encoder implemented over pdc_pdsch_encode (adapters/pdsch_encoder_cuda.cpp), driven through the reference's
    interface, against the codeword digests recorded from the compiled pdsch_encoder_impl (tests/golden/pdsch_enc_v1.npz)."""
    import hashlib

    from tools.enc_cases import CASES

    lib = _lib()
    g = np.load(os.path.join(HERE, "golden", "pdsch_enc_v1.npz"))
    for k, (bg, tbs, Qm, L, nsym, rv, Nref) in enumerate(CASES):
        tb = np.random.default_rng(500 + k).integers(0, 256, tbs // 8, dtype=np.uint8)
        cw = np.zeros(nsym * Qm, np.uint8)
        rc = lib.adt_pdsch_encode(ctypes.c_void_p(cw.ctypes.data), ctypes.c_void_p(tb.ctypes.data), ctypes.c_uint(tbs),
                                  ctypes.c_int(bg), ctypes.c_uint(rv), ctypes.c_uint(Qm), ctypes.c_uint(Nref),
                                  ctypes.c_uint(L), ctypes.c_uint(nsym))
        assert rc == 0
        got = np.frombuffer(hashlib.sha256(np.packbits(cw).tobytes()).digest(), np.uint8)
        assert np.array_equal(got, g["digests"][k]), CASES[k]


@pytest.mark.gpu
def test_cuda_pdsch_enc_accelerator_inside_the_reference_hw_encoder():
    """hal::hw_accelerator_pdsch_enc over pdc_pdsch_encode (adapters/hw_accelerator_pdsch_enc_cuda.cpp) inside the reference's
    UNMODIFIED pdsch_encoder_hw_impl (TB mode: the reference segments and computes the TB CRC on the host, hands over
    hal::hw_pdsch_encoder_configuration), against the digests recorded from the compiled pdsch_encoder_impl."""
    import hashlib

    from tools.enc_cases import CASES

    lib = _lib()
    g = np.load(os.path.join(HERE, "golden", "pdsch_enc_v1.npz"))
    for k, (bg, tbs, Qm, L, nsym, rv, Nref) in enumerate(CASES):
        tb = np.random.default_rng(500 + k).integers(0, 256, tbs // 8, dtype=np.uint8)
        cw = np.zeros(nsym * Qm, np.uint8)
        rc = lib.adt_pdsch_encode_hw(ctypes.c_void_p(cw.ctypes.data), ctypes.c_void_p(tb.ctypes.data), ctypes.c_uint(tbs),
                                     ctypes.c_int(bg), ctypes.c_uint(rv), ctypes.c_uint(Qm), ctypes.c_uint(Nref),
                                     ctypes.c_uint(L), ctypes.c_uint(nsym))
        assert rc == 0
        got = np.frombuffer(hashlib.sha256(np.packbits(cw).tobytes()).digest(), np.uint8)
        assert np.array_equal(got, g["digests"][k]), CASES[k]


@pytest.mark.gpu
@pytest.mark.parametrize("block,use_view", [(0, 0), (997, 0), (1536, 1)])
def test_cuda_ulsch_demultiplex_adapter_reproduces_the_golden_streams(block, use_view):
    """srsran::ulsch_demultiplex over pdc_ulsch_demultiplex (adapters/ulsch_demultiplex_cuda.cpp), driven through the reference's
    interface the way pusch_demodulator_impl does (blocks of soft bits + their scrambling bits, optionally written through
    get_next_block_view; CSI Part 2 configured up front or from the CSI Part 1 end-of-stream callback; the seed of the scrambling
    sequence is recovered from the sequence), against the streams recorded from the compiled ulsch_demultiplex_impl."""
    import hashlib

    from oracle.demux_oracle import CFG_FIELDS
    from tests import demux_cases as C

    lib = _lib()
    g = np.load(os.path.join(HERE, "golden", "ulsch_demux_v1.npz"))
    vp = ctypes.c_void_p
    for k, cfg in enumerate(C.cases()):
        cw = C.codeword(cfg, 100 + k)
        c = np.array([cfg[f] for f in CFG_FIELDS] + [block, use_view], np.uint32)
        outs = [np.zeros(cw.size, np.int8) for _ in range(4)]
        lens = np.zeros(4, np.uint32)
        rc = lib.adt_ulsch_demultiplex(vp(c.ctypes.data), vp(cw.ctypes.data), ctypes.c_uint(cw.size),
                                       *[vp(o.ctypes.data) for o in outs], ctypes.c_uint(cw.size), vp(lens.ctypes.data))
        assert rc == 0, (k, rc)
        for q in range(4):
            assert lens[q] == g["lengths"][k, q], (k, q)
            got = np.frombuffer(hashlib.sha256(outs[q][:int(lens[q])].tobytes()).digest(), np.uint8)
            assert np.array_equal(got, g["digests"][k, q]), (k, q, cfg)


def test_n3_patch_applies_and_compiles():
    """Product wiring: integration/srsran_cuda.patch (a `hal: cuda_hwacc:` section, the CUDA HAL factories beside the
    bbdev ones, "cuda" as LDPC decoder / rate dematcher type, the CMake option) applies to the reference tree and the patched
    translation units compile against the reference's headers and the adapters, with and without SRSRAN_CUDA_FOUND."""
    import subprocess

    if not os.path.isdir("/root/reference/apps/units/flexible_o_du/o_du_low"):
        pytest.skip("no reference tree on this machine")
    r = subprocess.run(["bash", os.path.join(os.path.dirname(HERE), "integration", "check_patch.sh")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "patch applies and the patched files compile" in r.stdout


@pytest.mark.gpu
def test_n3_patched_hal_factory_returns_the_cuda_accelerators():
    """The reference's make_du_low_hal_dependencies with the patch applied (compiled into the harness), given a `cuda_hwacc`
    configuration: the PUSCH decoder / PDSCH encoder HAL factories it returns are the CUDA ones and create accelerators."""
    lib = _lib()
    assert lib.adt_n3_hal_dependencies(1, 1, 1) == 7
    assert lib.adt_n3_hal_dependencies(1, 0, 0) == 5
    assert lib.adt_n3_hal_dependencies(0, 1, 1) == 6
