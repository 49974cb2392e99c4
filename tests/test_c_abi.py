"""The C ABI of include/pusch_dec_cuda.h as a contract.

CPU part (no GPU needed, no compute calls): the shared library loads, exports every function the header declares, the structure
layouts the Python binding assumes are the compiler's, the host-only helpers answer, and WITHOUT a CUDA device the product
refuses to start (PDC_ERR_NO_DEVICE: there is no CPU fallback).

GPU part: error codes instead of aborts (the reference's srsran_assert contracts: ldpc_decoder_impl.cpp:64-80,
ldpc_rate_dematcher_impl.cpp:52-60), ticket life cycle (enqueue / "not ready" / dequeue of hw_accelerator.h:37-57), HARQ arena
access, the rate-dematcher-only operator, thread safety of one context driven from several host threads (the reference drives
its accelerator objects concurrently from executor threads, factories.cpp:131-138), and device-pointer entry == host entry.
"""
import ctypes
import os
import re
import subprocess
import threading

import numpy as np
import pytest

from oracle.pyoracle import CRC16, CRC24B, FLAVOUR_AVX512, FLAVOUR_GENERIC, LIFTING_SIZES, kb_of, nshort_of

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pusch_dec_cuda.h")


def header_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pdc_[a-z0-9_]+)\s*\(", text)))


# ---- CPU -------------------------------------------------------------------------------------------------------------------
def test_library_exports_every_declared_function():
    from srsran_project_b200 import binding as B

    lib = B.load_library()
    declared = header_functions()
    assert len(declared) >= 19
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/pusch_dec_cuda.h but not exported"
    assert sorted(B.EXPORTED_SYMBOLS) == declared
    nm = subprocess.run(["nm", "-D", "--defined-only", B.library_path()], capture_output=True, text=True).stdout
    for name in declared:
        assert re.search(rf"\bT {name}\b", nm), name
    version = int(re.search(r"#define PDC_ABI_VERSION (\d+)", open(HEADER).read()).group(1))
    assert lib.pdc_abi_version() == version


def test_structure_layouts_match_the_compiler(tmp_path):
    """The numpy / ctypes mirrors in binding.py against sizeof / offsetof of the real header (gcc)."""
    from srsran_project_b200 import binding as B

    structs = {"pdc_cb_desc": B.DESC_DTYPE, "pdc_cb_result": B.RESULT_DTYPE, "pdc_tb_desc": B.TB_DESC_DTYPE,
               "pdc_tb_result": B.TB_RESULT_DTYPE, "pdc_demod_seg": B.DEMOD_SEG_DTYPE, "pdc_pdsch_tb_desc": B.PDSCH_TB_DTYPE,
               "pdc_ulsch_demux_desc": B.ULSCH_DEMUX_DTYPE}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(void) {"]
    for s, dt in structs.items():
        lines.append(f'printf("{s} %zu\\n", sizeof({s}));')
        for f in dt.names:
            lines.append(f'printf("{s}.{f} %zu\\n", offsetof({s}, {f}));')
    lines.append('printf("pdc_config %zu\\n", sizeof(pdc_config));')
    lines += ["return 0; }"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c11", "-o", str(exe), str(src)])
    got = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True).stdout.splitlines())
    for s, dt in structs.items():
        assert int(got[s]) == dt.itemsize, s
        for f in dt.names:
            assert int(got[f"{s}.{f}"]) == dt.fields[f][1], (s, f)
    assert int(got["pdc_config"]) == ctypes.sizeof(B.PdcConfig)


def test_host_only_helpers():
    from srsran_project_b200 import binding as B

    lib = B.load_library()
    for bg in (1, 2):
        for Z in LIFTING_SIZES:
            assert lib.pdc_cb_message_bits(bg, Z) == kb_of(bg) * Z
            assert lib.pdc_cb_buffer_len(bg, Z) == nshort_of(bg) * Z
        for Z in (0, 1, 17, 19, 385, 512):
            assert lib.pdc_cb_message_bits(bg, Z) == 0 and lib.pdc_cb_buffer_len(bg, Z) == 0
    assert lib.pdc_cb_message_bits(3, 384) == 0
    # layers the decoder can touch: a high-rate first transmission stays in the four core rows, a long one does not
    d = B.make_descs(2)
    d["bg"], d["Z"], d["Qm"], d["crc_type"], d["max_iter"], d["flags"] = 1, 384, 8, CRC24B, 6, B.PDC_F_NEW_DATA
    d["E"], d["nof_filler"] = [8960, 25000], 16
    layers = [lib.pdc_cb_max_layers(d[i:i + 1].ctypes.data) for i in range(2)]
    assert layers[0] == 4 and layers[1] == 46
    d["Z"] = 17
    assert lib.pdc_cb_max_layers(d[0:1].ctypes.data) == 0


def test_pdsch_segmentation_helper():
    """pdc_pdsch_segmentation (host-only) against the TX-chain restatement of TS 38.212 5.2.2 (tools/stimulus, pinned to the
    compiled reference by tests/test_enc_oracle.py) over a sweep of transport block sizes, and its contract."""
    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    rng = np.random.default_rng(12)
    sizes = [8, 24, 96, 288, 296, 3816, 3824, 3832, 8424, 8448, 8456, 16896, 25344, 1277968]
    sizes += [int(v) * 8 for v in rng.integers(1, 159746, 200)]
    for bg in (1, 2):
        for tbs in sizes:
            seg = S.segmentation(bg, tbs)
            C, Z, Kp, F = B.pdsch_segmentation(bg, tbs)
            assert (C, Z, F) == (seg["C"], seg["Z"], seg["F"]), (bg, tbs)
            assert Kp + F == seg["K"], (bg, tbs)
    for bg, tbs in ((0, 96), (3, 96), (1, 0), (1, 100), (1, 1277976)):
        with pytest.raises(B.PdcError):
            B.pdsch_segmentation(bg, tbs)


def test_no_device_means_no_context():
    """No CPU fallback: without a CUDA device the library refuses to create a context and says why."""
    from srsran_project_b200 import binding as B

    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(B.PdcError) as e:
        B.PuschDecCuda(device=0, max_batch_cbs=4, max_batch_llrs=1 << 16, harq_slots=0, nof_tickets=1)
    assert e.value.status == B.PDC_ERR_NO_DEVICE and "no CPU fallback" in str(e.value)


# ---- GPU -------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def acc():
    import srsran_project_b200 as pdc

    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=64, max_batch_llrs=4 << 20, harq_slots=16, nof_tickets=3)
    yield ctx
    ctx.close()


def valid_batch(B, n=4, Z=96, bg=2, seed=0, slot0=None):
    """n decodable BG2 code blocks (noisy valid codewords, CRC16)."""
    from tools import stimulus as S

    rng = np.random.default_rng(seed)
    K, F, E = kb_of(bg) * Z, 8, 1600
    d = B.make_descs(n)
    llrs = np.zeros(n * 1600, np.int8)
    for i in range(n):
        payload = rng.integers(0, 2, K - F - 16, dtype=np.uint8)
        c = S.crc(1, payload)
        msg = np.concatenate([payload, [(c >> (15 - k)) & 1 for k in range(16)], np.zeros(F, np.uint8)]).astype(np.uint8)
        tx = S.rate_match(bg, Z, S.encode(bg, Z, msg), 0, 2, 0, F, E)
        llrs[i * E:(i + 1) * E] = S.awgn_llrs(tx, 8.0, 3.0, 17 * seed + i)
    d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = bg, Z, 2, 0, CRC16, 6
    d["flags"] = B.PDC_F_NEW_DATA | B.PDC_F_EARLY_STOP
    d["E"], d["nof_filler"] = E, F
    d["llr_offset"] = np.arange(n) * E
    d["out_offset"] = np.arange(n) * ((K + 31) // 32 * 4)
    d["soft_offset"] = np.arange(n) * 25344
    if slot0 is not None:
        d["harq_slot"] = slot0 + np.arange(n)
    return d, llrs, n * ((K + 31) // 32 * 4)


@pytest.mark.gpu
def test_contract_violations_are_error_codes(acc):
    from srsran_project_b200 import binding as B

    d0, llrs, nb = valid_batch(B)
    want, _ = acc.run(d0, llrs, nb, np.zeros(4 * 25344, np.int8))
    assert (want["crc_pass"] == 1).all()
    bad = [("bg", 3), ("Z", 17), ("Z", 0), ("max_iter", 0), ("Qm", 3), ("rv", 4), ("crc_type", 9), ("E", 1601),
           ("nof_filler", 960), ("out_offset", 2), ("llr_offset", llrs.size), ("E", 0), ("Nref", 30000)]
    for field, value in bad:
        d = d0.copy()
        d[1][field] = value
        with pytest.raises(B.PdcError) as e:
            acc.submit(d, llrs, np.zeros(4 * 25344, np.int8))
        assert e.value.status == B.PDC_ERR_INVALID_ARG, (field, value)
    # host-exchanged soft buffers that do not cover every code block
    with pytest.raises(B.PdcError) as e:
        acc.submit(d0, llrs, np.zeros(3 * 25344, np.int8))
    assert e.value.status == B.PDC_ERR_INVALID_ARG
    # an arena slot that does not exist
    d = d0.copy()
    d["harq_slot"] = [0, 1, 2, 16]
    with pytest.raises(B.PdcError) as e:
        acc.submit(d, llrs)
    assert e.value.status == B.PDC_ERR_INVALID_ARG
    # nothing, and too much
    assert acc.lib.pdc_submit(acc.h, d0.ctypes.data, 0, llrs.ctypes.data, llrs.nbytes, None, 0,
                              ctypes.byref(ctypes.c_int64())) == B.PDC_ERR_INVALID_ARG
    big, big_llrs, _ = B.make_descs(65), llrs, nb
    with pytest.raises(B.PdcError) as e:
        acc.submit(big, big_llrs)
    assert e.value.status == B.PDC_ERR_CAPACITY
    with pytest.raises(B.PdcError) as e:
        acc.submit(d0, np.zeros((4 << 20) + 16, np.int8))
    assert e.value.status == B.PDC_ERR_CAPACITY
    # none of the rejected calls left anything behind: the same batch still decodes to the same result
    got, _ = acc.run(d0, llrs, nb, np.zeros(4 * 25344, np.int8))
    assert (got == want).all()


@pytest.mark.gpu
def test_ticket_life_cycle(acc):
    from srsran_project_b200 import binding as B

    d, llrs, nb = valid_batch(B, seed=1, slot0=0)
    tickets = [acc.submit(d, llrs) for _ in range(3)]  # nof_tickets = 3
    assert len(set(tickets)) == 3
    with pytest.raises(B.PdcError) as e:
        acc.submit(d, llrs)  # the caller has to dequeue first (enqueue_operation returning false)
    assert e.value.status == B.PDC_ERR_CAPACITY
    bits = np.zeros(nb, np.uint8)
    first = acc.fetch(tickets[0], 4, bits)
    while not acc.poll(tickets[1]):  # "not ready" is a status, not an error
        pass
    for t in tickets[1:]:
        assert (acc.fetch(t, 4, bits) == first).all()
    with pytest.raises(B.PdcError) as e:
        acc.fetch(tickets[0], 4, bits)  # already handed back
    assert e.value.status == B.PDC_ERR_BAD_TICKET
    with pytest.raises(B.PdcError) as e:
        acc.poll(123456)
    assert e.value.status == B.PDC_ERR_BAD_TICKET
    assert acc.submit(d, llrs) not in tickets  # tickets are never reused
    acc.fetch(acc.submit(d, llrs) - 1, 4, bits)


@pytest.mark.gpu
def test_harq_arena_access(acc):
    rng = np.random.default_rng(2)
    x = rng.integers(-120, 121, 25344, dtype=np.int8)
    acc.harq_write(5, x)
    assert (acc.harq_read(5, 25344) == x).all() and (acc.harq_read(5, 100) == x[:100]).all()
    acc.harq_free(5)
    assert (acc.harq_read(5, 25344) == 0).all()
    from srsran_project_b200 import binding as B

    for call in (lambda: acc.harq_free(16), lambda: acc.harq_read(-1, 10), lambda: acc.harq_write(3, np.zeros(25345, np.int8))):
        with pytest.raises(B.PdcError) as e:
            call()
        assert e.value.status == B.PDC_ERR_INVALID_ARG


@pytest.mark.gpu
@pytest.mark.parametrize("flavour", [FLAVOUR_AVX512, FLAVOUR_GENERIC])
def test_rate_dematcher_operator_alone(acc, oracle, flavour):
    """PDC_F_DEMATCH_ONLY = ldpc_rate_dematcher::rate_dematch: new data and combining over a stale buffer, every Qm, rv, LBRM,
    several laps; the soft buffer must equal the oracle's and nothing else is produced."""
    from srsran_project_b200 import binding as B

    rng = np.random.default_rng(3)
    cases = []
    for bg, Z in ((1, 384), (1, 22), (2, 96), (2, 7), (2, 208)):
        N, K = nshort_of(bg) * Z, kb_of(bg) * Z
        for Qm in (1, 2, 4, 6, 8):
            rv, new = int(rng.integers(0, 4)), bool(rng.integers(0, 2))
            F = int(rng.integers(0, Z))
            Nref = int(rng.choice([0, 0, N - int(rng.integers(1, N // 3))]))
            E = Qm * int(rng.integers(1, (3 * N) // Qm))
            cases.append((bg, Z, Qm, rv, new, F, Nref, E))
    n = len(cases)
    d = B.make_descs(n)
    soft = rng.integers(-120, 121, n * 25344, dtype=np.int8)
    want = soft.copy()
    lo, chunks = 0, []
    for i, (bg, Z, Qm, rv, new, F, Nref, E) in enumerate(cases):
        x = rng.integers(-120, 121, E, dtype=np.int8)
        chunks.append(x)
        d[i]["bg"], d[i]["Z"], d[i]["Qm"], d[i]["rv"], d[i]["crc_type"], d[i]["max_iter"] = bg, Z, Qm, rv, CRC16, 1
        d[i]["flags"] = B.PDC_F_DEMATCH_ONLY | (B.PDC_F_NEW_DATA if new else 0) | \
            (B.PDC_F_GENERIC if flavour == FLAVOUR_GENERIC else 0)
        d[i]["E"], d[i]["Nref"], d[i]["nof_filler"] = E, Nref, F
        d[i]["llr_offset"], d[i]["soft_offset"], d[i]["out_offset"] = lo, i * 25344, i * 1056
        lo += (E + 15) // 16 * 16
        buf = want[i * 25344:i * 25344 + nshort_of(bg) * Z]
        oracle.rate_dematch(flavour, buf, x, new, rv, Qm, Nref, F)
    llrs = np.zeros(lo, np.int8)
    for i, x in enumerate(chunks):
        llrs[int(d[i]["llr_offset"]):int(d[i]["llr_offset"]) + x.size] = x
    res, _ = acc.run(d, llrs, n * 1056, soft)
    assert (res["status"] == 0).all() and (res["crc_pass"] == 0).all() and (res["iterations"] == -1).all()
    for i, (bg, Z, *_rest) in enumerate(cases):
        N = nshort_of(bg) * Z
        got, exp = soft[i * 25344:i * 25344 + N], want[i * 25344:i * 25344 + N]
        assert (got == exp).all(), (cases[i], int((got != exp).sum()))
        assert (soft[i * 25344 + N:(i + 1) * 25344] == want[i * 25344 + N:(i + 1) * 25344]).all()  # nothing beyond N


@pytest.mark.gpu
def test_one_context_many_host_threads():
    """Four host threads share one context (its tickets): every submission gets its own results."""
    import srsran_project_b200 as pdc
    from srsran_project_b200 import binding as B

    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=16, max_batch_llrs=1 << 20, harq_slots=0, nof_tickets=4)
    try:
        work = [valid_batch(B, n=6, Z=(96, 104, 112, 120)[k], seed=10 + k) for k in range(4)]
        want = []
        for d, llrs, nb in work:
            soft = np.zeros(6 * 25344, np.int8)
            want.append(ctx.run(d, llrs, nb, soft) + (soft,))
        errors = []

        def worker(k):
            d, llrs, nb = work[k]
            try:
                for _ in range(25):
                    soft = np.zeros(6 * 25344, np.int8)
                    res, bits = ctx.run(d, llrs, nb, soft)
                    if not ((res == want[k][0]).all() and (bits == want[k][1]).all() and (soft == want[k][2]).all()):
                        errors.append((k, "mismatch"))
            except Exception as e:  # noqa: BLE001 - reported below
                errors.append((k, repr(e)))

        threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not errors, errors
    finally:
        ctx.close()


@pytest.mark.gpu
def test_device_pointer_entry_equals_host_entry(acc):
    import torch

    from srsran_project_b200 import binding as B

    d, llrs, nb = valid_batch(B, n=8, Z=384, bg=1, seed=4)
    soft = np.zeros(8 * 25344, np.int8)
    res, bits = acc.run(d, llrs, nb, soft)
    d_llrs = torch.from_numpy(llrs).cuda()
    d_soft = torch.zeros(8 * 25344, dtype=torch.int8, device="cuda")
    d_bits = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(8 * 8, dtype=torch.uint8, device="cuda")
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        acc.decode_device(d, d_llrs.data_ptr(), d_soft.data_ptr(), d_bits.data_ptr(), d_res.data_ptr(), side.cuda_stream)
    side.synchronize()
    assert (d_res.cpu().numpy().view(B.RESULT_DTYPE) == res).all()
    assert (d_bits.cpu().numpy() == bits).all() and (d_soft.cpu().numpy() == soft).all()


@pytest.mark.gpu
def test_pinned_source_may_be_overwritten_after_submit_unless_zero_copy(acc):
    """include/pusch_dec_cuda.h: by default pdc_submit stages the LLRs before it returns, page-locked source or not, so the
    caller may refill its buffer at once; with pdc_set_zero_copy the source is read in place (same results when left alone)."""
    import torch

    from srsran_project_b200 import binding as B

    d, llrs, nb = valid_batch(B, n=16, Z=384, bg=1, seed=21)
    want_res, want_bits = acc.run(d, llrs, nb, np.zeros(16 * 25344, np.int8))
    pinned = torch.empty(llrs.size, dtype=torch.int8).pin_memory()
    src = pinned.numpy()
    for zero_copy in (False, True):
        acc.set_zero_copy(zero_copy)
        try:
            for _ in range(8):
                src[:] = llrs
                t = acc.submit(d, src, np.zeros(16 * 25344, np.int8))
                if not zero_copy:
                    src[:] = 0x55  # the contract: the call has copied what it needs
                bits = np.zeros(nb, np.uint8)
                res = acc.fetch(t, 16, bits, np.zeros(16 * 25344, np.int8))
                assert (res == want_res).all() and (bits == want_bits).all(), zero_copy
        finally:
            acc.set_zero_copy(False)


@pytest.mark.gpu
def test_device_entry_from_two_threads_on_two_streams():
    """pdc_decode_device borrows a ticket's descriptor staging: calls from different host threads on different streams must not
    overwrite descriptors that earlier kernels are still reading (more calls than tickets, so staging is reused)."""
    import torch

    import srsran_project_b200 as pdc
    from srsran_project_b200 import binding as B

    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=64, max_batch_llrs=4 << 20, harq_slots=0, nof_tickets=2)
    try:
        work, want = [], []
        for k in range(2):
            d, llrs, nb = valid_batch(B, n=48, Z=(384, 352)[k], bg=1, seed=30 + k)
            want.append(ctx.run(d, llrs, nb, np.zeros(48 * 25344, np.int8)))
            work.append((d, torch.from_numpy(llrs).cuda(), nb))
        errors = []

        def worker(k):
            try:
                torch.cuda.set_device(0)
                d, d_llrs, nb = work[k]
                side = torch.cuda.Stream()
                outs = []
                for _ in range(12):  # back to back, no synchronisation in between: 12 launches per stream over 2 tickets
                    d_soft = torch.zeros(48 * 25344, dtype=torch.int8, device="cuda")
                    d_bits = torch.zeros(nb, dtype=torch.uint8, device="cuda")
                    d_res = torch.zeros(48 * 8, dtype=torch.uint8, device="cuda")
                    side.wait_stream(torch.cuda.current_stream())
                    ctx.decode_device(d, d_llrs.data_ptr(), d_soft.data_ptr(), d_bits.data_ptr(), d_res.data_ptr(),
                                      side.cuda_stream)
                    outs.append((d_res, d_bits, d_soft))
                side.synchronize()
                for d_res, d_bits, _ in outs:
                    if not ((d_res.cpu().numpy().view(B.RESULT_DTYPE) == want[k][0]).all() and
                            (d_bits.cpu().numpy() == want[k][1]).all()):
                        errors.append((k, "mismatch"))
            except Exception as e:  # noqa: BLE001 - reported below
                errors.append((k, repr(e)))

        threads = [threading.Thread(target=worker, args=(k,)) for k in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not errors, errors
    finally:
        ctx.close()
