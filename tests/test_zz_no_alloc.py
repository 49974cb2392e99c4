"""pdc_decode_device_tbs_static: the allocation-free variant of pdc_decode_device_tbs (scratch in a file-scope static device array,
at most PDC_STATIC_MAX_CBS code blocks per call). It has no counterpart in the reference.
Checked through the C ABI against the existing entry point, bit for bit (everything on this path is integer /
byte work: the tolerance is zero), on two calls in a row that share the scratch array: a seven-TB batch with a failing code block
and a wrong TB CRC, then a smaller and different batch that finds the first call's contents in the array.

Collected last on purpose."""
import numpy as np
import pytest

from tests.test_tb_assembly import TB_CASES, build_batch


@pytest.fixture(scope="module")
def acc():
    import srsran_project_b200 as pdc

    ctx = pdc.PuschDecCuda(device=0, max_batch_cbs=2048, max_batch_llrs=16 << 20, harq_slots=64, nof_tickets=2)
    yield ctx
    ctx.close()


def run_both(acc, B, b, fill):
    """-> [(cb results, cb bits, soft buffers, tb results, tb bytes)] of the existing entry point and of the static one."""
    import torch

    dev = "cuda:0"
    d_llrs = torch.from_numpy(b["llrs"]).to(dev)
    out = []
    for fn in (acc.decode_device_tbs, acc.decode_device_tbs_static):
        d_soft = torch.zeros(b["n"] * 25344, dtype=torch.int8, device=dev)
        d_bits = torch.full((b["bits_bytes"],), fill, dtype=torch.uint8, device=dev)
        d_res = torch.zeros(b["n"] * 8, dtype=torch.uint8, device=dev)
        d_tb = torch.full((b["tb_bytes"],), fill, dtype=torch.uint8, device=dev)
        d_tbres = torch.zeros(len(b["tbd"]) * 4, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        fn(b["descs"], b["tbd"], d_llrs.data_ptr(), d_soft.data_ptr(), d_bits.data_ptr(), d_res.data_ptr(), d_tb.data_ptr(),
           d_tbres.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        out.append((d_res.cpu().numpy().view(B.RESULT_DTYPE), d_bits.cpu().numpy(), d_soft.cpu().numpy(),
                    d_tbres.cpu().numpy().view(B.TB_RESULT_DTYPE), d_tb.cpu().numpy()))
    return out


def assert_same(want, got):
    for name, w, g in zip(("cb results", "cb bits", "soft buffers", "tb results", "tb bytes"), want, got):
        assert (w == g).all(), name


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_the_existing_one_twice(acc):
    from srsran_project_b200 import binding as B
    from tools import stimulus as S

    # call 1: seven TBs (C = 1..16, both base graphs); one code block of TB 3 is hopeless, TB 5 carries a wrong TB CRC
    sigma = lambda k, r: 40.0 if (k == 3 and r == 1) else 2.0
    b1 = build_batch(S, B, TB_CASES, sigma, False, seed0=11, corrupt={5})
    want, got = run_both(acc, B, b1, 0x5A)
    assert_same(want, got)
    tb = got[3]
    assert not tb[3]["written"] and tb[5]["written"] and not tb[5]["tb_crc_ok"]
    assert all(tb[k]["tb_crc_ok"] for k in (0, 1, 2, 4, 6))
    # call 2, right behind it: fewer and different transport blocks over the same scratch array (order: multi-CB TBs first, so
    # every descriptor, order entry and CRC share of call 1 is still lying in the array where call 2 does not overwrite it)
    b2 = build_batch(S, B, [TB_CASES[4], TB_CASES[2], TB_CASES[0]], lambda k, r: 2.0, False, seed0=77)
    want, got = run_both(acc, B, b2, 0x3C)
    assert_same(want, got)
    assert got[3]["tb_crc_ok"].all() and got[3]["written"].all()
    for k, tx in enumerate(b2["txs"]):
        o, nbytes = int(b2["tbd"][k]["tb_offset"]), int(b2["tbd"][k]["tbs_bits"]) // 8
        assert (got[4][o:o + nbytes] == np.packbits(tx["tb"])).all()
    # and call 1 again (the static scratch now holds call 2's contents)
    want, got = run_both(acc, B, b1, 0x5A)
    assert_same(want, got)


@pytest.mark.gpu
def test_static_scratch_entry_point_rejects_what_does_not_fit(acc):
    import torch

    from srsran_project_b200 import binding as B

    n = B.PDC_STATIC_MAX_CBS + 1
    descs = B.make_descs(n)
    descs["bg"], descs["Z"], descs["Qm"], descs["crc_type"], descs["max_iter"] = 2, 8, 2, B.PDC_CRC16, 6
    descs["flags"], descs["E"] = B.PDC_F_NEW_DATA | B.PDC_F_EARLY_STOP, 160
    descs["llr_offset"], descs["out_offset"], descs["soft_offset"] = np.arange(n) * 160, np.arange(n) * 12, np.arange(n) * 400
    tbd = np.zeros(n, B.TB_DESC_DTYPE)
    tbd["first_cb"], tbd["nof_cbs"], tbd["tbs_bits"], tbd["tb_offset"] = np.arange(n), 1, 64, np.arange(n) * 8
    d = torch.zeros(n * 400, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(B.PdcError) as e:
        acc.decode_device_tbs_static(descs, tbd, d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(),
                                     d.data_ptr(), None)
    assert e.value.status == B.PDC_ERR_CAPACITY and "PDC_STATIC_MAX_CBS" in str(e.value)
    torch.cuda.synchronize()
