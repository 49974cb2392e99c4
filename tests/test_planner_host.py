"""Host logic of the batch planner without a GPU (tests/host/plan_probe.cu, built here with nvcc against libpuschdec_cuda.so):
which kernel list a code block takes and what the runtime's bookkeeping of the HARQ arena (non-zero extent per slot, released
slots, slots to clear) says afterwards. SURVEY 8 config 2': the live stack's limited-buffer rate matching (Nref = 12611 < N) takes
the four-row kernel exactly when the stretch the reference's dematcher leaves untouched (ldpc_rate_dematcher_impl.cpp:198-200) is
known to hold zeros."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_DIR = os.path.join(ROOT, "srsran_project_b200")
SRC = os.path.join(ROOT, "tests", "host", "plan_probe.cu")

LIST_H2_BG1_Z384, LIST_H2G_BG1 = 2, 4  # batch_plan: [2] four-row kernel BG1 Z = 384, [4] many-layer kernel BG1
TILE = 0x80  # PDC_DEVF_FIRST_TX_TILE
N = 25344


@pytest.fixture(scope="module")
def probe(tmp_path_factory):
    if shutil.which("nvcc") is None or not os.path.exists(os.path.join(LIB_DIR, "libpuschdec_cuda.so")):
        pytest.skip("needs nvcc and the built libpuschdec_cuda.so")
    exe = str(tmp_path_factory.mktemp("plan_probe") / "plan_probe")
    subprocess.run(["nvcc", "-O1", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-I", os.path.join(LIB_DIR, "csrc"),
                    "-o", exe, SRC, "-L", LIB_DIR, "-lpuschdec_cuda", "-Xlinker", "-rpath=" + LIB_DIR], check=True,
                   capture_output=True, timeout=600)
    out = subprocess.run([exe], check=True, capture_output=True, text=True, timeout=60).stdout
    rows = {}
    for line in out.splitlines():
        f = line.split()
        rows.setdefault(f[0], []).append(dict(list=int(f[2]), flags=int(f[3]), max_layers=int(f[4]), nz=int(f[5]),
                                              released=int(f[6]), nzero=int(f[7]), status=int(f[8])))
    return rows


def test_limited_buffer_first_transmissions_take_the_four_row_kernel_on_empty_slots(probe):
    a, b = probe["clean"]
    assert a["list"] == b["list"] == LIST_H2_BG1_Z384 and a["flags"] & TILE and b["flags"] & TILE
    assert a["max_layers"] == b["max_layers"] == 4
    assert (a["nz"], b["nz"]) == (8960 + 16, 8992 + 16)


def test_stale_stretch_sends_the_code_block_to_the_general_kernel_until_it_is_covered(probe):
    a, b = probe["swapped"]
    assert a["list"] == LIST_H2_BG1_Z384 and a["nz"] == 9008          # longer than what the slot held: covers it
    assert b["list"] == LIST_H2G_BG1 and not b["flags"] & TILE        # shorter: [8976, 9008) keeps the old soft bits
    assert b["nz"] == 9008 and b["max_layers"] > 4
    (p,) = probe["persist"]
    assert p["list"] == LIST_H2G_BG1 and p["nz"] == 9008
    (c,) = probe["covered"]
    assert c["list"] == LIST_H2_BG1_Z384 and c["nz"] == 9008


def test_retransmissions_raise_the_bound_to_the_whole_buffer(probe):
    first, retx = probe["retx"]
    assert first["list"] == LIST_H2_BG1_Z384 and first["nz"] == 8976
    assert retx["list"] == LIST_H2G_BG1 and retx["nz"] == N
    (after,) = probe["after_retx"]
    assert after["list"] == LIST_H2G_BG1
    # [8976, N - (12611 - 8976)) keeps what the retransmission left, the tail beyond is zeroed
    assert after["nz"] == N - (12611 - 8976)


def test_released_slots_read_as_zeros_and_are_cleared_only_when_needed(probe):
    (r,) = probe["released_first_tx"]
    assert r["list"] == LIST_H2_BG1_Z384 and r["nz"] == 8976 and r["released"] == 0 and r["nzero"] == 0
    (c,) = probe["released_retx"]
    assert c["list"] == LIST_H2G_BG1 and c["nzero"] == 1 and c["released"] == 0 and c["nz"] == N


def test_unlimited_buffer_rewrites_everything_and_rejected_batches_change_nothing(probe):
    (u,) = probe["no_limit"]
    assert u["list"] == LIST_H2_BG1_Z384 and u["flags"] & TILE and u["nz"] == 8976
    rej = probe["rejected"]
    assert all(x["status"] != 0 for x in rej) and rej[0]["nz"] == 777
    (h,) = probe["host_soft"]
    assert h["list"] == LIST_H2G_BG1 and h["status"] == 0
