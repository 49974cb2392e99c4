"""N > 1 host logic on CPU: two gloo ranks run bench.py's rank plumbing (Ranks, tb_shard, build_workload, make_descs).
The data path has no collective (code blocks are independent): the only cross-rank traffic is the
barrier and the max / sum of scalars, which is what is checked here. Also checks the reference arm's rank handling."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(script_args, world, extra_env=None):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1", **(extra_env or {}))
        procs.append(subprocess.Popen([sys.executable] + script_args, env=env, cwd=ROOT, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            p.kill()
            raise
        assert p.returncode == 0, e[-2000:]
        outs.append(o)
    return outs


def test_two_ranks_shard_by_transport_block_and_reduce():
    outs = _spawn([os.path.join(ROOT, "tests", "gloo_worker.py")], 2)
    assert outs[1].strip() == ""  # only rank 0 prints
    r = json.loads(outs[0].strip().splitlines()[-1])
    assert r["world"] == 2 and r["sum_rank"] == 1.0
    assert r["tbs"] == [[0, 1], [2, 3]]  # disjoint transport blocks per rank (weak scaling)
    assert r["max_digest"] != r["min_digest"]  # the ranks decoded different LLRs
    assert r["sum_ok"] == 2 * 2 * 3  # every sampled code block of both ranks decoded to the transmitted bits
    assert r["sum_info_bits"] == 2 * 2 * 1277992
    assert r["max_sec"] <= r["sum_sec"] and r["max_sec"] >= r["sum_sec"] / 2
    assert r["harq_slots_rank_local"]


def test_reference_arm_runs_on_rank0_only():
    from oracle import pyoracle as O

    if not O.ref_available():
        pytest.skip("oracle/_ref not built")
    outs = _spawn([os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                   "--tbs-per-step", "2"], 2)
    assert outs[1].strip() == ""
    line = json.loads(outs[0].strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "reference" and line["e2e"]["h2d_bytes_per_step"] == 0
    for k in ("metric", "unit", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "config"):
        assert k in line


def test_numa_cpulist_parsing_and_noop_binding(monkeypatch):
    """Host placement helper (srsran_project_b200/numa.py): sysfs cpulist parsing, and binding is a no-op when the device's CPUs
    are all the process may use anyway (the single-socket box) or cannot be told."""
    from srsran_project_b200 import numa

    assert numa._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa._parse_cpulist("") == set()
    allowed = os.sched_getaffinity(0)
    monkeypatch.setattr(numa, "device_local_cpus", lambda device: set(allowed))
    assert numa.bind_to_device_socket(0) is None
    monkeypatch.setattr(numa, "device_local_cpus", lambda device: None)
    assert numa.bind_to_device_socket(0) is None
    assert os.sched_getaffinity(0) == allowed
