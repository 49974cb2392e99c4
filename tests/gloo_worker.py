"""Worker of tests/test_multi_rank.py: one process per rank (gloo, CPU). Runs bench.py's N > 1 host logic — the Ranks
helper, the transport-block shard of this rank, the workload builder and descriptor builder — with the CPU oracle standing
in for the GPU as the "decoder" of the shard (test infrastructure; the product path never does this). Rank 0 prints one
JSON line with the aggregate, like bench.py does."""
import hashlib
import json
import os
import sys
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from oracle import pyoracle as O  # noqa: E402


def main():
    ranks = bench.Ranks("gloo", "cpu")
    args = types.SimpleNamespace(tbs_per_step=2, input_sets=1, population="awgn", sigma=2.2)
    wl = bench.build_workload(args, ranks.rank)
    descs = bench.make_descs(wl, 0, wl["n_cbs"])
    C = wl["seg"]["C"]
    # Decode the first 3 code blocks of each of this rank's transport blocks with the oracle.
    orc = O.Oracle()
    ranks.barrier()
    t0 = time.perf_counter()
    ok = 0
    for t in range(wl["n_tbs"]):
        for r in range(3):
            i = t * C + r
            d = descs[i]
            o = int(wl["offs"][i])
            buf = np.zeros(bench.N_SOFT, np.int8)
            it, bits = orc.codeblock_decode(O.FLAVOUR_AVX512, 1, int(d["Z"]), buf, wl["llr_sets"][0][o:o + int(d["E"])],
                                            True, 0, 8, 0, int(d["nof_filler"]), O.CRC24B, 6, True)
            same = (np.unpackbits(bits)[:wl["seg"]["Kp"]] == wl["msgs"][i, :wl["seg"]["Kp"]]).all()
            ok += int(it > 0 and same)
    sec = time.perf_counter() - t0
    ranks.barrier()
    digest = int(hashlib.sha256(wl["llr_sets"][0].tobytes()).hexdigest()[:12], 16)
    out = {
        "world": ranks.world,
        "max_sec": ranks.max(sec),
        "sum_sec": ranks.sum(sec),
        "sum_ok": ranks.sum(float(ok)),
        "sum_info_bits": ranks.sum(float(wl["info_bits"])),
        "sum_rank": ranks.sum(float(ranks.rank)),
        "max_digest": ranks.max(float(digest)),
        "min_digest": -ranks.max(-float(digest)),
        "tbs": [list(bench.tb_shard(r, wl["n_tbs"])) for r in range(ranks.world)],
        "harq_slots_rank_local": bool((descs["harq_slot"] == np.arange(wl["n_cbs"])).all()),
    }
    if ranks.rank == 0:
        print(json.dumps(out), flush=True)
    ranks.close()


if __name__ == "__main__":
    main()
