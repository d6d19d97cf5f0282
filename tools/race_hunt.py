#!/usr/bin/env python
"""Repeat the batch test case that failed once (3000 slots x 100 B, group size 3, 4 queries) and report which
path (bits = BITS-mode expansion + batch scan, fused = per-query kernels) disagrees with the oracle, and where."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from tests.helpers import Handle, make_queries  # noqa: E402

reps = int(os.environ.get("REPS", "300"))
cases = [(3000, 100, 3, 4), (5000, 512, 1, 5), (777, 16, 1, 8), (1 << 13, 1024, 1, 9)]
if os.environ.get("CASES"):
    cases = [cases[int(i)] for i in os.environ["CASES"].split(",")]
for db_size, slot_bytes, W, nq in cases:
    data = oracle.synth_fill(21, 0, db_size, slot_bytes)
    rows = db_size // W
    rng = np.random.default_rng(nq)
    idxs = [int(x) for x in rng.integers(0, rows, size=nq)]
    pairs = [make_queries(100 + i, db_size, W, idx) for i, idx in enumerate(idxs)]
    qs = [p[1][i & 1] for i, p in enumerate(pairs)]
    want = np.stack([oracle.answer(p[0][i & 1], data, db_size, slot_bytes).reshape(-1) for i, p in enumerate(pairs)])
    bad = {"bits": 0, "fused": 0}
    for rep in range(reps):
        with Handle(db_size, slot_bytes) as h:
            h.fill(21)
            for mode in ("bits", "fused"):
                os.environ["PIRB_BATCH_MODE"] = mode
                got = np.asarray(h.answer_batch(qs)).reshape(nq, -1)
                if not np.array_equal(got, want):
                    bad[mode] += 1
                    if bad[mode] <= 3:
                        qi = [i for i in range(nq) if not np.array_equal(got[i], want[i])]
                        diff = got[qi[0]] ^ want[qi[0]]
                        print("  rep %d mode %s: queries %s differ; first: %d differing bytes, is a single row's XOR: %s" % (
                            rep, mode, qi, int(np.count_nonzero(diff)),
                            any(np.array_equal(diff, data[r * W * slot_bytes:(r + 1) * W * slot_bytes]) for r in range(rows))), flush=True)
    print("case %s: %d reps, mismatches %s" % ((db_size, slot_bytes, W, nq), reps, bad), flush=True)
