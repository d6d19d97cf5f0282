#!/usr/bin/env python
"""Like race_hunt.py, but the cases are INTERLEAVED, so that a fresh handle's device buffers reuse memory that a
different shape has just written (an uninitialised read would then see foreign data, as it does inside pytest)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from tests.helpers import Handle, make_queries  # noqa: E402

reps = int(os.environ.get("REPS", "300"))
cases = [(3000, 100, 3, 4), (5000, 512, 1, 5), (777, 16, 1, 8), (1 << 13, 1024, 1, 9), (40000, 256, 1, 70), (1 << 12, 2048, 2, 17)]
prep = []
for db_size, slot_bytes, W, nq in cases:
    data = oracle.synth_fill(21, 0, db_size, slot_bytes)
    rows = db_size // W
    rng = np.random.default_rng(nq)
    idxs = [int(x) for x in rng.integers(0, rows, size=nq)]
    pairs = [make_queries(100 + i, db_size, W, idx) for i, idx in enumerate(idxs)]
    qs = [p[1][i & 1] for i, p in enumerate(pairs)]
    want = np.stack([np.asarray(oracle.answer(p[0][i & 1], data, db_size, slot_bytes)).reshape(-1) for i, p in enumerate(pairs)])
    prep.append((db_size, slot_bytes, W, nq, qs, want, data, rows))
bad = {}
for rep in range(reps):
    for ci, (db_size, slot_bytes, W, nq, qs, want, data, rows) in enumerate(prep):
        with Handle(db_size, slot_bytes) as h:
            h.fill(21)
            for mode in ("bits", "fused"):
                os.environ["PIRB_BATCH_MODE"] = mode
                got = np.asarray(h.answer_batch(qs)).reshape(nq, -1)
                if not np.array_equal(got, want):
                    key = (ci, mode)
                    bad[key] = bad.get(key, 0) + 1
                    if bad[key] <= 3:
                        qi = [i for i in range(nq) if not np.array_equal(got[i], want[i])]
                        diff = got[qi[0]] ^ want[qi[0]]
                        rb = W * slot_bytes
                        single = [r for r in range(rows) if np.array_equal(diff, data[r * rb:(r + 1) * rb])]
                        print("  rep %d case %s mode %s: queries %s differ; first: %d differing bytes; equals row(s) %s" % (
                            rep, cases[ci], mode, qi, int(np.count_nonzero(diff)), single[:4]), flush=True)
print("interleaved %d reps x %d cases: mismatches %s" % (reps, len(cases), bad), flush=True)
