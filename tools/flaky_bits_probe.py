#!/usr/bin/env python
"""Probe for the intermittent mismatch of the two-phase batch path on 336-byte rows.

The mismatch shows up early in the life of a process (repetitions 13 and 20 of 40 in two of 45 pytest runs, never in
three long-running loops), so this script is meant to be started MANY times (tools/flaky_loop.sh style), each start
doing a short burst.  Per repetition on a fresh handle it runs
  1. pir_answer_batch through the bits path   (answer_kernel<0> on two side streams + batch_scan_kernel),
  2. pir_expand_bits for every query          (answer_kernel<0> on one stream + unpack),
  3. pir_answer_with_bits with the ORACLE's bits (scan only),
and reports which of the three disagrees with the oracle, to tell phase 1 from phase 2."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from tests.helpers import Handle, make_queries  # noqa: E402

db_size, slot_bytes, W, nq = 3000, 100, 3, 4
reps = int(os.environ.get("REPS", "40"))
data = oracle.synth_fill(21, 0, db_size, slot_bytes)
rows = db_size // W
rng = np.random.default_rng(nq)
idxs = [int(x) for x in rng.integers(0, rows, size=nq)]
pairs = [make_queries(100 + i, db_size, W, idx) for i, idx in enumerate(idxs)]
qs = [p[1][i & 1] for i, p in enumerate(pairs)]
os_ = [p[0][i & 1] for i, p in enumerate(pairs)]
want = np.stack([np.asarray(oracle.answer(o, data, db_size, slot_bytes)).reshape(-1) for o in os_])
want_bits = [np.asarray(oracle.expand_shared_query(o, db_size)).astype(np.uint8) for o in os_]
os.environ["PIRB_BATCH_MODE"] = "bits"
bad = {"batch": 0, "expand": 0, "scan": 0}
for rep in range(reps):
    with Handle(db_size, slot_bytes) as h:
        h.fill(21)
        got = np.asarray(h.answer_batch(qs)).reshape(nq, -1)
        if not np.array_equal(got, want):
            bad["batch"] += 1
            qi = [i for i in range(nq) if not np.array_equal(got[i], want[i])]
            diff = got[qi[0]] ^ want[qi[0]]
            rb = W * slot_bytes
            single = [r for r in range(rows) if np.array_equal(diff, data[r * rb:(r + 1) * rb])]
            print("rep %d: batch answer of queries %s differs; %d bytes differ; the difference is row(s) %s" % (
                rep, qi, int(np.count_nonzero(diff)), single[:4]), flush=True)
        for i, q in enumerate(qs):
            if not np.array_equal(np.asarray(h.expand_bits(q)), want_bits[i][:rows]):
                bad["expand"] += 1
                print("rep %d: expanded bits of query %d differ" % (rep, i), flush=True)
            if not np.array_equal(np.asarray(h.answer_with_bits(W, want_bits[i][:rows])).reshape(-1), want[i]):
                bad["scan"] += 1
                print("rep %d: scan of the oracle's bits of query %d differs" % (rep, i), flush=True)
print("flaky_bits_probe: %d reps, mismatches %s" % (reps, bad), flush=True)
