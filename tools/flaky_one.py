"""One process start of tools/flaky_loop.sh: fresh handles over padded (100-byte) slots, both batch paths and single
queries against the oracle's bytes, 6 repetitions.  Prints 'start <k>: ok' or the mismatches."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
from tests.helpers import Handle, make_queries  # noqa: E402

k = int(sys.argv[1])
bad = 0
for db_size, S, W, nq in ((3000, 100, 3, 4), (4097, 50, 1, 5)):
    data = oracle.synth_fill(21, 0, db_size, S)
    rows = db_size // W
    rng = np.random.default_rng(nq + k)
    pairs = [make_queries(100 + i + 10 * k, db_size, W, int(idx)) for i, idx in enumerate(rng.integers(0, rows, size=nq))]
    qs = [p[1][i & 1] for i, p in enumerate(pairs)]
    want = np.stack([np.asarray(oracle.answer(p[0][i & 1], data, db_size, S)).reshape(-1) for i, p in enumerate(pairs)])
    for rep in range(6):
        with Handle(db_size, S) as h:
            h.fill(21)
            for mode in (2, 1):
                h.set_option("batch_mode", mode)
                got = np.asarray(h.answer_batch(qs)).reshape(nq, -1)
                if not np.array_equal(got, want):
                    bad += 1
                    print("start %d: MISMATCH shape %s rep %d mode %d" % (k, (db_size, S, W), rep, mode))
            if not np.array_equal(h.answer(qs[0]), want[0]):
                bad += 1
                print("start %d: MISMATCH single query shape %s rep %d" % (k, (db_size, S, W), rep))
print("start %d: %s" % (k, "ok" if bad == 0 else "FAILED (%d mismatches)" % bad))
sys.exit(1 if bad else 0)
