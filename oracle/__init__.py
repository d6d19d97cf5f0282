"""CPU oracle for the DPF-PIR answer path — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  Nothing under pir_b200/ does.  See oracle/pir_oracle.h for the parity status.
"""
from .loader import (  # noqa: F401
    ORACLE_DIR,
    PirQuery,
    build,
    lib,
    gen_two_server,
    evaluate2p,
    expand_shared_query,
    expand_tree,
    answer_with_bits,
    answer,
    answers_over_synth,
    synth_fill,
    tape,
    num_bits_for,
    query_from_bytes,
)
