"""pir_b200 — B200-native server-side answer path of the two-server DPF PIR of mvmcconnell/pir.

Layout: csrc/ (CUDA kernels + the C ABI of include/pir_b200.h), lib/ (the built .so), and the host
mirror of the reference's Go API for this path: `pir_b200.pir` (db.go / slot.go / query.go) and
`pir_b200.dpf` (dpf/*.go), `pir_b200.aspir` (aspir.go:227-295), `pir_b200.keyword` (keyword.go).  No CPU fallback exists.
"""
from . import _capi  # noqa: F401
from . import dpf, pir  # noqa: F401
from . import aspir  # noqa: F401,E402
from . import keyword  # noqa: F401,E402
