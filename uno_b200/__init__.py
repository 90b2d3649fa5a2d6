"""uno_b200 — B200-native sparse symmetric-indefinite KKT solver behind Uno's
DirectSymmetricIndefiniteLinearSolver interface.  The product is uno_b200/lib/libb200kkt.so
(C ABI in include/b200kkt.h); this package only holds the build script, the ctypes binding used
by tests/bench, and the synthetic KKT generators."""
from .capi import B200KKTError, KKTSolver, load_library  # noqa: F401
