"""import-only stand-in for `ray` (not installable offline): src/CGNN.py imports `tune` and `CLIReporter` at module
level for its hyper-parameter search script; the model classes the goldens exercise never touch them.
TEST INFRASTRUCTURE ONLY (oracle/make_golden.py)."""
from . import tune  # noqa: F401
