"""import-only stand-in for `torchdiffeq` (not installed in this image): lets src/CGNN.py be imported for its
ODEFunc / get_sym_adj; the goldens never integrate through it.  TEST INFRASTRUCTURE ONLY (oracle/make_golden.py)."""


def odeint(*a, **k):
    raise RuntimeError('torchdiffeq is a stub (oracle/ref_stubs): goldens cover single evaluations only')


odeint_adjoint = odeint
