"""import-only stand-in for the `data` module src/CGNN.py imports (`get_dataset`, `set_train_val_test_split`): the
fork does not ship a src/data.py at all (upstream's downloads Planetoid / OGB datasets, unavailable offline).
TEST INFRASTRUCTURE ONLY (oracle/make_golden.py)."""


def get_dataset(*a, **k):
    raise RuntimeError('data.get_dataset is a stub (oracle/ref_stubs)')


def set_train_val_test_split(*a, **k):
    raise RuntimeError('data.set_train_val_test_split is a stub (oracle/ref_stubs)')
