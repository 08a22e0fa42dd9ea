import os
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


import pytest  # noqa: E402


@pytest.fixture
def emulated_abi(monkeypatch):
    """run the product's Python host side against tests/abi_emulator.py (host-memory restatement of the C ABI)
    instead of libgrandb200.so: plumbing tests without a GPU.  Test infrastructure only - see that module."""
    import contextlib
    import torch
    import grand_b200  # noqa: F401
    from grand_b200 import _lib, plan
    from abi_emulator import Emulator
    emu = Emulator()
    monkeypatch.setattr(_lib, 'load', lambda: emu)
    monkeypatch.setattr(_lib, 'require_cuda', lambda *a: None)
    monkeypatch.setattr(_lib, 'stream', lambda: 0)
    monkeypatch.setattr(torch.cuda, 'device', lambda dev: contextlib.nullcontext())
    plan._PLAN_CACHE.clear()
    yield emu
    plan._PLAN_CACHE.clear()
