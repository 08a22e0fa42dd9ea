"""Import-only stand-in: `MessagePassing` is used by the reference as a plain nn.Module base
(`src/base_classes.py:116,137`; no `propagate` call exists anywhere in the reference)."""
from torch import nn


class MessagePassing(nn.Module):
    def __init__(self, *args, **kwargs):
        super().__init__()
