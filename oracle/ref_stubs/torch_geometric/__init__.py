"""Import-only stand-in (see torch_scatter stub): the reference subclasses MessagePassing purely for nn.Module-ness (`src/base_classes.py:3,116,137`)."""
