"""see torch_geometric/__init__.py"""
