"""Import shim: `import proteinbert_b200` -> the package in ./proteinbert-pytorch-replication_b200/
(whose directory name is not a valid python identifier)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
sys.modules[__name__] = importlib.import_module("proteinbert-pytorch-replication_b200")
