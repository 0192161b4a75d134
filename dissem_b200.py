"""Alias: `import dissem_b200` loads the package directory `distributed-llm-dissemination_b200/`
(whose name, fixed by the project layout, is not a Python identifier)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
sys.modules[__name__] = importlib.import_module("distributed-llm-dissemination_b200")
