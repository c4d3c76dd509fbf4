"""Import alias: the package directory name contains hyphens, so `import nsof_b200 as nsof` is the
convenient spelling of importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")
sys.modules[__name__] = _pkg
