"""Import helper: the package directory name contains hyphens (mandated layout), so it cannot be
named in an ``import`` statement.  ``from lgattn_pkg import pkg`` gives the package module."""
import importlib
import sys
from pathlib import Path

_ROOT = Path(__file__).resolve().parent
if str(_ROOT) not in sys.path:
    sys.path.insert(0, str(_ROOT))

PKG_NAME = "lightglue-with-flashattentionv2-tensorrt_b200"
pkg = importlib.import_module(PKG_NAME)
