"""Re-export of oracle/ref_shim.py (the import shim for the unmodified reference) for the golden generators."""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..", "..")))
from oracle.ref_shim import *  # noqa: F401,F403,E402
from oracle.ref_shim import REF_ROOT, available, load, load_operators  # noqa: F401,E402
