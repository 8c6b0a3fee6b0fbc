"""Makes `gaussian_process_b200` importable from the drop-in modules wherever they are put on sys.path."""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
