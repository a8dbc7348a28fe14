"""Re-exports the host-batch builder of tests/golden/make_golden.py for other tests."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden import host_batch  # noqa: E402,F401
