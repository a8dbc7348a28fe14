"""foremast_1_b200 — B200-native scoring path of Foremast's brain.

Host-side mirror of the brain's interface (env knobs, config strings, document
statuses) over the C ABI of libforemast_cuda.so.  See DESIGN.md.
"""
from . import _abi  # noqa: F401

__all__ = ["_abi"]
