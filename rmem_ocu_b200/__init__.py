"""rmem_ocu_b200 -- B200-native (sm_100a) implementation of RMem's per-frame memory-reading hot path
(AOT / DeAOT propagation with a restricted long-term memory bank), behind the reference's engine API.
"""
from ._lib import RmemError, build  # noqa: F401

__all__ = ["RmemError", "build"]
