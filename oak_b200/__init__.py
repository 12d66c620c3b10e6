"""oak_b200 — B200-native batched Generation I battle engine for the hot path of lab-oak/oak.

The product is the CUDA library `oak_b200/liboakb200.so` behind the C ABI in include/oakb200.h; this
package is the thin Python host side (ctypes + torch for device memory and streams). There is no CPU
implementation here: importing works anywhere, but creating a `BattleEngine` without a CUDA device or
without the built library raises.
"""
from .engine import BattleEngine, OakB200Error, load_library  # noqa: F401

__all__ = ["BattleEngine", "OakB200Error", "load_library"]
