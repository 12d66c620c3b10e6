"""oak_b200 — B200-native batched Generation I battle engine for the hot path of lab-oak/oak.

The product is the CUDA library `oak_b200/liboakb200.so` behind the C ABI in include/oakb200.h; this
package is the thin Python host side (ctypes + torch for device memory and streams). There is no CPU
implementation here: importing works anywhere, but creating a `BattleEngine` without a CUDA device or
without the built library raises.
"""
import os as _os

# hardware work queues per process: the tree search runs ~32 streams at once and the default of 8 makes them wait for each
# other (oakb200_create explains); must be in the environment before CUDA initialises, an explicit setting wins
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from .engine import BattleEngine, OakB200Error, load_library  # noqa: F401

__all__ = ["BattleEngine", "OakB200Error", "load_library"]
