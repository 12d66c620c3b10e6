#!/usr/bin/env python3
"""Golden outputs of THE REFERENCE'S OWN int8 main-net layers (nn/battle/quantized/{affine,clipped_relu}.h compiled in
place with -mavx2 into oracle/_ref/libref_oak.so by oracle/ref_shim.cpp: ref_quantized_main) for a seeded clamp network and
256 random u8 battle embeddings. Writes tests/golden/nn_discrete_golden.npz. Run in the build container only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oak_ref import RefQuantizedMain, make_net_blob, ref_library  # noqa: E402

ref = ref_library()
assert ref is not None, "build oracle/_ref first (make -C oracle)"
rng = np.random.default_rng(2025)
blob = make_net_blob(rng, activation=2, scale=4.0)
N = 256
q = rng.integers(0, 128, (N, 768)).astype(np.uint8)
q[rng.random((N, 768)) < 0.5] = 0
idx1 = rng.integers(0, 315, (N, 9)).astype(np.uint16)
idx2 = rng.integers(0, 315, (N, 9)).astype(np.uint16)
value, l1, l2 = np.zeros(N, np.float32), np.zeros((N, 9), np.float32), np.zeros((N, 9), np.float32)
rq = RefQuantizedMain(ref, blob)
for i in range(N):
    value[i], l1[i], l2[i] = rq(q[i], idx1[i], idx2[i])
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "nn_discrete_golden.npz"), blob=blob, q=q, idx1=idx1, idx2=idx2, value=value,
                    l1=l1, l2=l2)
print("wrote", N, "cases; value range", float(value.min()), float(value.max()))
