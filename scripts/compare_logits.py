#!/usr/bin/env python
"""Compare two logits dumps of tests/harness/llama_logits.cpp (CPU backend only vs the B200 plugin loaded)."""
import sys

import numpy as np

a = np.fromfile(sys.argv[1], dtype=np.float32)
b = np.fromfile(sys.argv[2], dtype=np.float32)
assert a.size == b.size and a.size > 0, (a.size, b.size)
same = a.view(np.uint32) == b.view(np.uint32)
diff = np.abs(a.astype(np.float64) - b.astype(np.float64))
scale = max(float(np.abs(a).max()), 1e-30)
print(f"logits: {a.size} values, bit-identical {same.mean() * 100:.4f} %, max |diff| {diff.max():.3e} "
      f"(max |logit| {scale:.3e}, relative {diff.max() / scale:.3e}), argmax agreement "
      f"{(a.reshape(-1, 32000).argmax(1) == b.reshape(-1, 32000).argmax(1)).mean() * 100:.1f} %")
sys.exit(0 if diff.max() / scale < 1e-5 else 1)
