"""Compare two Radiance .hdr files (linear radiance): used to check a multi-GPU render against the 1-GPU render of
the same samples.  Usage: compare_hdr.py a.hdr b.hdr -> one JSON line."""
import json
import sys

import cv2
import numpy as np

a = cv2.imread(sys.argv[1], cv2.IMREAD_UNCHANGED).astype(np.float64)
b = cv2.imread(sys.argv[2], cv2.IMREAD_UNCHANGED).astype(np.float64)
assert a.shape == b.shape, (a.shape, b.shape)
lit = (a.sum(axis=2) > 0) | (b.sum(axis=2) > 0)
rel = np.abs(a - b).sum(axis=2)[lit] / np.maximum(a.sum(axis=2)[lit], 1e-30)
print(json.dumps({"shape": list(a.shape), "lit_pixels": int(lit.sum()), "identical_pixels": int((a == b).all(axis=2).sum()),
                  "mean_rel_err": float(rel.mean()), "max_rel_err": float(rel.max()),
                  "mean_a": float(a.mean()), "mean_b": float(b.mean())}))
