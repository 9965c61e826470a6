#!/usr/bin/env python
"""compare_pfm.py a.pfm b.pfm -> one JSON line: identical bits? relative RMSE (over finite pixels)."""
import json, sys
import numpy as np


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data


a, b = read_pfm(sys.argv[1]), read_pfm(sys.argv[2])
ok = np.isfinite(a).all(-1) & np.isfinite(b).all(-1)
x, y = a[ok].astype(np.float64), b[ok].astype(np.float64)
rr = float(np.sqrt(np.mean((x - y) ** 2)) / max(np.sqrt(np.mean(y ** 2)), 1e-30))
print(json.dumps({"a": sys.argv[1], "b": sys.argv[2], "shape": list(a.shape), "bit_identical": bool(a.tobytes() == b.tobytes()),
                  "words_equal": float(np.mean(a.view(np.uint32) == b.view(np.uint32))), "rel_rmse": rr, "finite": float(ok.mean())}))
