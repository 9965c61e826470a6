"""Shared helpers for the parity tests: seeded inputs, bit-exact comparison that treats NaNs as equal."""
import numpy as np

import mrt_b200 as m


def bits_equal(a, b):
    """Bit-for-bit equality of two arrays, except that any NaN matches any NaN (x86 and CUDA give NaNs different
    payload/sign bits for the same invalid operation)."""
    a = np.ascontiguousarray(a); b = np.ascontiguousarray(b)
    if a.dtype.fields:
        return all(bits_equal(a[name], b[name]) for name in a.dtype.names if not name.startswith("_"))
    if a.dtype.kind == "f":
        an, bn = np.isnan(a), np.isnan(b)
        if not np.array_equal(an, bn):
            return False
        return np.array_equal(a.view(np.uint32)[~an], b.view(np.uint32)[~bn])
    return np.array_equal(a, b)


def mismatch_report(a, b, limit=5):
    out = []
    for name in a.dtype.names or [None]:
        x = a[name] if name else a
        y = b[name] if name else b
        if name and name.startswith("_"):
            continue
        if not bits_equal(x, y):
            bad = np.argwhere(~((x == y) | (np.isnan(x) & np.isnan(y)) if x.dtype.kind == "f" else (x == y)))
            out.append(f"{name}: {len(bad)} mismatches, first at {bad[:limit].tolist()}: {x[tuple(bad[0])]} vs {y[tuple(bad[0])]}")
    return "; ".join(out)


def setup_pair(scene, w, h, seed=1, max_path_length=8, renderer_kwargs=None, oracle_threads=0):
    """A GPU Renderer and a CPU Oracle on the same scene, size and initial noise slots."""
    import oracle as orc
    r = m.Renderer(max_path_length=max_path_length, **(renderer_kwargs or {}))
    o = orc.Oracle(max_path_length=max_path_length, threads=oracle_threads)
    r.init_raytracing(scene); o.init_raytracing(scene)
    r.drawable_size_will_change(w, h); o.drawable_size_will_change(w, h)
    init = m.noise_tile(seed, 0xFFFFFFFF)           # all three slots start identical (Renderer.mm:115-126)
    for s in range(3):
        r.update_noise(s, init); o.update_noise(s, init)
    return r, o


def random_rays(n, seed, dtype=None):
    """Rays from inside/around the Cornell volume in random directions, as 32-byte records (or RAY-sized)."""
    rng = np.random.default_rng(seed)
    dtype = dtype or np.dtype([("origin", "<f4", 3), ("minDistance", "<f4"), ("direction", "<f4", 3), ("maxDistance", "<f4")])
    rays = np.zeros(n, dtype=dtype)
    rays["origin"] = rng.uniform([-0.99, 0.01, -0.99], [0.99, 1.99, 2.3], size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["direction"] = d.astype(np.float32)
    rays["minDistance"] = 0.0
    rays["maxDistance"] = np.inf
    return rays
