"""Synthetic Euromillions-shaped inputs.

There is no network for scraping (the reference's acquisition stage, Main.java:37-67), so the
inputs follow the reference's row schema — a draw is a date + 5 mains (1..50) + 2 stars (1..12),
Main.java:69,86-102 — generated from an index-addressable splitmix64 so that C++, CUDA, Java and
Python produce identical data.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(seed, idx):
    """z = seed + 0x9E3779B97F4A7C15*(idx+1); two xorshift-multiply rounds; u64 wraparound."""
    idx = np.asarray(idx, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * (idx + np.uint64(1))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def flat(n_rows, n_features, n_classes=50, seed=42, row_start=0, chunk=1 << 18):
    """Flat mode (cfg2/cfg3/cfg5): x[i][j] = 1 + splitmix64(seed, i*F+j) mod 50 (as f64);
    label[i] = (x[i][0] + x[i][3] + splitmix64(seed ^ 0xA5A5A5A5, i) mod 5) mod C."""
    X = np.empty((n_rows, n_features), np.float64)
    y = np.empty(n_rows, np.float64)
    F = np.uint64(n_features)
    cols = np.arange(n_features, dtype=np.uint64)
    for a in range(0, n_rows, chunk):
        b = min(n_rows, a + chunk)
        rows = np.arange(row_start + a, row_start + b, dtype=np.uint64)
        idx = rows[:, None] * F + cols[None, :]
        v = (splitmix64(seed, idx) % np.uint64(50)).astype(np.int64) + 1
        X[a:b] = v
        e = (splitmix64(seed ^ 0xA5A5A5A5, rows) % np.uint64(5)).astype(np.int64)
        c3 = v[:, 3] if n_features > 3 else 0
        y[a:b] = (v[:, 0] + c3 + e) % n_classes
    return X, y


def draws(n_draws=1500, seed=20200614):
    """Draw mode: ball vector b[d][0..6] = 5 sorted distinct mains in 1..50, 2 sorted distinct stars in 1..12."""
    out = np.zeros((n_draws, 7), np.int64)
    for d in range(n_draws):
        pool = list(range(1, 51))
        for j in range(5):
            k = j + int(splitmix64(seed, 16 * d + j)) % (50 - j)
            pool[j], pool[k] = pool[k], pool[j]
        out[d, :5] = sorted(pool[:5])
        pool = list(range(1, 13))
        for j in range(2):
            k = j + int(splitmix64(seed, 16 * d + 8 + j)) % (12 - j)
            pool[j], pool[k] = pool[k], pool[j]
        out[d, 5:] = sorted(pool[:2])
    return out


def draw_history(n_draws=1500, lag=10, ball=0, seed=20200614):
    """cfg1/cfg4: row t in [lag, n_draws): x[(l-1)*7+q] = b[t-l][q], l=1..lag; label = b[t][ball]-1.
    Returns X (N x 7*lag f64), y (N f64), n_classes (50 for mains, 12 for stars)."""
    b = draws(n_draws, seed)
    n = n_draws - lag
    X = np.zeros((n, 7 * lag), np.float64)
    for l in range(1, lag + 1):
        X[:, (l - 1) * 7:(l - 1) * 7 + 7] = b[lag - l:n_draws - l]
    y = (b[lag:, ball] - 1).astype(np.float64)
    return X, y, (50 if ball < 5 else 12)


class DeviceFlat:
    """flat() generated in HBM by the engine (emrf_synth_flat): device pointers of an n_rows x n_features float64 matrix
    and its labels, for shapes whose host copy would only exist to cross PCIe (BASELINE configs[2]: 10 GB)."""

    def __init__(self, ctx, n_rows, n_features, n_classes=50, seed=42, row_start=0):
        import ctypes as C
        from . import _lib
        L = _lib.lib()
        self.ctx, self.n_rows, self.n_features = ctx, int(n_rows), int(n_features)
        self.X, self.y = C.c_void_p(), C.c_void_p()
        _lib.check(L.emrf_dev_alloc(ctx.h, self.n_rows * self.n_features * 8, C.byref(self.X)))
        _lib.check(L.emrf_dev_alloc(ctx.h, self.n_rows * 8, C.byref(self.y)))
        _lib.check(L.emrf_synth_flat(ctx.h, self.n_rows, self.n_features, n_classes, seed, row_start, self.X, self.y))

    def to_host(self):
        import ctypes as C
        import numpy as np
        from . import _lib
        L = _lib.lib()
        X = np.empty((self.n_rows, self.n_features), np.float64)
        y = np.empty(self.n_rows, np.float64)
        _lib.check(L.emrf_memcpy_d2h(self.ctx.h, C.c_void_p(X.ctypes.data), self.X, X.nbytes))
        _lib.check(L.emrf_memcpy_d2h(self.ctx.h, C.c_void_p(y.ctypes.data), self.y, y.nbytes))
        return X, y

    def free(self):
        from . import _lib
        L = _lib.lib()
        for p in (self.X, self.y):
            if p and p.value:
                L.emrf_dev_free(self.ctx.h, p)
        self.X = self.y = None
