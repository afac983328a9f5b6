"""numpy restatement of `/root/reference/jobs/preprocess.py:23-41` (oracle, test infrastructure only).

PARITY UNPINNED for the statistics: the reference computes `mean`/`stddev` inside
Spark 3.4.1 (`preprocess.py:33`; docker-compose.yml:57 pins the image), which cannot
run here (no pyspark, no JVM) and the reference holds no golden vectors.  What is
restated is Spark's documented semantics for those aggregates: nulls are skipped,
arithmetic is float64, `stddev` is `stddev_samp` (n-1 denominator) and is null when
fewer than two non-null rows exist.  The element-wise part (`:23-25`, `:36`, `:38-41`)
is plain IEEE float64 arithmetic and is restated exactly.
"""
from __future__ import annotations

import os

import numpy as np

IMPUTE_NONE = 0   # reference behaviour: null -> null -> NaN in the trainer (train_lightning_ddp.py:31,45)
IMPUTE_MEAN = 1   # extension: null -> column mean == 0.0 in normalised space


def label_encode(strings) -> np.ndarray:
    """`when(col("Rain") == "rain", 1).otherwise(0)` -- preprocess.py:23-25.

    Case-sensitive exact match; null and every other string -> 0.  Spark writes the
    literal ints as int32 (the trainer widens to int64 at train_lightning_ddp.py:46).
    """
    return np.fromiter((1 if s == "rain" else 0 for s in strings), dtype=np.int32, count=len(strings))


def column_stats(col: np.ndarray) -> tuple[int, float, float | None]:
    """`df.select(mean(c), stddev(c)).first()` -- preprocess.py:33.

    Returns (non-null count, mean, sample stddev) in float64; stddev is None when
    count < 2 (Spark returns null).  Two-pass formula; the CUDA kernel uses a Chan
    merge of per-tile (count, mean, M2), so the two agree to a few float64 ulps only.
    """
    v = np.asarray(col, dtype=np.float64)
    v = v[~np.isnan(v)]
    n = int(v.size)
    if n == 0:
        return 0, float("nan"), None
    mean = float(np.mean(v, dtype=np.float64))
    if n < 2:
        return n, mean, None
    m2 = float(np.sum((v - mean) ** 2, dtype=np.float64))
    return n, mean, float(np.sqrt(m2 / (n - 1)))


def std_guard(std: float | None) -> float | None:
    """`std_val = stats['std'] if stats['std'] != 0 else 1.0` -- preprocess.py:36 (None passes through)."""
    if std is None:
        return None
    return std if std != 0 else 1.0


def zscore(col: np.ndarray, mean: float, std_val: float | None) -> np.ndarray:
    """`(col - mean) / std_val` in float64, null-propagating -- preprocess.py:38-41."""
    v = np.asarray(col, dtype=np.float64)
    if std_val is None:                       # col / null -> null column (unguarded edge)
        return np.full_like(v, np.nan)
    return (v - mean) / std_val


def preprocess_table(raw: np.ndarray, label_strings, impute: int = IMPUTE_NONE,
                     stats: list[tuple[float, float]] | None = None) -> dict:
    """Whole `preprocess.py:23-41` + the trainer's casts (`train_lightning_ddp.py:45-46`).

    raw: float64 [F, N] columnar, NaN = null.  `stats` (list of (mean, std_val)) may be
    injected so the element-wise path can be compared bit-for-bit with given statistics.
    Returns float64 `norm` [F, N] (what Spark writes), float32 row-major `features` [N, F]
    and int64 `labels` [N] (what WeatherDataset holds), the per-row null mask (bit j set
    <=> feature j was null) and, for IMPUTE_MEAN, the ascending flat indices row*F+col of
    the imputed cells.
    """
    f, n = raw.shape
    assert f <= 32
    out_stats = []
    norm = np.empty((f, n), dtype=np.float64)
    for j in range(f):
        if stats is None:
            cnt, mean, std = column_stats(raw[j])
            std_val = std_guard(std)
        else:
            mean, std_val = stats[j]
            cnt = int((~np.isnan(raw[j])).sum())
        out_stats.append((cnt, mean, std_val))
        norm[j] = zscore(raw[j], mean, std_val)
    isnull = np.isnan(raw)                                   # [F, N]
    nullmask = np.zeros(n, dtype=np.uint32)
    for j in range(f):
        nullmask |= (isnull[j].astype(np.uint32) << np.uint32(j))
    features = norm.T.astype(np.float32)                      # FloatTensor(float64 ndarray) at :45
    imputed_idx = np.zeros(0, dtype=np.int64)
    if impute == IMPUTE_MEAN:
        features[isnull.T] = 0.0
        imputed_idx = np.flatnonzero(isnull.T.reshape(-1)).astype(np.int64)
    labels32 = label_encode(label_strings)
    return {"stats": out_stats, "norm": norm, "features": np.ascontiguousarray(features),
            "labels": labels32.astype(np.int64), "labels32": labels32,
            "nullmask": nullmask, "imputed_idx": imputed_idx}


def write_parquet_dir(data_dir: str, names: list[str], norm: np.ndarray, labels32: np.ndarray,
                      n_parts: int = 2) -> str:
    """Lay out `data/processed/data.parquet/part-*.parquet` the way preprocess.py:44-51 does.

    A DIRECTORY of part files + `_SUCCESS`, float64 `<name>_norm` columns (NaN stored as
    parquet null, as Spark does for SQL null) and int32 `label_encoded`.
    """
    import pyarrow as pa
    import pyarrow.parquet as pq

    out = os.path.join(data_dir, "data.parquet")
    os.makedirs(out, exist_ok=True)
    n = norm.shape[1]
    bounds = np.linspace(0, n, n_parts + 1).astype(int)
    for p in range(n_parts):
        lo, hi = bounds[p], bounds[p + 1]
        cols = {}
        for j, name in enumerate(names):
            v = norm[j, lo:hi]
            cols[f"{name}_norm"] = pa.array(v, type=pa.float64(), mask=np.isnan(v))
        cols["label_encoded"] = pa.array(labels32[lo:hi], type=pa.int32())
        pq.write_table(pa.table(cols), os.path.join(out, f"part-{p:05d}-synthetic.snappy.parquet"),
                       compression="snappy")
    open(os.path.join(out, "_SUCCESS"), "w").close()
    return out
