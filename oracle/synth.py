"""Seeded synthetic `weather.csv`-shaped tables (oracle / test infrastructure only).

The reference ships no data (`/root/reference/.gitignore:39` ignores `data/`;
`README.md:69` tells the user to supply `data/raw/weather.csv`).  The only facts
the code pins are the column names and types it touches:
  * five numeric feature columns `Temperature, Humidity, Wind_Speed, Cloud_Cover,
    Pressure` (`jobs/preprocess.py:29`), read by Spark with `inferSchema`
    (`:18`) -> DoubleType, nullable;
  * one string column `Rain`, compared with the literal "rain" (`:23-25`).
Everything else here (distributions, null rate, the label rule) is ours and is
chosen so the loss actually decreases and every guard in the path is exercised.
"""
from __future__ import annotations

import numpy as np

FEATURE_NAMES_5 = ["Temperature", "Humidity", "Wind_Speed", "Cloud_Cover", "Pressure"]
_MU = [20.0, 65.0, 10.0, 50.0, 1013.0]
_SIGMA = [8.0, 20.0, 6.0, 30.0, 7.0]


def feature_names(f: int) -> list[str]:
    """F=5 -> the reference's exact names (preprocess.py:29); wider -> extra synthetic sensors."""
    names = list(FEATURE_NAMES_5[: min(f, 5)])
    names += [f"Sensor_{j:02d}" for j in range(5, f)]
    return names


def synth_table(seed: int, n: int, f: int, null_rate: float = 0.0,
                const_col: int | None = None, odd_label_rate: float = 0.001) -> dict:
    """Return a raw table: columnar float64 [F, N] (NaN = null) + a list of label strings.

    Values are rounded to 2 decimals like a CSV field; the float32 view of the same
    table is `raw.astype(np.float32)` (used by the fp32-input kernel variant).
    Labels: "rain"/"no rain" from a noisy linear rule on the first three columns,
    plus ~0.1% of None / "Rain" / "RAIN" / "rain " / "" to pin the `otherwise(0)`
    branch of preprocess.py:23-25 (case-sensitive exact match, null -> 0).
    """
    rng = np.random.default_rng(seed)
    raw = np.empty((f, n), dtype=np.float64)
    for j in range(f):
        mu = _MU[j % 5] + 3.0 * (j // 5)
        sg = _SIGMA[j % 5] * (1.0 + 0.25 * (j // 5))
        raw[j] = np.round(rng.normal(mu, sg, size=n), 2)
    if const_col is not None:
        raw[const_col] = 42.5                                   # std == 0 -> guard at preprocess.py:36
    z = [(raw[j] - _MU[j]) / _SIGMA[j] for j in range(min(3, f))]
    score = -1.6 + 0.9 * z[1 % len(z)] - 0.7 * z[0] + 0.5 * z[2 % len(z)] + rng.normal(0, 0.6, size=n)
    is_rain = score > 0.0
    labels = np.where(is_rain, "rain", "no rain").astype(object)
    if odd_label_rate > 0:
        odd = rng.random(n) < odd_label_rate
        pool = np.array([None, "Rain", "RAIN", "rain ", "", " rain", "rainy", "rai"], dtype=object)
        labels[odd] = pool[rng.integers(0, len(pool), size=int(odd.sum()))]
    if null_rate > 0:
        nulls = rng.random((f, n)) < null_rate
        if const_col is not None:
            nulls[const_col] = False
        raw[nulls] = np.nan
    return {"names": feature_names(f), "raw": raw, "labels": list(labels), "seed": seed}


def to_arrow_strings(strings) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Arrow `utf8` layout of a string column: (offsets int32 [N+1], data uint8, validity bits).

    Validity is LSB-first, 1 = valid -- the layout Spark/pyarrow hand to a columnar consumer.
    """
    n = len(strings)
    offsets = np.zeros(n + 1, dtype=np.int32)
    chunks = []
    valid = np.zeros(n, dtype=bool)
    pos = 0
    for i, s in enumerate(strings):
        if s is not None:
            b = s.encode("utf-8")
            chunks.append(b)
            pos += len(b)
            valid[i] = True
        offsets[i + 1] = pos
    data = np.frombuffer(b"".join(chunks), dtype=np.uint8).copy() if pos else np.zeros(0, np.uint8)
    return offsets, data, np.packbits(valid, bitorder="little")
