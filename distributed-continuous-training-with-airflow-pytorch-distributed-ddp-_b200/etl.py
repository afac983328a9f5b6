"""GPU version of the reference's ETL arithmetic (jobs/preprocess.py:15-51): host reads the CSV into Arrow
columns, the device does label encode + statistics + z-score (K1 kernels), host writes the parquet directory."""
from __future__ import annotations

import os

import numpy as np
import torch

from .ops import Context

FEATURE_COLS = ["Temperature", "Humidity", "Wind_Speed", "Cloud_Cover", "Pressure"]      # preprocess.py:29


def arrow_string_buffers(arr):
    """pyarrow StringArray (possibly chunked) -> (offsets int32, data uint8, validity uint8|None) numpy views."""
    import pyarrow as pa

    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    arr = arr.cast(pa.string())
    n = len(arr)
    validity, offsets, data = arr.buffers()
    off = np.frombuffer(offsets, dtype=np.int32, count=n + 1 + arr.offset)[arr.offset:]
    dat = np.frombuffer(data, dtype=np.uint8) if data is not None else np.zeros(1, np.uint8)
    val = None
    if validity is not None and arr.null_count > 0:
        if arr.offset % 8 != 0:
            bits = np.unpackbits(np.frombuffer(validity, dtype=np.uint8), bitorder="little")[arr.offset:arr.offset + n]
            val = np.packbits(bits, bitorder="little")
        else:
            val = np.frombuffer(validity, dtype=np.uint8)[arr.offset // 8:]
    return off.copy(), dat.copy() if dat.size else np.zeros(1, np.uint8), None if val is None else val.copy()


def preprocess_table(table, feature_cols=None, device=0, label_col: str = "Rain"):
    """pyarrow Table -> dict(norm float64 [F,N] numpy with NaN = null, labels32 int32 numpy, stats [F,4])."""
    import pyarrow as pa
    import pyarrow.compute as pc

    feature_cols = feature_cols or FEATURE_COLS
    f = len(feature_cols)
    ctx = Context(f, device=device)
    dev = ctx.device
    cols = []
    for name in feature_cols:
        col = table[name]
        if isinstance(col, pa.ChunkedArray):
            col = col.combine_chunks()
        col = pc.cast(col, pa.float64())
        v = col.to_numpy(zero_copy_only=False)                # nulls -> NaN
        cols.append(torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64)).to(dev))
    stats = ctx.col_stats(cols)                               # preprocess.py:32-36
    off, dat, val = arrow_string_buffers(table[label_col])
    labels64, labels32 = ctx.label_encode(torch.from_numpy(off).to(dev), torch.from_numpy(dat).to(dev),
                                          None if val is None else torch.from_numpy(val).to(dev), want_int32=True)  # :23-25
    # the parquet contract is float64 `*_norm`: (x-mean)/std in float64 is a host-side one-liner on the stats the
    # GPU produced; the fp32 training layout comes from wx_normalize_pack when the trainer ingests the table
    st = stats.cpu().numpy()
    norm = np.empty((f, len(table)), dtype=np.float64)
    for j, c in enumerate(cols):
        norm[j] = (c.cpu().numpy() - st[j, 1]) / st[j, 3]     # preprocess.py:38-41
    feats, nullmask = ctx.normalize_pack(cols, stats)
    return {"norm": norm, "labels32": labels32.cpu().numpy(), "labels": labels64, "stats": st,
            "features": feats, "nullmask": nullmask, "ctx": ctx}


def write_parquet_dir(out_dir: str, feature_cols, norm: np.ndarray, labels32: np.ndarray) -> str:
    """`final_df.write.mode("overwrite").parquet(output_path)` (preprocess.py:44-51): a directory with part files."""
    import pyarrow as pa
    import pyarrow.parquet as pq

    os.makedirs(out_dir, exist_ok=True)
    for name in os.listdir(out_dir):                          # mode("overwrite")
        if name.startswith("part-") or name == "_SUCCESS":
            os.remove(os.path.join(out_dir, name))
    cols = {f"{n}_norm": pa.array(norm[j], type=pa.float64(), mask=np.isnan(norm[j])) for j, n in enumerate(feature_cols)}
    cols["label_encoded"] = pa.array(labels32, type=pa.int32())
    pq.write_table(pa.table(cols), os.path.join(out_dir, "part-00000-wxddp.snappy.parquet"), compression="snappy")
    open(os.path.join(out_dir, "_SUCCESS"), "w").close()
    return out_dir


def main() -> int:
    import pyarrow.csv as pacsv

    input_path = os.environ.get("WX_RAW_CSV", "/opt/spark/data/raw/weather.csv")            # preprocess.py:15
    out_root = os.environ.get("WX_PROCESSED_DIR", "/opt/spark/data/processed")
    print("=" * 80)
    print("Step 1: Weather Data Preprocessing (wxddp, B200)")
    print("=" * 80)
    print(f"Reading data from: {input_path}")
    table = pacsv.read_csv(input_path)                                                      # header + inferred types (:18)
    print("Encoding target variable 'Rain'...")
    print("Normalizing features...")
    res = preprocess_table(table)
    output_path = os.path.join(out_root, "data.parquet")                                    # :44
    print(f"Saving processed data to: {output_path}")
    write_parquet_dir(output_path, FEATURE_COLS, res["norm"], res["labels32"])
    print("✓ Preprocessing Complete!")
    return 0
