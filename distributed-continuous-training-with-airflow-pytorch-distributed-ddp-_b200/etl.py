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


def preprocess_table(table, feature_cols=None, device=0, label_col: str = "Rain", want_features: bool = True):
    """pyarrow Table -> dict(norm float64 [F,N] numpy (what Spark writes: nulls stay null), validity bitmaps, labels32
    int32 numpy, stats [F,4]).  The Arrow value buffers and validity bitmaps go to the device as they are (pinned,
    chunked, on a copy stream); statistics (preprocess.py:32-36), the float64 z-score (:38-41) and the label compare
    (:23-25) are kernels -- the host computes nothing on the values."""
    import pyarrow as pa
    import pyarrow.compute as pc

    from .ingest import ColumnUploader, _arrow_column_buffers, _bitmap_slice

    feature_cols = feature_cols or FEATURE_COLS
    f = len(feature_cols)
    n = len(table)
    ctx = Context(f, device=device)
    dev = ctx.device
    up = ColumnUploader(dev)
    cols, validity, host_validity = [], [], []
    for name in feature_cols:
        col = table[name]
        if col.type != pa.float64():
            col = pc.cast(col, pa.float64())                  # inferSchema may have typed an all-integer column as long
        vals, bits, boff, nulls = _arrow_column_buffers(col, np.float64)
        d = torch.empty(n, dtype=torch.float64, device=dev)
        up.put(d, vals)
        cols.append(d)
        if nulls:
            hb = np.ascontiguousarray(_bitmap_slice(bits, boff, n))
            dv = torch.empty(hb.size, dtype=torch.uint8, device=dev)
            up.put(dv, hb)
            validity.append(dv)
            host_validity.append(hb)
        else:
            validity.append(None)
            host_validity.append(None)
    off, dat, val = arrow_string_buffers(table[label_col])
    d_off = torch.empty(off.size, dtype=torch.int32, device=dev)
    d_dat = torch.empty(dat.size, dtype=torch.uint8, device=dev)
    up.put(d_off, off)
    up.put(d_dat, dat)
    d_val = None
    if val is not None:
        d_val = torch.empty(val.size, dtype=torch.uint8, device=dev)
        up.put(d_val, val)
    up.wait()
    vlist = validity if any(v is not None for v in validity) else None
    stats = ctx.col_stats(cols, validity=vlist)               # preprocess.py:32-36
    labels64, labels32 = ctx.label_encode(d_off, d_dat, d_val, want_int32=True)            # :23-25
    norm_d = ctx.zscore_f64(cols, stats, validity=vlist)      # :38-41, float64 as Spark writes it
    norm = torch.empty(f, n, dtype=torch.float64).pin_memory()
    norm.copy_(norm_d, non_blocking=True)
    lab = torch.empty(n, dtype=torch.int32).pin_memory()
    lab.copy_(labels32, non_blocking=True)
    out = {"labels": labels64, "stats": None, "ctx": ctx, "validity": host_validity}
    if want_features:
        feats, nullmask = ctx.normalize_pack(cols, stats, validity=vlist)
        out["features"], out["nullmask"] = feats, nullmask
    torch.cuda.current_stream(dev).synchronize()
    out["norm"], out["labels32"], out["stats"] = norm.numpy(), lab.numpy(), stats.cpu().numpy()
    return out


def write_parquet_dir(out_dir: str, feature_cols, norm: np.ndarray, labels32: np.ndarray, validity=None) -> str:
    """`final_df.write.mode("overwrite").parquet(output_path)` (preprocess.py:44-51): a directory with part files.
    validity: per column an Arrow validity bitmap (uint8, LSB first) or None; it becomes the column's null bitmap as it
    is.  Without bitmaps NaN cells are written as nulls (Spark writes SQL null for them)."""
    import pyarrow as pa
    import pyarrow.parquet as pq

    os.makedirs(out_dir, exist_ok=True)
    for name in os.listdir(out_dir):                          # mode("overwrite")
        if name.startswith("part-") or name == "_SUCCESS":
            os.remove(os.path.join(out_dir, name))
    n = norm.shape[1]
    cols = {}
    for j, name in enumerate(feature_cols):
        v = np.ascontiguousarray(norm[j])
        vb = None if validity is None else validity[j]
        if vb is not None:
            cols[f"{name}_norm"] = pa.Array.from_buffers(pa.float64(), n, [pa.py_buffer(vb), pa.py_buffer(v)])
        elif validity is not None and not np.isnan(v[:1]).any() and not np.isnan(v[-1:]).any():
            cols[f"{name}_norm"] = pa.Array.from_buffers(pa.float64(), n, [None, pa.py_buffer(v)])
        else:
            cols[f"{name}_norm"] = pa.array(v, type=pa.float64(), mask=np.isnan(v))
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
    res = preprocess_table(table, want_features=False)
    output_path = os.path.join(out_root, "data.parquet")                                    # :44
    print(f"Saving processed data to: {output_path}")
    write_parquet_dir(output_path, FEATURE_COLS, res["norm"], res["labels32"], validity=res["validity"])
    print("✓ Preprocessing Complete!")
    return 0
