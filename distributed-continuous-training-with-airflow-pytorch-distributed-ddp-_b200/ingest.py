"""Columnar ingest: Arrow buffers -> pinned staging -> HBM, chunk by chunk on a copy stream, with the Arrow validity
bitmaps handed to the K1 kernels as they are (no NaN rewrite, no arithmetic on the host).

Either side of the hot path the reference moves data as a DIRECTORY of snappy parquet part files with nullable float64
`*_norm` columns and an int32 `label_encoded` column (written at jobs/preprocess.py:44-51, read with pandas at
jobs/train_lightning_ddp.py:19,31,37,45-46).  Decoding parquet stays with pyarrow on the host cores; from the decoded
Arrow buffers on, every byte moves by memcpy / cudaMemcpyAsync and every value is computed on the GPU:

    part file -> pyarrow Table -> per column: data buffer (+ validity bitmap) -> pinned chunk -> H2D on the copy stream
              -> device columns [N] (+ bitmaps [ceil(N/8)])  -> wx_col_stats / wx_normalize_pack / wx_zscore_f64
"""
from __future__ import annotations

import glob
import os
import time

import numpy as np
import torch

from ._lib import IMPUTE_NONE, WxError
from .ops import Context


def _arrow_column_buffers(col, np_dtype):
    """pyarrow (Chunked)Array of a fixed-width type -> (values numpy view, validity bytes or None, bit offset, nulls)."""
    import pyarrow as pa

    if isinstance(col, pa.ChunkedArray):
        col = col.combine_chunks() if col.num_chunks != 1 else col.chunk(0)
    validity, data = col.buffers()
    n, off = len(col), col.offset
    itemsize = np.dtype(np_dtype).itemsize
    vals = np.frombuffer(data, dtype=np_dtype, count=n, offset=off * itemsize)
    if validity is None or col.null_count == 0:
        return vals, None, 0, 0
    return vals, np.frombuffer(validity, dtype=np.uint8), off, col.null_count


def _bitmap_slice(bits: np.ndarray, bit_offset: int, n: int) -> np.ndarray:
    """Arrow validity bits [bit_offset, bit_offset + n) as a byte-aligned bitmap (a byte copy when the offset allows)."""
    if bit_offset % 8 == 0:
        return bits[bit_offset // 8: bit_offset // 8 + (n + 7) // 8]
    b = np.unpackbits(bits[bit_offset // 8: (bit_offset + n + 7) // 8], bitorder="little")[bit_offset % 8: bit_offset % 8 + n]
    return np.packbits(b, bitorder="little")


class ColumnUploader:
    """Double-buffered pinned staging + a copy stream: `put(dst_device_view, src_numpy)` returns as soon as the bytes
    are in pinned memory; the H2D copy of chunk k overlaps the host memcpy of chunk k+1."""

    def __init__(self, device, chunk_bytes: int = 32 << 20):
        self.device = device
        self.chunk_bytes = chunk_bytes
        self.stream = torch.cuda.Stream(device=device)
        self.stage = [torch.empty(chunk_bytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
        self.free = [torch.cuda.Event() for _ in range(2)]
        self.k = 0
        self.bytes = 0

    def put(self, dst: torch.Tensor, src: np.ndarray) -> None:
        src = np.ascontiguousarray(src).view(np.uint8).reshape(-1)
        d8 = dst.view(torch.uint8).reshape(-1)
        if d8.numel() != src.size:
            raise WxError(f"upload size mismatch: {d8.numel()} device bytes vs {src.size} host bytes")
        for lo in range(0, src.size, self.chunk_bytes):
            hi = min(lo + self.chunk_bytes, src.size)
            slot = self.k & 1
            if self.k >= 2:
                self.free[slot].synchronize()                    # the copy that last read this staging slot is done
            st = self.stage[slot][: hi - lo]
            st.numpy()[:] = src[lo:hi]                            # memcpy into pinned memory
            with torch.cuda.stream(self.stream):
                d8[lo:hi].copy_(st, non_blocking=True)
                self.free[slot].record(self.stream)
            self.k += 1
            self.bytes += hi - lo

    def wait(self) -> None:
        """Make the current stream wait for every copy issued so far."""
        torch.cuda.current_stream(self.device).wait_stream(self.stream)


class DeviceTable:
    """Columns of a weather table resident in HBM: float64 (or float32) columns [N], optional Arrow validity bitmaps,
    int32 labels."""

    def __init__(self, names, cols, validity, labels32, n):
        self.names, self.cols, self.validity, self.labels32, self.n = names, cols, validity, labels32, n

    @property
    def any_nulls(self) -> bool:
        return any(v is not None for v in self.validity)


def _part_files(parquet_dir: str) -> list[str]:
    if os.path.isfile(parquet_dir):
        return [parquet_dir]
    files = sorted(f for f in glob.glob(os.path.join(parquet_dir, "*.parquet")) if not os.path.basename(f).startswith(("_", ".")))
    if not files:
        raise FileNotFoundError(f"no parquet part files under {parquet_dir}")
    return files


def read_parquet_columns(parquet_dir: str, device, columns=None, label_col: str = "label_encoded",
                         suffix: str | None = "_norm", chunk_bytes: int = 32 << 20) -> tuple[DeviceTable, dict]:
    """A Spark-style parquet directory -> device columns.  Feature columns: `columns`, or every column ending in
    `suffix` in file order (train_lightning_ddp.py:37).  Returns (table, timing dict)."""
    import pyarrow.parquet as pq

    t0 = time.perf_counter()
    files = _part_files(parquet_dir)
    metas = [pq.ParquetFile(f) for f in files]
    schema = metas[0].schema_arrow
    names = list(columns) if columns is not None else [c for c in schema.names if suffix is None or c.endswith(suffix)]
    if not names:
        raise ValueError("CRITICAL ERROR: No columns ending with '_norm' found. Check Spark logic.")
    rows = [m.metadata.num_rows for m in metas]
    n = int(sum(rows))
    dev = torch.device(device) if not isinstance(device, torch.device) else device
    np_dt = {"double": np.float64, "float": np.float32}[str(schema.field(names[0]).type)]
    t_dt = torch.float64 if np_dt == np.float64 else torch.float32
    cols = [torch.empty(n, dtype=t_dt, device=dev) for _ in names]
    vbytes = [None] * len(names)                       # host bitmaps, allocated on the first null seen
    labels = torch.empty(n, dtype=torch.int32, device=dev)
    up = ColumnUploader(dev, chunk_bytes)
    # parquet decode (pyarrow, GIL released) of the next parts runs on worker threads while this thread stages and
    # uploads the current one
    from concurrent.futures import ThreadPoolExecutor

    def _read(pf):
        t1 = time.perf_counter()
        tab = pf.read(columns=names + [label_col])
        return tab, time.perf_counter() - t1

    t_decode = 0.0
    row0 = 0
    pool = ThreadPoolExecutor(max_workers=2)
    pending = [pool.submit(_read, pf) for pf in metas[:2]]
    for idx in range(len(metas)):
        tab, td = pending.pop(0).result()
        t_decode += td
        if idx + 2 < len(metas):
            pending.append(pool.submit(_read, metas[idx + 2]))
        k = tab.num_rows
        for j, name in enumerate(names):
            vals, bits, boff, nulls = _arrow_column_buffers(tab[name], np_dt)
            up.put(cols[j][row0:row0 + k], vals)
            if nulls:
                if vbytes[j] is None:
                    vbytes[j] = np.full((n + 7) // 8, 0xFF, dtype=np.uint8)
                _merge_bits(vbytes[j], row0, _bitmap_slice(bits, boff, k), k)
        lv, lbits, _, lnulls = _arrow_column_buffers(tab[label_col], np.int32)
        if lnulls:
            raise WxError("label_encoded holds nulls; preprocess.py:24 never writes any")
        up.put(labels[row0:row0 + k], lv)
        row0 += k
    validity = []
    for vb in vbytes:
        if vb is None:
            validity.append(None)
        else:
            dv = torch.empty(vb.size, dtype=torch.uint8, device=dev)
            up.put(dv, vb)
            validity.append(dv)
    pool.shutdown(wait=True)
    up.wait()
    torch.cuda.current_stream(dev).synchronize()
    dt = time.perf_counter() - t0
    info = {"rows": n, "files": len(files), "bytes_h2d": up.bytes, "seconds": dt, "decode_seconds": t_decode,
            "gb_per_s": up.bytes / dt / 1e9,
            "decode_note": "decode_seconds is summed over the reader threads; it overlaps staging and upload"}
    return DeviceTable(names, cols, validity, labels, n), info


def _merge_bits(dst: np.ndarray, row0: int, src: np.ndarray, k: int) -> None:
    """Write the k validity bits of `src` (byte-aligned bitmap) into `dst` starting at bit row0."""
    if row0 % 8 == 0:
        full = k // 8
        dst[row0 // 8: row0 // 8 + full] = src[:full]
        if k % 8:
            keep = np.uint8((1 << (k % 8)) - 1)
            dst[row0 // 8 + full] = (dst[row0 // 8 + full] & ~keep) | (src[full] & keep)
        return
    bits = np.unpackbits(src, bitorder="little")[:k]
    lo, hi = row0 // 8, (row0 + k + 7) // 8
    cur = np.unpackbits(dst[lo:hi], bitorder="little")
    cur[row0 % 8: row0 % 8 + k] = bits
    dst[lo:hi] = np.packbits(cur, bitorder="little")


def load_processed(data_path: str, device=0) -> dict:
    """What `WeatherDataset(data_path)` holds (train_lightning_ddp.py:16-46), built on the GPU: the processed parquet
    directory goes to HBM as columns, and `features` fp32 [N,F] row-major / `labels` int64 [N] come out of
    wx_normalize_pack with identity statistics -- the float64 -> float32 cast of :45 (nulls -> NaN, as pandas gives)
    and the int32 -> int64 widening of :46, with no pandas DataFrame and no host pass over the values."""
    parquet_path = os.path.join(data_path, "data.parquet")
    if not os.path.exists(parquet_path):
        raise FileNotFoundError(f"CRITICAL ERROR: Data not found at {parquet_path}.\n"
                                "Did the Spark preprocessing step finish successfully?")
    table, info = read_parquet_columns(parquet_path, device)
    f = len(table.names)
    ctx = Context(f, device=table.cols[0].device)
    ident = torch.tensor([[float(table.n), 0.0, 0.0, 1.0]] * f, dtype=torch.float64, device=ctx.device)
    feats, _ = ctx.normalize_pack(table.cols, ident, impute=IMPUTE_NONE,
                                  validity=table.validity if table.any_nulls else None, want_nullmask=False)
    labels = table.labels32.to(torch.int64)
    ctx.close()
    return {"features": feats, "labels": labels, "feature_cols": table.names, "info": info}
