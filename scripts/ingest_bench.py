"""Columnar ingest rate: a Spark-style parquet directory -> HBM -> fp32 row-major features.
Writes two synthetic directories (S2: 145,460 x 20; ~1 GB: 5.9 M x 20 float64 + int32 labels, 1% nulls), then times
wxddp.ingest.load_processed (parquet decode by pyarrow on the host cores, pinned chunked H2D on a copy stream, cast +
pack kernel) next to the reference's loader (pandas.read_parquet + torch casts, train_lightning_ddp.py:31,45-46)."""
import json, os, shutil, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pyarrow as pa, pyarrow.parquet as pq, torch
from wxddp import ingest
from wxddp.module import WeatherDataset

def write_dir(root, n, f, parts, null_rate=0.01, seed=3):
    out = os.path.join(root, "data.parquet"); os.makedirs(out)
    rng = np.random.default_rng(seed)
    bounds = np.linspace(0, n, parts + 1).astype(int)
    for p in range(parts):
        k = bounds[p + 1] - bounds[p]
        cols = {}
        for j in range(f):
            v = rng.standard_normal(k)
            cols[f"Sensor_{j:02d}_norm"] = pa.array(v, mask=rng.random(k) < null_rate)
        cols["label_encoded"] = pa.array((rng.random(k) < 0.13).astype(np.int32))
        pq.write_table(pa.table(cols), os.path.join(out, f"part-{p:05d}.snappy.parquet"), compression="snappy")
    open(os.path.join(out, "_SUCCESS"), "w").close()

res = []
for tag, n, f, parts in (("S2 145,460 x 20", 145_460, 20, 2), ("1 GB 5,900,000 x 20", 5_900_000, 20, 8)):
    tmp = tempfile.mkdtemp()
    try:
        write_dir(tmp, n, f, parts)
        ingest.load_processed(tmp, device=0)                       # warm-up (CUDA context, page cache)
        t0 = time.perf_counter(); got = ingest.load_processed(tmp, device=0); torch.cuda.synchronize(); t_gpu = time.perf_counter() - t0
        t0 = time.perf_counter(); ds = WeatherDataset(tmp); x = ds.features.cuda(); y = ds.labels.cuda(); torch.cuda.synchronize(); t_ref = time.perf_counter() - t0
        same = bool(torch.equal(got["features"].view(torch.int32), x.view(torch.int32)))
        info = got["info"]
        res.append({"table": tag, "bytes_h2d": info["bytes_h2d"], "wxddp_seconds": t_gpu, "decode_seconds": info["decode_seconds"],
                    "wxddp_gb_per_s": info["bytes_h2d"] / t_gpu / 1e9,
                    "reference_loader_seconds": t_ref, "speedup": t_ref / t_gpu, "bit_identical": same})
        print(res[-1], flush=True)
    finally:
        shutil.rmtree(tmp)
print(json.dumps({"ingest": res, "host_cores": os.cpu_count()}))
