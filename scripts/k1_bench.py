"""K1 streaming kernels vs the HBM roofline: GB/s of algorithmic bytes at table sizes >> L2."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, wxddp
PEAK = 6550.7
f = int(os.environ.get("F", 20)); n = int(os.environ.get("N", 1 << 24))
c = wxddp.Context(f)
REPS = int(os.environ.get("REPS", 10))
def timeit(fn, reps=None):
    reps = reps or REPS
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
res = {}
for dt, nm in ((torch.float32, "f32"), (torch.float64, "f64")):
    raw = (torch.randn(f, n, device="cuda") * 5 + 20).to(dt)
    raw[:, ::97] = float("nan")
    cols = [raw[j] for j in range(f)]
    es = raw.element_size()
    ms = timeit(lambda: c.col_stats(cols)); gb = n * f * es / ms / 1e6
    res[f"col_stats_{nm}"] = (ms, gb)
    stats = c.col_stats(cols)
    out = torch.empty(n, f, device="cuda"); nmk = None
    ms = timeit(lambda: c.normalize_pack(cols, stats, impute=1, out=out)); gb = n * (f * es + f * 4 + 4) / ms / 1e6
    res[f"normalize_pack_{nm}"] = (ms, gb)
    del raw, cols, out
    torch.cuda.empty_cache()
# label encode: strings "rain"/"no rain" in Arrow layout
rng = np.random.default_rng(0)
isr = rng.random(n) < 0.13
lens = np.where(isr, 4, 7).astype(np.int32)
off = np.zeros(n + 1, np.int32); np.cumsum(lens, out=off[1:])
data = np.empty(int(off[-1]), np.uint8)
pat_r = np.frombuffer(b"rain", np.uint8); pat_n = np.frombuffer(b"no rain", np.uint8)
# vectorised fill
pos = off[:-1]
for k in range(7):
    m = (lens > k)
    data[pos[m] + k] = np.where(isr[m], pat_r[min(k, 3)], pat_n[k])
offd, datad = torch.from_numpy(off).cuda(), torch.from_numpy(data).cuda()
ms = timeit(lambda: c.label_encode(offd, datad)); gb = n * (4 + 8) / ms / 1e6
res["label_encode"] = (ms, gb)
lab = c.label_encode(offd, datad)
assert int(lab.sum().item()) == int(isr.sum())
for k, (ms, gb) in res.items():
    print(f"{k:22s} N={n} F={f}: {ms:8.3f} ms  {gb:8.1f} GB/s  {100*gb/PEAK:5.1f}% of measured HBM peak")
print(json.dumps({k: {"ms": v[0], "gbs": v[1], "frac": v[1] / PEAK} for k, v in res.items()}))
