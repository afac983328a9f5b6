"""Where does the small-batch step time go?  Host call cost vs kernel time, TC vs FFMA step kernel, raw vs packed input."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, wxddp
from bench import synth_raw_on_device
from wxddp.engine import StepEngine
F = 20
dev = torch.device("cuda", 0)
raw, y = synth_raw_on_device(torch, 1 << 22, F, 1234, dev)
cols = [raw[j] for j in range(F)]
for use_tc in (True, False):
    eng = StepEngine(F, device=dev, use_tc=use_tc)
    eng.fit_stats(cols)
    feats, _ = eng.ctx.normalize_pack(cols, eng.stats, impute=1, want_nullmask=False)
    for B in (256, 1024, 4096, 16384, 65536):
        nwin = raw.shape[1] // B
        for name, fn in (("raw", lambda i: eng.step_raw(cols, y, row0=(i % nwin) * B, batch=B)),
                         ("packed", lambda i: eng.step(feats, y, row0=(i % nwin) * B, batch=B))):
            for i in range(10): fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 300
            t0 = time.perf_counter(); e0.record()
            for i in range(n): fn(i)
            e1.record(); t_host = time.perf_counter() - t0
            torch.cuda.synchronize()
            print(f"use_tc={use_tc} B={B:6d} {name:6s}: device {1e3*e0.elapsed_time(e1)/n:7.2f} us/step, host enqueue {1e6*t_host/n:7.2f} us/step")
