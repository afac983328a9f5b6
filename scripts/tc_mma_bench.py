import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, wxddp
c = wxddp.Context(20)
out = torch.zeros(2, dtype=torch.int64, device="cuda")
names = {8: "G1 layout, warp-collective issue (elect.sync)", 9: "G2 layout, warp-collective issue (elect.sync)", 4: "G1 layout, 4 independent accumulators", 5: "G1 layout, hoisted descriptors, 1 accumulator", 6: "G2 layout, 2 independent accumulators", 0: "G1 K-major SW128 (k inside 128B row) M128", 1: "G1 K-major no-swizzle K-blocks M128", 2: "G2 MN-major SW128_B32 M64", 3: "G2 K-major no-swizzle M64"}
for variant, ns in ((5, (64,)), (8, (64, 32, 128, 256)), (2, (24,)), (9, (24, 8, 64))):
    for n in ns:
        for reps in (64,):
            c.lib.wxdev_tc_mma_bench(c._h, variant, n, reps, out.data_ptr(), None); torch.cuda.synchronize()
            c.lib.wxdev_tc_mma_bench(c._h, variant, n, reps, out.data_ptr(), None); torch.cuda.synchronize()
            o = out.cpu().tolist()
            print(f"{names[variant]:45s} N={n:3d}: issue {o[0]/(reps*4):7.1f} cyc/MMA, total {o[1]/(reps*4):7.1f} cyc/MMA")
