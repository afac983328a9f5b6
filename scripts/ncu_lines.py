"""Read an ncu report here (no GPU): per-source-line instruction counts and stall samples of one kernel, plus the
headline counters.  usage: python scripts/ncu_lines.py gpurun_out/prof_tc.ncu-rep <object.o> <kernel-substring> [warps_per_tile_pass]"""
import collections, csv, io, re, subprocess, sys, os, tempfile

rep, obj, kname = sys.argv[1], sys.argv[2], sys.argv[3]
per = float(sys.argv[4]) if len(sys.argv) > 4 else 32768.0
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
amap, line, on = {}, None, False
for l in dis.splitlines():
    if l.lstrip().startswith(".text."):
        on = kname in l
        continue
    if not on:
        continue
    m = re.search(r'File "([^"]+)", line (\d+)', l)
    if m:
        line = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(\S+)", l)
    if m:
        amap[int(m.group(1), 16)] = (line, m.group(2))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
for k in ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
          "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
          "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
          "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread"]:
    if k in hdr:
        print(f"{k:75s} {rows[2][hdr.index(k)]} {rows[1][hdr.index(k)]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
ia, iex, ism = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
base, blocks, opc, stalls = None, collections.OrderedDict(), collections.Counter(), collections.Counter()
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[0] == "Address":
        break
    addr = int(r[ia], 16)
    base = addr if base is None else base
    ln, op = amap.get(addr - base, ((None, 0), "?"))
    ex, sm = int(r[iex] or 0), int(r[ism] or 0)
    b = blocks.setdefault(ln, [0, 0]); b[0] += ex; b[1] += sm
    opc[op.split(".")[0]] += ex
    waitline = ln and ln[0] == "tc_common.cuh" and 20 <= ln[1] <= 70
    for i in stall_cols:
        v = int(r[i] or 0)
        if v and not waitline:
            stalls[hdr[i]] += v
tot = sum(b[0] for b in blocks.values())
print(f"instructions per row-thread (all warps / {per:.0f}): {tot / per:.0f}")
for k, (ex, sm) in sorted(blocks.items(), key=lambda kv: (str(kv[0][0]) if kv[0] else "", kv[0][1] if kv[0] else 0)):
    if ex / per >= 8:
        print(f"  {str(k):32s} {ex / per:6.0f} instr/row  {sm:5d} samples")
print("opcodes/row:", [(k, round(v / per)) for k, v in opc.most_common(24)])
ts = sum(stalls.values())
print("stall reasons outside barrier waits:", [(k, f"{100 * v / ts:.0f}%") for k, v in stalls.most_common(10)])
