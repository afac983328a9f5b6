"""Count the Blackwell-specific SASS mnemonics per kernel of libwxddp.so (cuobjdump -sass; no GPU needed):
UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UBLKCP = TMA, SYNCS = mbarrier, LDGSTS = cp.async,
HMMA = legacy mma.sync (there must be none).  usage: python scripts/sass_mnemonics.py [lib] > profiles/r2_sass_mnemonics.txt"""
import collections, os, re, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
lib = sys.argv[1] if len(sys.argv) > 1 else None
if lib is None:
    import wxddp._build as b
    lib = b.LIB_PATH
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
PAT = [("UTC*MMA", r"\bUTC\w*MMA\b"), ("UTCBAR", r"\bUTCBAR\b"), ("LDTM", r"\bLDTM\b"), ("STTM", r"\bSTTM\b"),
       ("UTMALDG", r"\bUTMALDG\b"), ("UTMASTG", r"\bUTMASTG\b"), ("UBLKCP", r"\bUBLKCP\b"), ("SYNCS", r"\bSYNCS\b"),
       ("LDGSTS", r"\bLDGSTS\b"), ("HMMA", r"\bHMMA\b"), ("FFMA2", r"\bFFMA2\b")]
counts, name, n_inst = {}, None, collections.Counter()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        name = re.sub(r"\(.*", "", name)
        counts[name] = collections.Counter()
        continue
    if name is None or not re.match(r"\s+/\*[0-9a-f]+\*/", line):
        continue
    n_inst[name] += 1
    for key, pat in PAT:
        if re.search(pat, line):
            counts[name][key] += 1
print(f"# {os.path.basename(lib)}: SASS mnemonics per kernel (cuobjdump -sass, sm_100a).  Kernels without any of them are omitted.")
print("kernel".ljust(86) + "instr".rjust(7) + "".join(k.rjust(9) for k, _ in PAT))
for k in sorted(counts):
    c = counts[k]
    if not any(c[key] for key, _ in PAT if key not in ("LDGSTS", "FFMA2", "SYNCS")):
        continue
    print(k[:85].ljust(86) + str(n_inst[k]).rjust(7) + "".join(str(c[key]).rjust(9) for key, _ in PAT))
tot = collections.Counter()
for c in counts.values():
    tot.update(c)
print("total over the library".ljust(93) + "".join(str(tot[key]).rjust(9) for key, _ in PAT))
