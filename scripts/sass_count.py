"""Per-kernel SASS instruction mix of the built library (development aid)."""
import re, subprocess, sys, collections
lib = sys.argv[1] if len(sys.argv) > 1 else "starframe_b200/csrc/libsfp_b200.so"
pat = sys.argv[2] if len(sys.argv) > 2 else "narrowphase|substeps_coop|batch_frame"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
name, cnt = None, collections.defaultdict(collections.Counter)
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1); continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        op = m.group(1).split(".")[0]
        cnt[name]["total"] += 1; cnt[name][op] += 1
for n, c in cnt.items():
    if re.search(pat, n):
        short = re.sub(r"^_ZN\d+_GLOBAL__N__\w+?_cu_[0-9a-f]+", "", n)[:60]
        print(f"{short:60s} total {c['total']:6d} FFMA {c['FFMA']:5d} FMUL {c['FMUL']:5d} FADD {c['FADD']:5d} MUFU {c['MUFU']:4d} LDG {c['LDG']:4d} STG {c['STG']:4d}")
