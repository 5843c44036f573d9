"""Per-kernel count of the SASS mnemonics that prove a Blackwell-native kernel (B200_PROFILING.md): UTC*MMA (tcgen05.mma),
LDTM/STTM (tcgen05.ld/st), UTMALDG/UTMASTG/UBLKCP (TMA), SYNCS (mbarrier), MULTIMEM, peer/vector LDG/STG, REDG/ATOMG."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "distributed_model_parallel_b200", "_lib", "libdmp_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
keys = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "HMMA", "MULTIMEM", "LDG.E.128", "STG.E.128",
        "LDG.E.64", "REDG", "ATOMG", "LDGSTS", "BAR.SYNC", "SHFL"]
fn, counts = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        try:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        except Exception:
            fn = m.group(1)
        fn = re.sub(r"\(anonymous namespace\)::", "", fn).split("(")[0].replace("void ", "")
        counts[fn] = collections.Counter()
        continue
    if fn is None:
        continue
    m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        for k in keys:
            if op.startswith(k):
                counts[fn][k] += 1
print("# SASS mnemonic counts per kernel of libdmp_b200.so (sm_100a).  UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA load,")
print("# SYNCS = mbarrier ops, UTCBAR = tcgen05.commit, REDG = global reduction (split-K / weight-gradient accumulation).")
for fn, c in counts.items():
    if c:
        print("%-78s %s" % (fn[:78], "  ".join("%s=%d" % (k, c[k]) for k in keys if c[k])))
