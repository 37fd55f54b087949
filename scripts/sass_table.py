"""cuobjdump -sass libmbeast_sm100.so | python scripts/sass_table.py > profiles/r02_sass_mnemonics.txt
Counts of async-copy / tensor-core / atomic instructions per kernel (evidence that TMA, tcgen05 and TMEM are what runs)."""
import collections, re, subprocess, sys
keys = ["UBLKCP", "UTMALDG", "UTMASTG", "UTCHMMA", "LDTM", "UTCBAR", "UTCATOMSWS", "LDGSTS", "SYNCS", "ATOMG", "REDG", "RED", "FFMA", "MUFU"]
cur, counts = None, collections.defaultdict(collections.Counter)
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    if cur:
        for k in keys:
            if re.search(r"\b" + k + r"[\. ;]", line):
                counts[cur][k] += 1
want = sys.argv[1:] or ["actor_head_fwd", "cell_grad", "pack_cells", "vtrace_tma_kernel", "head_sample_exp", "linear_tf32x3", "cell_fill", "pack_actor"]
print("# cuobjdump -sass microbeast_b200/libmbeast_sm100.so (authoring container, sm_100a): instruction counts per kernel")
print("# UBLKCP = cp.async.bulk; UTMALDG / UTMASTG = cp.async.bulk.tensor load / store; UTCHMMA = tcgen05.mma; LDTM = tcgen05.ld;")
print("# UTCBAR = tcgen05.commit; SYNCS = mbarrier; LDGSTS = cp.async; ATOMG / REDG = global atomics")
for fn, c in sorted(counts.items()):
    dem = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip() or fn
    if any(x in dem for x in want):
        print(f"{dem[:120]:120s} " + " ".join(f"{k}={v}" for k, v in c.items() if v))
