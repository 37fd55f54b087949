"""gpurun_out/prof_r2.ncu-rep + launches_r2.csv -> profiles/r02_ncu_kernels.csv, r02_ncu_stalls.md, r02_step_launches_c2.csv,
traffic.json (ncu --set full: per-launch, cold cache, serialised -- never a bench value)."""
import collections, csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
sys.argv = [sys.argv[0], "r02_skip"]
rep = os.path.join(G, "prof_r2.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout


def stall_summary(src):
    out, seen = [], set()
    for blk in src.split('"Kernel Name",')[1:]:
        lines = blk.splitlines(); name = lines[0].strip('",')
        if name in seen:
            continue
        seen.add(name)
        rs = list(csv.reader(io.StringIO("\n".join(lines[1:])))); h = rs[0]; col = {x: i for i, x in enumerate(h)}
        stalls = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
        tot = collections.Counter(); data = []; ops = collections.Counter(); ninst = 0
        for k, r in enumerate(rs[1:]):
            try:
                n = int(r[col["# Samples"]])
            except Exception:
                continue
            for x in stalls:
                tot[x] += int(r[col[x]] or 0)
            best = max(stalls, key=lambda x: int(r[col[x]] or 0))
            ex = int(r[col["Instructions Executed"]] or 0)
            s = r[col["Source"]].strip(); t = s.split()
            op = (t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")).split(".")[0]
            ops[op] += ex; ninst += ex
            data.append((n, k, ex, best[6:], s[:80]))
        total = sum(d[0] for d in data)
        out.append(f"## {name}\nsamples {total}, warp-instructions executed {ninst}\n")
        out.append("stall reasons: " + ", ".join(f"{k[6:]} {100 * v / max(total, 1):.1f}%" for k, v in tot.most_common(8)) + "\n")
        out.append("instruction mix: " + ", ".join(f"{k} {100 * v / max(ninst, 1):.1f}%" for k, v in ops.most_common(14)) + "\n")
        out.append("hottest SASS (samples, share, #executed, dominant stall, instruction):\n```")
        for n, k, ex, b, s in sorted(data, reverse=True)[:10]:
            out.append(f"{n:7d} {100 * n / max(total, 1):5.1f}% x{ex:>9d} {b:16s} {s}")
        out.append("```\n")
    return out


rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
keep = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
keep = [k for k in keep if k in idx]
N, C = 20736, 256
alg = {"pack_cells": ("pack_cells", "byte", (78 + 56 + 16) * N * C), "actor_head_fwd": ("actor_head_fwd", "flop (3 x TF32)", 3 * 2.0 * N * 80 * C * 256),
       # list form: ~6 % of the pairs are live at the "dense" preset; issued flops = 128-entry tiles x N 80 x K 256 x 3 passes
       "actor_head_list_fwd": ("actor_head_list_fwd", "flop (3 x TF32, live pairs only)", None),
       "cell_grad": ("cells_bwd", "flop (fp32 FMA x2)", None), "head_kernel": ("head_sample", "byte", 474 * C * N + 8 * N),
       "vtrace": ("vtrace_loss", "byte", 37 * 65536 * 512 + 4 * 65536),
       # tensor-core convolutions: issued TF32 flops = 3 passes x 2 x pixels x 9 taps x cin x cout
       "TcCfg<32, 32, 4>": ("conv3x3_tc_32x32x4", "flop (3 x TF32)", 3 * 2.0 * N * 16 * 9 * 32 * 32),
       "TcCfg<16, 16, 8>": ("conv3x3_tc_16x16x8", "flop (3 x TF32)", 3 * 2.0 * N * 64 * 9 * 16 * 16),
       # first convolution from bits: bytes = one int32 per cell in, 16 fp32 per cell out (fwd); the reverse + 7 KB (wgrad)
       "conv1_bits_fwd": ("conv1_bits_fwd", "byte", N * 256 * (4 + 64)), "conv1_bits_wgrad": ("conv1_bits_wgrad", "byte", N * 256 * (4 + 64))}
traffic = json.load(open(os.path.join(P, "traffic.json")))
with open(os.path.join(P, "r02_ncu_kernels.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow(keep + ["entry_point", "algorithmic_unit", "algorithmic_amount", "dram_bytes_total", "algorithmic_rate"])
    w.writerow([units[idx[k]] for k in keep] + ["", "", "", "byte", "GB/s or TFLOP/s"])
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        key = next((k for k in alg if k in name), None)
        if key is None:
            continue
        ep, unit, amount = alg[key]

        def val(k):
            v = float(r[idx[k]].replace(",", "")); u = units[idx[k]]
            return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9,
                        "second": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0}.get(u, 1)
        tot = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
        dur = val("gpu__time_duration.sum")
        rate = ""
        if amount:
            rate = f"{amount / dur / (1e9 if unit == 'byte' else 1e12):.1f}"
        traffic.setdefault(ep, {})["c2" if ep != "vtrace_loss" else "sweep_T512_R65536"] = tot
        w.writerow([r[idx[k]] for k in keep] + [ep, unit, amount or "", int(tot), rate])
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
open(os.path.join(P, "r02_ncu_stalls.md"), "w").write(
    "# ncu --set full, source page summary (r02)\n\nCommand: `ncu --set full --clock-control none --import-source on "
    "-k regex:'pack_cells_kernel|actor_head_fwd_kernel|actor_head_list_fwd_kernel|cell_grad_kernel|head_kernel|vtrace_tma|conv3x3_tc_kernel|conv1_bits' -s 10 -c 10 python "
    "scripts/profile_r2.py 2` (config 2 shapes: 20 736 frames x 256 cells; T=512 x R=65 536 for V-trace; 20 736 frames for the convolutions; one launch each, "
    "cold cache).\n\n" + "\n".join(stall_summary(src)))

rows = [r for r in csv.reader(open(os.path.join(G, "launches_r2.csv"))) if len(r) > 10]
h = rows[0]; ix = {x: i for i, x in enumerate(h)}
data = []
for r in rows[1:]:
    try:
        data.append((r[ix["Kernel Name"]], float(r[ix["Metric Value"]].replace(",", ""))))
    except Exception:
        pass
# one default step = from one launch of the list-form actor head to the next (the dense head only runs in bench.py's
# separate info.roofline_tensor pass at the end)
f = [i for i, d in enumerate(data) if "actor_head_list_fwd_kernel" in d[0]]
if len(f) < 3:
    f = [i for i, d in enumerate(data) if "actor_head_fwd_kernel" in d[0]]
span = data[f[2]:f[3]] if len(f) > 3 else data[f[-2]:f[-1]]
tot = sum(d[1] for d in span)
agg = collections.defaultdict(lambda: [0, 0.0])
for n, t in span:
    agg[n][0] += 1; agg[n][1] += t
with open(os.path.join(P, "r02_step_launches_c2.csv"), "w") as fo:
    w = csv.writer(fo); w.writerow(["share_of_step_pct", "total_us", "launches", "kernel"])
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        w.writerow([f"{100 * t / tot:.2f}", f"{t / 1e3:.1f}", c, n[:160]])
ours = sum(t for n, t in span if "mb::" in n)
print(f"step: {len(span)} launches, {tot / 1e6:.2f} ms summed (cold-cache, serialised), ours {100 * ours / tot:.2f}%")
