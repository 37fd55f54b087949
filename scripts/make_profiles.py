"""Turn the scratch artefacts in gpurun_out/ into the tracked summaries under profiles/ (round tag r01).

    ncu -i gpurun_out/prof_kernels.ncu-rep --page raw --csv    > /tmp/raw.csv
    ncu -i gpurun_out/prof_kernels.ncu-rep --page source --csv --print-source sass > /tmp/src_sass.csv
    python scripts/make_profiles.py r01
"""
import collections, csv, io, json, os, subprocess, sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out"); P = os.path.join(ROOT, "profiles")
# ---- stall reasons + hottest SASS per kernel ----
def stall_summary(src):
  out = []
  seen = set()
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

rep = os.path.join(G, "prof_kernels.ncu-rep")
if os.path.isfile(rep):      # head / V-trace kernels (scripts/gpu_profile.sh)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout

    # ---- per-kernel metrics (ncu --set full, one launch each, cold cache) ----
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    keep = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
            "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
    keep = [k for k in keep if k in idx]
    alg = {"0": ("head_eval_fwd", 446 * 256 * 20736 + 8 * 20736), "1": ("head_eval_bwd", 758 * 256 * 20736 + 8 * 20736),
           "2": ("head_sample", 474 * 256 * 20736 + 8 * 20736)}
    traffic = {}
    with open(os.path.join(P, f"{tag}_ncu_kernels.csv"), "w") as f:
        w = csv.writer(f)
        w.writerow(keep + ["entry_point", "algorithmic_bytes", "dram_bytes_total", "traffic_over_algorithmic"])
        w.writerow([units[idx[k]] for k in keep] + ["", "byte", "byte", ""])
        for r in rows[2:]:
            name = r[idx["Kernel Name"]]
            def val(k):
                v = float(r[idx[k]].replace(",", "")); u = units[idx[k]]
                return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
            tot = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
            if "vtrace" in name:
                ep, ab = "vtrace_loss", 37 * 65536 * 512 + 4 * 65536
                traffic.setdefault(ep, {})["sweep_T512_R65536"] = tot
            else:
                mode = name.split(",")[1].strip()
                ep, ab = alg[mode]
                traffic.setdefault(ep, {})["c2"] = tot
            w.writerow([r[idx[k]] for k in keep] + [ep, ab, int(tot), f"{tot / ab:.4f}"])
    json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)

    open(os.path.join(P, f"{tag}_ncu_stalls.md"), "w").write(
        f"# ncu --set full, source page summary ({tag})\n\nCommand: `ncu --set full --clock-control none --import-source on "
        "-k regex:'head_kernel|vtrace_tma' -s 4 -c 4 python scripts/profile_kernels.py 3` (config 2 shapes for the head "
        "kernels, T=512 x R=65536 for V-trace; one launch each, cold cache).\n\n" + "\n".join(stall_summary(src)))


# ---- encoder-side kernels (csrc/convgrad.cu, csrc/gemm.cu): second report ----
rep2 = os.path.join(G, "prof_encoder.ncu-rep")
if os.path.isfile(rep2):
    raw2 = subprocess.run(["ncu", "-i", rep2, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    src2 = subprocess.run(["ncu", "-i", rep2, "--page", "source", "--csv", "--print-source", "sass"],
                          capture_output=True, text=True).stdout
    rows2 = list(csv.reader(io.StringIO(raw2)))
    hdr2, units2 = rows2[0], rows2[1]
    idx2 = {h: i for i, h in enumerate(hdr2)}
    keep2 = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
             "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum",
             "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
             "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
             "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
             "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
             "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "smsp__inst_executed.sum"]
    keep2 = [k for k in keep2 if k in idx2]
    N = 20736
    with open(os.path.join(P, f"{tag}_ncu_encoder_kernels.csv"), "w") as f:
        w = csv.writer(f)
        w.writerow(keep2 + ["useful_flops", "useful_tflops"])
        w.writerow([units2[idx2[k]] for k in keep2] + ["flop", "TFLOP/s"])
        for r in rows2[2:]:
            name = r[idx2["Kernel Name"]]
            dur_ms = float(r[idx2["gpu__time_duration.sum"]].replace(",", ""))
            u = units2[idx2["gpu__time_duration.sum"]]
            dur_s = dur_ms * {"msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0}.get(u, 1e-3)
            flops = ""
            if "conv3x3" in name and "<" in name:
                t = [int(x) for x in name.split("<")[-1].split(">")[0].replace("(int)", "").split(",")[:3]]
                flops = 2.0 * N * t[0] * t[1] * (3 * t[2] - 2) ** 2
            elif "linear_tf32x3" in name:
                flops = 2.0 * N * 19968 * 256
            w.writerow([r[idx2[k]] for k in keep2] + [flops, f"{flops / dur_s / 1e12:.2f}" if flops else ""])
    open(os.path.join(P, f"{tag}_ncu_encoder_stalls.md"), "w").write(
        f"# ncu --set full, source page summary of the encoder-side kernels ({tag})\n\nCommand: `ncu --set full "
        "--clock-control none --import-source on -k regex:'conv3x3|linear_tf32x3|maxpool3s2|active_cells' -s 13 -c 13 "
        "python scripts/profile_encoder_kernels.py 2` (config 2 shapes, N = 20 736 frames; one launch each, cold "
        "cache).\n\n" + "\n".join(stall_summary(src2)))

if os.path.isfile(os.path.join(G, "launches.csv")):
    # ---- launch list of the learner step ----
    rows = [r for r in csv.reader(open(os.path.join(G, "launches.csv"))) if len(r) > 10]
    h = rows[0]; ix = {x: i for i, x in enumerate(h)}
    data = []
    for r in rows[1:]:
        try:
            data.append((r[ix["Kernel Name"]], float(r[ix["Metric Value"]].replace(",", ""))))
        except Exception:
            pass
    f = [i for i, d in enumerate(data) if "head_kernel" in d[0] and ", 0," in d[0]]
    span = data[f[-2]:f[-1]]
    tot = sum(d[1] for d in span)
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, t in span:
        agg[n][0] += 1; agg[n][1] += t
    with open(os.path.join(P, f"{tag}_step_launches_c2.csv"), "w") as fo:
        w = csv.writer(fo); w.writerow(["share_of_step_pct", "total_us", "launches", "kernel"])
        for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            w.writerow([f"{100 * t / tot:.2f}", f"{t / 1e3:.1f}", c, n[:160]])
    ours = sum(t for n, t in span if "mb::" in n or "head_kernel" in n or "vtrace" in n)
    print(f"step: {len(span)} launches, {tot / 1e6:.2f} ms summed (cold-cache, serialised), ours {100 * ours / tot:.2f}%")

for src_name, dst_name in (("step_dense.txt", "step_warm_profile_c2.txt"), ("encoder_layers.txt", "encoder_layers.txt"),
                           ("wgrad_time.txt", "conv_kernels_vs_cudnn.txt"), ("gemm_time.txt", "actor_linear_vs_cublas.txt")):
    p = os.path.join(G, src_name)
    if os.path.isfile(p):
        open(os.path.join(P, f"{tag}_{dst_name}"), "w").write(open(p).read())
for name in ("sweep_head.jsonl", "sweep_vtrace.jsonl", "bench.jsonl", "bench_ref.jsonl"):
    p = os.path.join(G, name)
    if os.path.isfile(p):
        open(os.path.join(P, f"{tag}_{name}"), "w").write(open(p).read())
print("profiles written:", sorted(os.listdir(P)))
