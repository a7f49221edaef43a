"""Turns gpurun_out/prof_<cfg>_<kernel>_<R>.ncu-rep and gpurun_out/launches_<cfg>_<R>.csv (tools/profile_round.sh) into the text
summaries committed under profiles/: python tools/summarize_profiles.py <R> [<out prefix>]"""
import collections, csv, glob, json, os, re, subprocess, sys
R = sys.argv[1] if len(sys.argv) > 1 else "r2"
OUT = sys.argv[2] if len(sys.argv) > 2 else R
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg", "sm__cycles_active.avg", "sm__inst_executed.avg.per_cycle_elapsed",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__thread_inst_executed_pred_on_per_inst_executed.ratio"]
TRAFFIC = {}
for rep in sorted(glob.glob(f"gpurun_out/prof_*_{R}.ncu-rep")):
    m = re.match(rf"gpurun_out/prof_(c\d|fit)_(k_\w+?)_{R}\.ncu-rep", rep)
    if not m:
        continue
    cfg, kern = m.group(1), m.group(2)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    name = f"profiles/{OUT}_ncu_full_{kern}_{cfg}.txt"
    with open(name, "w") as f:
        what = "tools/time_fit.py C1 (gcr_fit_score_models / gcr_lo_round calls)" if cfg == "fit" else f"bench.py config {cfg.upper()} (winner-only run)"
        f.write(f"# ncu --set full --clock-control none --import-source on, kernel {kern}, {what}, {rep}\n")
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            f.write(f"kernel: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}\n")
            for h, u, v in zip(hdr, units, r):
                if h in KEYS or ("issue_stalled" in h and "per_issue_active" in h):
                    f.write(f"{h:90s} {u:16s} {v}\n")
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tr = sum(float(v.replace(",", "")) * scale.get(u, 1.0) for h, u, v in zip(hdr, units, r) if h in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            TRAFFIC.setdefault(cfg.upper(), {})[kern] = {"dram_bytes_per_launch": tr, "source": name}
        # stall samples by opcode (source page)
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
        srows = list(csv.reader(src.splitlines()))
        hi = next((i for i, r in enumerate(srows) if "# Samples" in r), None)
        if hi is not None:
            sh = {h: i for i, h in enumerate(srows[hi])}
            tot, st = collections.Counter(), collections.defaultdict(collections.Counter)
            cols = [c for c in srows[hi] if c.startswith("stall_") and "Not Issued" not in c]
            for r in srows[hi + 1:]:
                try:
                    n = int(r[sh["# Samples"]])
                except Exception:
                    continue
                t = re.sub(r"^@!?U?P[0-9T]+\s+", "", r[sh["Source"]].strip())
                op = t.split()[0] if t else "?"
                tot[op] += n
                for c in cols:
                    try:
                        st[op][c] += int(r[sh[c]])
                    except Exception:
                        pass
            allsamp = sum(tot.values())
            f.write(f"\n# warp stall samples by opcode (source page, {allsamp} samples): opcode, share, dominant stall reasons\n")
            for op, n in tot.most_common(12):
                f.write(f"{op:18s} {100.0 * n / max(allsamp, 1):5.1f} %  " + " ".join(f"{c[6:]}:{st[op][c]}" for c in cols if st[op][c] > 0.05 * n) + "\n")
    print("wrote", name)
if TRAFFIC:
    old = json.load(open("profiles/traffic.json")) if os.path.exists("profiles/traffic.json") else {}
    for cfg, d in TRAFFIC.items():
        old.setdefault(cfg, {}).update(d)
    json.dump(old, open("profiles/traffic.json", "w"), indent=1)
for path in sorted(glob.glob(f"gpurun_out/launches_*_{R}.csv")):
    cfg = re.match(rf"gpurun_out/launches_(c\d|fit)_{R}\.csv", path).group(1)
    rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0].isdigit()]
    agg = collections.OrderedDict()
    for r in rows:
        agg.setdefault((r[4].split("(")[0], r[8], r[7]), []).append(float(r[-1]))
    tot = sum(sum(v) for k, v in agg.items() if "gcr::" in k[0])
    name = f"profiles/{OUT}_launch_list_{cfg}.txt"
    with open(name, "w") as f:
        if cfg == "fit":
            f.write("# ncu --metrics gpu__time_duration.sum --clock-control none  REPS=3 python tools/time_fit.py C1: gcr_fit_score_models (refit shape, 50-fit batch,\n")
            f.write("# one small fit) and gcr_lo_round, three calls each.  Per-launch times are cold-cache and serialised.\n")
        else:
            f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none  python bench.py {'--config C2 ' if cfg == 'c2' else ''}--steps .. --warmup 3 (config {cfg.upper()})\n")
            f.write("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes.  The list covers every loop of bench.py: device-resident\n")
            f.write("# steps, the winner check, the kernel-timing pass and the host-buffer (e2e) steps.\n")
        f.write(f"{'kernel':60s} {'grid':>16s} {'block':>14s} {'launches':>9s} {'mean_us':>10s} {'share_of_gcr':>12s}\n")
        for (n, g, b), v in agg.items():
            share = sum(v) / tot if "gcr::" in n else float("nan")
            f.write(f"{n[:60]:60s} {g:>16s} {b:>14s} {len(v):9d} {sum(v) / len(v) / 1e3:10.2f} {share:12.3f}\n")
    print(open(name).read())
