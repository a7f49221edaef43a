"""Turns gpurun_out/*.ncu-rep and launch lists into the text summaries committed under profiles/."""
import csv, subprocess, sys, os, collections
R = sys.argv[1] if len(sys.argv) > 1 else "r01"
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg", "sm__cycles_active.avg", "sm__inst_executed.avg.per_cycle_elapsed",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__thread_inst_executed_pred_on_per_inst_executed.ratio"]
OUT = sys.argv[2] if len(sys.argv) > 2 else R
import json
TRAFFIC = {}
for cfg, kern in (("c2", "k_score_resident"), ("c2", "k_score_fast"), ("c3", "k_score_fast"), ("c2", "k_sample_solve"), ("c2", "k_tail")):
    rep = f"gpurun_out/prof_{cfg}_{kern}_{R}.ncu-rep"
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(f"profiles/{OUT}_ncu_full_{kern}_{cfg}.txt", "w") as f:
        f.write(f"# ncu --set full --clock-control none, kernel {kern}, bench.py config {cfg.upper()} ({rep})\n")
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            f.write(f"kernel: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}\n")
            for h, u, v in zip(hdr, units, r):
                if h in KEYS or "issue_stalled" in h and "per_issue_active" in h:
                    f.write(f"{h:90s} {u:16s} {v}\n")
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tr = 0.0
            for h, u, v in zip(hdr, units, r):
                if h in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tr += float(v.replace(",", "")) * scale.get(u, 1.0)
            TRAFFIC.setdefault(cfg.upper(), {})[kern] = {"dram_bytes_per_launch": tr, "source": f"profiles/{OUT}_ncu_full_{kern}_{cfg}.txt"}
if TRAFFIC:
    old = json.load(open("profiles/traffic.json")) if os.path.exists("profiles/traffic.json") else {}
    for cfg, d in TRAFFIC.items():
        old.setdefault(cfg, {}).update(d)
    json.dump(old, open("profiles/traffic.json", "w"), indent=1)
rows = [r for r in csv.reader(open(f"gpurun_out/launches_c2_{R}.csv")) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    k = (r[4].split("(")[0], r[8], r[7])
    agg.setdefault(k, []).append(float(r[-1]))
tot = sum(sum(v) for k, v in agg.items() if "gcr::" in k[0])
with open(f"profiles/{OUT}_launch_list_c2.txt", "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none python bench.py --steps 5 --warmup 3 (config C2)\n")
    f.write("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
    f.write("# the list covers every loop of bench.py: device-resident steps, the kernel-timing pass and the host-buffer (e2e) steps,\n")
    f.write("# in which k_sample_solve reads the replayed samples in place over PCIe (about twice its device-resident duration)\n")
    f.write(f"{'kernel':60s} {'grid':>16s} {'block':>14s} {'launches':>9s} {'mean_us':>9s} {'share_of_gcr':>12s}\n")
    for (n, g, b), v in agg.items():
        share = sum(v) / tot if "gcr::" in n else float("nan")
        f.write(f"{n[:60]:60s} {g:>16s} {b:>14s} {len(v):9d} {sum(v)/len(v)/1e3:9.2f} {share:12.3f}\n")
print(open(f"profiles/{OUT}_launch_list_c2.txt").read())
