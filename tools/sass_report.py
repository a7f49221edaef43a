"""Hot-loop SASS evidence for profiles/ (runs in the container): python tools/sass_report.py <lib.so> <kernel substring> > profiles/<name>.txt
For every loop of the kernel that holds MUFU instructions and no other loop (the scoring loops: one MUFU.LG2 per scale evaluation, one MUFU.RCP per
orientation evaluation, or per PAIR with the shared reciprocal): the per-evaluation instruction tally and the full listing; then the TMA /
mbarrier staging instructions of the kernel with their addresses."""
import collections, re, sys, os
sys.path.insert(0, os.path.dirname(__file__))
import sass_loops as s
lib, pat = sys.argv[1], sys.argv[2]
name, rows = s.function_sass(lib, pat)
ai = {a: i for i, (a, _) in enumerate(rows)}
loops = []
for i, (a, t) in enumerate(rows):
    m = re.search(r"\bBRA\S*\s+(?:.*?)(0x[0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in ai:
            loops.append((ai[tgt], i))
inner = [l for l in loops if not any(o != l and l[0] <= o[0] and o[1] <= l[1] for o in loops)]
print(f"# {name}\n# {len(rows)} SASS instructions; built with nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false; cuobjdump -sass {os.path.basename(lib)}")
whole = collections.Counter(s.opcode(t) for _, t in rows)
print("# whole kernel, Blackwell-specific / staging opcodes:", {k: v for k, v in sorted(whole.items()) if k.startswith(("UBLKCP", "SYNCS", "FFMA2", "FADD2", "FMNMX3", "FFMA.SAT", "MUFU", "ACQBULK", "ELECT"))})
for b, e in inner:
    ops = collections.Counter(s.opcode(t) for _, t in rows[b:e + 1])
    lg2, rcp = ops.get("MUFU.LG2", 0), ops.get("MUFU.RCP", 0)
    if not (lg2 or rcp):
        continue
    n = e - b + 1
    f64 = any(k.startswith("DFMA") for k in ops)
    sat = ops.get("FFMA.SAT", 0)
    evals = sat if sat else (lg2 or rcp)  # winner-only loops: one FFMA.SAT per evaluation
    kind = ("scale" if lg2 else "orientation") + (" (precise tile: FP64 front end)" if f64 else " (FP32 tile)")
    print(f"\n## loop 0x{rows[b][0]:04x}..0x{rows[e][0]:04x}: {kind}; {n} instructions per iteration, {evals} evaluations per iteration"
          + (" (the loop lays out the degree-2 and the degree-3 polynomial variant side by side, half of the evaluations each; a pass runs one of them)" if ops.get("BRA.U", 0) else "")
          + f" -> {n / max(evals, 1):.2f} issue slots per evaluation")
    print("# per evaluation: " + ", ".join(f"{k} {v / max(evals, 1):.3f}" for k, v in ops.most_common()))
    for a, t in rows[b:e + 1]:
        print(f"    /*{a:04x}*/ {t}")
print("\n## TMA bulk copies and mbarrier operations (feature staging ring)")
for a, t in rows:
    if s.opcode(t).startswith(("UBLKCP", "SYNCS", "ACQBULK", "ELECT", "R2UR")):
        print(f"    /*{a:04x}*/ {t}")
