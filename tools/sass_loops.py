"""Hot-loop excerpts of a kernel's SASS (runs in the container, no GPU):
    python tools/sass_loops.py <lib.so> <substring of the mangled kernel name> [--min 40] [--full]
Finds every backward branch of the function, takes the innermost loops (no other loop inside), and prints for each its
address range, the instruction histogram, the number of LDS.128 / MUFU in it (one MUFU per evaluation in both scoring
loops, so issue slots per evaluation = instructions / MUFU count) and, with --full, the instruction listing.  The
staging sequence (UBLKCP = cp.async.bulk, SYNCS = mbarrier) is listed separately."""
import collections
import re
import subprocess
import sys


def function_sass(lib, pat):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    on, rows, name = False, [], None
    for line in out.splitlines():
        if "Function :" in line:
            on = pat in line
            if on:
                name = line.split("Function :")[1].strip()
                rows = []
            continue
        if not on:
            continue
        m = re.match(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?)\s*;?\s*/\*", line)
        if m:
            rows.append((int(m.group(1), 16), m.group(2).rstrip(" ;")))
    return name, rows


def opcode(text):
    t = re.sub(r"^@!?U?P[0-9T]+\s+", "", text)
    return t.split()[0] if t else "?"


def main():
    lib, pat = sys.argv[1], sys.argv[2]
    min_len = int(sys.argv[sys.argv.index("--min") + 1]) if "--min" in sys.argv else 40
    full = "--full" in sys.argv
    name, rows = function_sass(lib, pat)
    print(f"function {name}: {len(rows)} instructions")
    addr_index = {a: i for i, (a, _) in enumerate(rows)}
    loops = []
    for i, (a, t) in enumerate(rows):
        m = re.search(r"\bBRA\S*\s+(?:.*?)(0x[0-9a-f]+)", t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt <= a and tgt in addr_index:
                loops.append((addr_index[tgt], i))
    inner = [l for l in loops if not any(o != l and l[0] <= o[0] and o[1] <= l[1] for o in loops)]
    whole = collections.Counter(opcode(t) for _, t in rows)
    stage = {k: v for k, v in whole.items() if k.startswith(("UBLKCP", "SYNCS", "UTMA", "ACQBULK", "ARRIVES"))}
    print("staging / barrier instructions in the function:", dict(sorted(stage.items())))
    for b, e in sorted(inner):
        n = e - b + 1
        if n < min_len:
            continue
        ops = collections.Counter(opcode(t) for _, t in rows[b:e + 1])
        mufu = sum(v for k, v in ops.items() if k.startswith("MUFU"))
        kind = "scale (MUFU.LG2)" if any(k == "MUFU.LG2" for k in ops) else ("orientation (MUFU.RCP)" if any(k == "MUFU.RCP" for k in ops) else "other")
        f64 = sum(v for k, v in ops.items() if k.startswith(("DFMA", "DADD", "DMUL")))
        print(f"\nloop 0x{rows[b][0]:04x}..0x{rows[e][0]:04x}: {n} instructions, {kind}{', precise (FP64 front)' if f64 else ''}")
        if mufu:
            print(f"  evaluations per iteration (= MUFU count) {mufu}; issue slots per evaluation {n / mufu:.2f}")
        print("  " + ", ".join(f"{k} {v}" for k, v in ops.most_common()))
        if mufu:
            print("  per evaluation: " + ", ".join(f"{k} {v / mufu:.2f}" for k, v in ops.most_common()))
        if full:
            for a, t in rows[b:e + 1]:
                print(f"    /*{a:04x}*/ {t}")


if __name__ == "__main__":
    main()
