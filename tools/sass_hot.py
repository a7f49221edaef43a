"""Digest of the scoring loops of a kernel: python tools/sass_hot.py <lib.so> <mangled-name substring> [max loop length]"""
import collections, re, sys
sys.path.insert(0, __import__("os").path.dirname(__file__))
import sass_loops as s
name, rows = s.function_sass(sys.argv[1], sys.argv[2])
lim = int(sys.argv[3]) if len(sys.argv) > 3 else 700
ai = {a: i for i, (a, _) in enumerate(rows)}
print(name, len(rows), "instructions")
for i, (a, t) in enumerate(rows):
    m = re.search(r"\bBRA\S*\s+(?:.*?)(0x[0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in ai:
            b = ai[tgt]
            ops = collections.Counter(s.opcode(x) for _, x in rows[b:i + 1])
            n = i - b + 1
            if any(k.startswith("MUFU") for k in ops) and n < lim:
                print(hex(tgt), hex(a), n, dict(ops.most_common(14)))
