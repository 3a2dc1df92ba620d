"""Per-source-line instruction histogram of a kernel's time loop, from `nvdisasm --print-line-info` (needs -lineinfo).
    cuobjdump -xelf all obj.o; nvdisasm --print-line-info X.cubin | python tools/sass_lines.py <function-substring> [top]
The loop is the longest backward branch of the function that contains MUFU instructions (as in tools/sass_loop.py);
tools/sass_branches.py does the same for the branches only, following the inlining chain."""
import collections, re, sys
flt, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
lines = sys.stdin.read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and flt in l)
end = next((i for i in range(start + 1, len(lines)) if lines[i].lstrip().startswith(".section")), len(lines))
body, cur = [], None
for l in lines[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)), (m.group(3) or "").split("/")[-1], int(m.group(4) or 0))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        body.append((int(m.group(1), 16), m.group(2), cur))
best = None
for a, t, _ in body:
    m = re.search(r"BRA\S*\s+(?:.*\s)?`?\(?\.L_x_(\d+)\)?", t)
labels = {}
for i, l in enumerate(lines[start:end]):
    m = re.match(r"\.L_x_(\d+):", l.strip())
    if m:
        # address of the next instruction
        for l2 in lines[start + i + 1:end]:
            m2 = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l2)
            if m2:
                labels[m.group(1)] = int(m2.group(1), 16)
                break
for a, t, _ in body:
    m = re.search(r"BRA\S*\s+.*\.L_x_(\d+)", t)
    if m and m.group(1) in labels:
        tgt = labels[m.group(1)]
        # the time loop is the longest backward branch whose body holds SFU work (an unrolled attribute-load loop
        # ahead of it can be longer in a large topology)
        if tgt < a and (best is None or a - tgt > best[1] - best[0]) and any(tgt <= a2 <= a and "MUFU" in t2 for a2, t2, _ in body):
            best = (tgt, a)
lo, hi = best
per_line, per_op = collections.Counter(), collections.defaultdict(collections.Counter)
for a, t, c in body:
    if lo <= a <= hi and c:
        key = (c[0], c[1])
        per_line[key] += 1
        op = re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0]
        per_op[key][op] += 1
tot = sum(per_line.values())
print(f"loop {lo:#x}-{hi:#x}: {tot} instructions with line info")
byfile = collections.Counter()
for (f, ln), c in per_line.items():
    byfile[f] += c
print(dict(byfile))
for (f, ln), c in per_line.most_common(top):
    print(f"{f}:{ln:<5d} {c:4d}  {dict(per_op[(f, ln)].most_common(5))}")
