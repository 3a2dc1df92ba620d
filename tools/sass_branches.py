"""Which source lines do the branches of a kernel's time loop come from?  (needs -lineinfo)
    cuobjdump -xelf all obj.o; nvdisasm --print-line-info-inline X.cubin > k.txt
    python tools/sass_branches.py k.txt <function-substring> <loop lo hex> <loop hi hex>      (the range: tools/sass_loop.py)
Prints, per (innermost hgf_ops / hgf_math line, enclosing hgf_ops line), the number of BRA instructions in the range."""
import collections, re, sys
path, flt, lo, hi = sys.argv[1], sys.argv[2], int(sys.argv[3], 16), int(sys.argv[4], 16)
lines = open(path).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and flt in l)
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith(".text.") or lines[i].lstrip().startswith(".section")), len(lines))
addr_re = re.compile(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);")
chain, body, pending = [], [], []
for l in lines[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        pending.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    m = addr_re.match(l)
    if m:
        if pending:
            chain, pending = pending, []
        body.append((int(m.group(1), 16), m.group(2), tuple(chain)))
per = collections.Counter()
for a, t, c in body:
    if lo <= a <= hi and re.match(r"(@!?U?P\d+\s+)?BRA", t):
        key = next((x for x in c if x[0] in ("hgf_ops.cuh", "hgf_math.cuh")), c[0] if c else None)
        outer = next((x for x in c if x[0] == "hgf_ops.cuh"), None)
        per[(key, outer)] += 1
for k, v in per.most_common(40):
    print(v, k)
