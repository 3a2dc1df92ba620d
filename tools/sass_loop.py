"""Static instruction mix of the time loop of a kernel: finds the longest backward branch in each function of
a cuobjdump -sass listing and counts the instructions inside it by class (MUFU = SFU pipe).
    cuobjdump -sass obj.o | python tools/sass_loop.py [name-filter]"""
import re, sys, collections
flt = sys.argv[1] if len(sys.argv) > 1 else ""
funcs, cur = [], None
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = (m.group(1), []); funcs.append(cur); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m and cur:
        cur[1].append((int(m.group(1), 16), m.group(2)))
for name, ins in funcs:
    if flt not in name: continue
    # the time loop: the backward branch whose body holds the most SFU instructions (ties: the longest) -- the
    # parameter-loading prologue of a wide topology is a longer loop, but has none
    best, best_key = None, None
    for addr, text in ins:
        m = re.search(r"BRA\S*\s+(?:.*\s)?0x([0-9a-f]+)", text)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < addr:
                key = (sum(1 for a, t in ins if tgt <= a <= addr and "MUFU" in t), addr - tgt)
                if best is None or key > best_key: best, best_key = (tgt, addr), key
    if not best:
        print(name, "no loop found", len(ins)); continue
    body = [t for a, t in ins if best[0] <= a <= best[1]]
    c = collections.Counter()
    for t in body:
        t = re.sub(r"^@!?U?P\d+\s+", "", t)
        op = t.split()[0].split(".")[0]
        c[op] += 1
    fp = sum(c[k] for k in ("FFMA", "FMUL", "FADD", "FFMA2", "FMUL2", "FADD2"))
    cmp_ = c['FSETP'] + c['FSEL'] + c['FMNMX']
    print(f"{name[:110]}\n  loop {best[0]:#x}-{best[1]:#x}: {len(body)} instr; MUFU {c['MUFU']}; FFMA/FMUL/FADD {fp}; "
          f"FSETP/FSEL/FMNMX {cmp_}; other {len(body)-fp-c['MUFU']-cmp_}")
    print("  ", dict(c.most_common(14)))
