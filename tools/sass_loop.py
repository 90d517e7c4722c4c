"""Static look at the main loop of a kernel in an object file: python tools/sass_loop.py <obj> <kernel-substring> [dump]
Finds the widest backward branch, counts the instructions inside it by opcode (predicated-off side paths included)."""
import collections, re, subprocess, sys
obj, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
funcs = out.split("Function : ")
for f in funcs[1:]:
    name = f.split("\n", 1)[0]
    if pat not in name:
        continue
    ins = []
    for line in f.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    best = None
    for a, t in ins:
        m = re.search(r"BRA(?:\.U)?(?:\.ANY)?\s+(?:[!U]*P\d,\s*)?`?\(?\.?L?_?x?_?(\w+)\)?", t)
        m2 = re.search(r"BRA\S*\s.*?0x([0-9a-f]+)", t)
        if m2:
            tgt = int(m2.group(1), 16)
            if tgt < a and (best is None or a - tgt > best[1] - best[0]):
                best = (tgt, a)
    print(name[:100], "instructions", len(ins), "loop", best and (hex(best[0]), hex(best[1])))
    if best:
        body = [t for a, t in ins if best[0] <= a <= best[1]]
        hist = collections.Counter(re.sub(r"^@!?U?P\d\s+", "", t).split()[0].split(".")[0] for t in body)
        print("  loop instructions", len(body), dict(hist.most_common(40)))
        if len(sys.argv) > 3:
            for a, t in ins:
                if best[0] <= a <= best[1]:
                    print(f"  {a:05x} {t}")
