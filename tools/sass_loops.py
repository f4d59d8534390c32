"""Static view of a kernel's SASS: instruction count of every backward-branch loop body (developer tool)."""
import re, subprocess, sys
lib, fun = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks = out.split("Function : ")
for b in blocks:
    if not b.startswith("_Z") or fun not in b.split("\n")[0]: continue
    ins = []
    for line in b.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m: ins.append((int(m.group(1), 16), m.group(2)))
    addr_index = {a: i for i, (a, _) in enumerate(ins)}
    print(b.split("\n")[0], len(ins), "instructions")
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA\S*\s+(?:\S+,\s*)?`?\(?\.?L_x_\d+\)?|BRA\S*\s+.*0x([0-9a-f]+)", t)
        m2 = re.search(r"0x([0-9a-f]+)", t) if "BRA" in t else None
        if m2:
            tgt = int(m2.group(1), 16)
            if tgt <= a and tgt in addr_index:
                body = ins[addr_index[tgt]:i + 1]
                from collections import Counter
                c = Counter(x[1].split()[1].split(".")[0] if x[1].startswith("@") else x[1].split()[0].split(".")[0] for x in body)
                print("  loop %#x..%#x: %d instr  %s" % (tgt, a, len(body), dict(c.most_common(12))))
