"""Developer tool: fills the @PLACEHOLDER@ fields of DESIGN.md / README.md from the committed profiles/ files."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
def load(name):
    p = os.path.join(ROOT, "profiles", name)
    return json.loads(open(p).read().splitlines()[-1]) if os.path.exists(p) else None
vals = json.load(open(os.path.join(ROOT, "tools", "design_values.json")))
for fn in ("DESIGN.md", "README.md"):
    p = os.path.join(ROOT, fn)
    s = open(p).read()
    for k, v in vals.items():
        s = s.replace("@%s@" % k, str(v))
    open(p, "w").write(s)
    import re
    left = set(re.findall(r"@[A-Z0-9_]+@", s))
    if left: print(fn, "unfilled:", left)
