"""Static SASS mnemonic counts per kernel of the built library (cuobjdump -sass; no GPU needed):
   python tools/sass_mnemonics.py > profiles/r02_sass_mnemonics.txt"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "chanamq_b200", "libchanagpu.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
def width(tok, base):
    if not tok.startswith(base):
        return None
    if ".128" in tok:
        return "128"
    if ".64" in tok:
        return "64"
    if any(x in tok for x in (".U8", ".S8", ".U16", ".S16")):
        return None
    return "32"


COLS = [("LDG.128", lambda t: width(t, "LDG.E") == "128"), ("LDG.64", lambda t: width(t, "LDG.E") == "64"),
        ("LDG.32", lambda t: width(t, "LDG.E") == "32"), ("STG.128", lambda t: width(t, "STG.E") == "128"),
        ("STG.EF", lambda t: t.startswith("STG.E") and ".EF" in t), ("STG.32", lambda t: width(t, "STG.E") == "32" and ".EF" not in t),
        ("LDS.128", lambda t: t.startswith("LDS.128")), ("LDGSTS", lambda t: t.startswith("LDGSTS")), ("ATOMS", lambda t: t.startswith("ATOMS")),
        ("UBLKCP", lambda t: t.startswith("UBLKCP")), ("SYNCS", lambda t: t.startswith("SYNCS")), ("SHFL", lambda t: t.startswith("SHFL")),
        ("VOTE", lambda t: t.startswith("VOTE")), ("POPC", lambda t: t.startswith("POPC"))]
kernels, cur = [], None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).split("::")[-1]
        cur = {"name": name, "n": 0, **{c: 0 for c, _ in COLS}}
        kernels.append(cur)
        continue
    m = re.search(r"/\*[0-9a-f]{4,6}\*/\s+(.*?);", line)
    if cur is None or not m:
        continue
    toks = m.group(1).split()
    tok = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
    cur["n"] += 1
    for c, f in COLS:
        if f(tok):
            cur[c] += 1
print("SASS mnemonics per kernel of chanamq_b200/libchanagpu.so (cuobjdump -sass, sm_100a; static counts; tools/sass_mnemonics.py).")
print("UBLKCP.S.G = cp.async.bulk global->shared (the staged long list of warp_weave in k_write_rest); SYNCS.* = its mbarrier;")
print("LDGSTS = cp.async 4-byte copies global->shared (the staging of warp_kway: the slices' lists start anywhere);")
print("STG.EF = the streaming (evict-first) stores of queue ids; LDG.128 in k_pull = the 16-byte loads from the peers' slice buffers;")
print("LDG.128 in k_probe = the 16-byte tag buckets and group heads, in k_pack24 = four ids per load.  k_write's own id traffic is")
print("32 bits per lane on purpose: 128-bit stores (four consecutive positions per lane) were built and measured slower (DESIGN.md section 5).")
print()
print("%-22s %6s " % ("kernel", "instrs") + " ".join("%-8s" % c for c, _ in COLS))
for k in kernels:
    print("%-22s %6d " % (k["name"][:22], k["n"]) + " ".join("%-8d" % k[c] for c, _ in COLS))
