"""Per-kernel SASS instruction mix of a built library: fmaheavy-pipe cycles estimate for the integer kernels.
Usage: python tools/sass_mix.py [lib.so] [name-filter]"""
import collections, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "upmem--openfhe_b200/lib/libfhe_b200.so"
flt = sys.argv[2] if len(sys.argv) > 2 else "ntt"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, mix = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); mix[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        mix[cur][m.group(1)] += 1
for k, c in mix.items():
    if flt not in k: continue
    name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()[:70]
    wide = sum(v for o, v in c.items() if o.startswith("IMAD.WIDE") or o.startswith("IMAD.HI"))
    fake = sum(v for o, v in c.items() if o.startswith(("IMAD.IADD", "IMAD.X", "IMAD.MOV", "IMAD.SHL")))
    narrow = sum(v for o, v in c.items() if o.startswith("IMAD")) - wide - fake
    alu = sum(v for o, v in c.items() if o.startswith(("IADD3", "SEL", "ISETP", "LOP3", "SHF", "LEA", "MOV", "VIADD", "IADD", "PRMT")))
    tot = sum(c.values())
    print(f"{name:70s} total {tot:5d} wide {wide:4d} imad {narrow:4d} imad-as-add {fake:4d} alu {alu:4d}  fmaheavy-cycles {wide*4+(narrow+fake)*2}")
