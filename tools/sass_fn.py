"""Developer tool: print the SASS of one kernel of libpgt_cuda.so (instruction lines only) and the opcode mix of a line range.

    python tools/sass_fn.py <mangled-name-substring> [lo hi]
"""
import collections, re, subprocess, sys
txt = subprocess.run(["cuobjdump", "-sass", "pyglasstools_b200/libpgt_cuda.so"], capture_output=True, text=True).stdout.splitlines()
starts = [i for i, l in enumerate(txt) if "Function :" in l]
body = []
for si, i in enumerate(starts):
    if sys.argv[1] in txt[i]:
        body = txt[i:(starts[si + 1] if si + 1 < len(starts) else len(txt))]
        break
ins = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).strip() for l in body if re.search(r"/\*[0-9a-f]{4,5}\*/\s+\S", l)]
def op(l):
    t = l.split("*/", 1)[1].strip().split()
    o = t[1] if t[0].startswith("@") else t[0]
    return re.sub(r"\..*", "", o)
if len(sys.argv) > 3:
    lo, hi = int(sys.argv[2]), int(sys.argv[3])
    for k in range(lo, hi):
        print(k, ins[k])
    print(hi - lo, collections.Counter(op(l) for l in ins[lo:hi]).most_common())
else:
    for k, l in enumerate(ins):
        print(k, l)
