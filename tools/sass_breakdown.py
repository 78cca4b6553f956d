"""Static SASS instruction count of one kernel, attributed to (a) the top-level source line of the kernel body and
(b) the innermost inlined function's file.  Usage:
  cuobjdump -xelf all libchs.so; nvdisasm --print-line-info-inline X.cubin > all.sass
  python tools/sass_breakdown.py all.sass '<mangled kernel name substring>' [bucket]"""
import re, sys, collections
path, key = sys.argv[1], sys.argv[2]
bucket = int(sys.argv[3]) if len(sys.argv) > 3 else 20
lines = open(path).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and key in l)
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith(".text.") or lines[i].startswith(".section")), len(lines))
ins = re.compile(r"^\s*/\*[0-9a-f]{4,6}\*/\s+(\S+)")
loc = re.compile(r'//## File "([^"]+)", line (\d+)( inlined at)?')
top = collections.Counter(); inner = collections.Counter(); total = 0
cur_top = cur_inner = None; group = []
for l in lines[start:end]:
    m = loc.search(l)
    if m:
        group.append((m.group(1).split("/")[-1], int(m.group(2)), bool(m.group(3))))
        continue
    m = ins.match(l)
    if m:
        if group:
            cur_inner = group[0][:2]
            cur_top = next(((f, n) for f, n, inl in group if not inl), group[-1][:2])
            group = []
        total += 1
        top[(cur_top[0], cur_top[1] // bucket * bucket)] += 1
        inner[cur_inner[0]] += 1
print("total instructions", total)
print("-- by top-level line bucket")
for (f, n), c in sorted(top.items()): print(f"{f}:{n:4d}  {c}")
print("-- by innermost file")
for f, c in inner.most_common(): print(f, c)
