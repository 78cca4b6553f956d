"""Per-source-line table (instructions, stall samples, top stall reason) of an .ncu-rep, and per-phase totals.
    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [n_lines]"""
import collections
import csv
import io
import subprocess
import sys


def load(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
    src = list(csv.reader(io.StringIO(out)))
    hrow = next(r for r in src if "Address" in r and "# Samples" in r)
    ia, ie, isamp = hrow.index("Address"), hrow.index("Instructions Executed"), hrow.index("# Samples")
    stall_cols = [(i, c) for i, c in enumerate(hrow) if c.startswith("stall_") and "Not Issued" not in c]
    inst, samp = collections.Counter(), collections.Counter()
    stalls = collections.defaultdict(collections.Counter)
    text = {}
    cur, fname = None, None
    for r in src:
        if len(r) == 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif len(r) > ie and r[0].isdigit():
            cur = (fname, int(r[0]))
            text[cur] = r[1].strip()[:100]
        elif len(r) > ie and r[ia].startswith("0x") and cur:
            try:
                inst[cur] += int(r[ie]); samp[cur] += int(r[isamp])
            except ValueError:
                continue
            for i, c in stall_cols:
                if i < len(r) and r[i].isdigit():
                    stalls[cur][c] += int(r[i])
    return inst, samp, stalls, text


if __name__ == "__main__":
    inst, samp, stalls, text = load(sys.argv[1])
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    ti, ts = sum(inst.values()), sum(samp.values())
    print(f"total warp inst {ti}, samples {ts}")
    byfile_i, byfile_s = collections.Counter(), collections.Counter()
    for k in inst:
        byfile_i[k[0]] += inst[k]; byfile_s[k[0]] += samp[k]
    for f in byfile_i:
        print(f"  {f:22s} inst {100 * byfile_i[f] / ti:5.1f}%  samples {100 * byfile_s[f] / ts:5.1f}%")
    print("top lines by samples:")
    for k, v in samp.most_common(n):
        top = ", ".join(f"{c[6:]}={x}" for c, x in stalls[k].most_common(3))
        print(f"  {100 * v / ts:5.1f}% smp {100 * inst[k] / ti:5.1f}% inst  {k[0]}:{k[1]}  {text[k][:70]}  [{top}]")
