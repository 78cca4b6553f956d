"""Digests an .ncu-rep (read with ncu on the build box, no GPU needed) into a short text summary for profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r1_step_kernel.txt
"""
import collections
import csv
import io
import subprocess
import sys

RAW = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum",
       "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
       "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
       "launch__block_size", "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
       "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
       "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
       "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.sum",
       "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum",
       "sm__inst_executed_pipe_fp64.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
       "lts__t_sector_hit_rate.pct"]


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main(rep, out):
    lines = []
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr = rows[0]
    names = [r[hdr.index("Kernel Name")][:60] for r in rows[2:]]
    lines.append(f"report: {rep}")
    lines.append("launches captured: " + "; ".join(names))
    for m in RAW:
        if m in hdr:
            i = hdr.index(m)
            lines.append(f"{m} [{rows[1][i]}]: " + ", ".join(r[i] for r in rows[2:]))
    src = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"]))))
    hrow = next((r for r in src if "Address" in r and "# Samples" in r), None)
    if hrow:
        ia, ie, isamp = hrow.index("Address"), hrow.index("Instructions Executed"), hrow.index("# Samples")
        stall_cols = [(i, c) for i, c in enumerate(hrow) if c.startswith("stall_")]
        inst, samp, stalls = collections.Counter(), collections.Counter(), collections.Counter()
        cur, fname = None, None
        for r in src:
            if len(r) == 2 and r[0] == "File Path":
                fname = r[1].split("/")[-1]
            elif len(r) > ie and r[0].isdigit():
                cur = (fname, int(r[0]), r[1].strip()[:90])
            elif len(r) > ie and r[ia].startswith("0x") and cur:
                try:
                    inst[cur] += int(r[ie]); samp[cur] += int(r[isamp])
                except ValueError:
                    continue
                for i, c in stall_cols:
                    if i < len(r) and r[i].isdigit():
                        stalls[c] += int(r[i])
        ti, ts = max(1, sum(inst.values())), max(1, sum(samp.values()))
        lines.append(f"\nwarp instructions (all captured launches): {ti}; stall samples: {ts}")
        lines.append("stall reasons: " + ", ".join(f"{k}={100 * v / max(1, sum(stalls.values())):.1f}%" for k, v in stalls.most_common(8)))
        lines.append("\ntop source lines by executed warp instructions:")
        for k, n in inst.most_common(20):
            lines.append(f"  {100 * n / ti:5.1f}% inst {100 * samp[k] / ts:5.1f}% samples  {k[0]}:{k[1]}  {k[2]}")
        lines.append("\ntop source lines by stall samples:")
        for k, n in samp.most_common(15):
            lines.append(f"  {100 * n / ts:5.1f}% samples  {k[0]}:{k[1]}  {k[2]}")
    with open(out, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
