"""Summarise an .ncu-rep (run where ncu is installed; no GPU needed):
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--lines 30]
Prints the headline metrics per captured launch, the SASS opcode mix and the hottest source lines."""
import collections
import csv
import io
import re
import subprocess
import sys

WANT = re.compile(r"^(dram__bytes_(read|write)\.sum|gpu__time_duration\.sum|launch__grid_size|launch__block_size|"
                  r"launch__registers_per_thread|launch__occupancy_limit_\w+|launch__shared_mem_per_block_dynamic|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"l1tex__throughput\.avg\.pct_of_peak_sustained_elapsed|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"lts__t_sector_hit_rate\.pct|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
                  r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum|"
                  r"smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|"
                  r"smsp__thread_inst_executed\.sum|"
                  r"smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio)$")


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    nlines = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 30
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("=== launch", r[0], r[hdr.index("Kernel Name")][:90])
        for h, u, v in zip(hdr, units, r):
            if WANT.match(h):
                try:
                    if "stalled" in h and float(v) < 0.15:
                        continue
                except ValueError:
                    pass
                print("  %-90s %-12s %s" % (h, u, v))
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]))))
    cur, hdr, agg, ops, first = None, None, {}, collections.Counter(), None
    for r in rows:
        if not r:
            continue
        if r[0] == "Function Name":
            if first is None:
                first = r[1]
            elif r[1] != first:
                pass
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            iI = hdr.index("Instructions Executed")
            continue
        if hdr is None:
            continue
        try:
            n = int(r[iI])
        except (ValueError, IndexError):
            continue
        if r[0] != "":
            key = (cur, int(r[0]))
            agg[key] = (agg.get(key, (0, ""))[0] + n, r[1].strip()[:100])
        else:
            m = r[3].split()
            if m:
                o = m[1] if m[0].startswith("@") and len(m) > 1 else m[0]
                ops[o.split(".")[0]] += n
    tot = sum(v[0] for v in agg.values()) or 1
    print("=== SASS opcode mix (warp-level instructions, all captured launches)")
    t2 = sum(ops.values()) or 1
    for k, v in ops.most_common(16):
        print("  %-10s %12d %5.1f%%" % (k, v, 100.0 * v / t2))
    print("=== hottest source lines")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:nlines]:
        print("  %-22s %12d %5.1f%%  %s" % ("%s:%d" % k, v[0], 100.0 * v[0] / tot, v[1]))


if __name__ == "__main__":
    main()
