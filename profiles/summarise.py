"""Turns ncu captures brought back in gpurun_out/ into the text summaries committed here.

  python profiles/summarise.py r01 purchase=gpurun_out/purchase_r1h.ncu-rep seller_production=gpurun_out/sellerprod_r1f.ncu-rep

For every kernel=<report>: profiles/<round>_<kernel>_ncu.txt (launch shape, time, DRAM bytes, occupancy, pipe and stall
metrics of the --set full capture), profiles/<round>_<kernel>_source_lines.txt (instructions and stall samples per source
line, from the source page) and an entry in profiles/<round>_traffic.json (dram bytes per launch, read by bench.py).
"""
import collections
import csv
import json
import os
import subprocess
import sys

WANT = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct"]
STALLS = ["no_instruction", "barrier", "short_scoreboard", "long_scoreboard", "wait", "math_pipe_throttle", "mio_throttle", "lg_throttle",
          "branch_resolving", "dispatch_stall", "not_selected", "selected"]
UNIT = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rnd = sys.argv[1]
    here = os.path.dirname(os.path.abspath(__file__))
    tpath = os.path.join(here, rnd + "_traffic.json")
    traffic = json.load(open(tpath)) if os.path.exists(tpath) else {}
    for spec in sys.argv[2:]:
        name, rep = spec.split("=")
        rows = list(csv.reader(ncu(["-i", rep, "--page", "raw", "--csv"]).splitlines()))
        d = {h: (u, v) for h, u, v in zip(rows[0], rows[1], rows[2])}
        lines = ["# ncu --set full --clock-control none --import-source on; kernel %s; capture %s" % (name, os.path.basename(rep))]
        lines += ["%s [%s] = %s" % (w, d[w][0], d[w][1]) for w in WANT if w in d]
        for s in STALLS:
            k = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % s
            if k in d:
                lines.append("stall %s per issue = %s" % (s, d[k][1]))
        open(os.path.join(here, "%s_%s_ncu.txt" % (rnd, name)), "w").write("\n".join(lines) + "\n")
        b = sum(float(d[k][1]) * UNIT[d[k][0]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        traffic[name] = {"dram_bytes_per_launch": b, "source": "profiles/%s_%s_ncu.txt" % (rnd, name)}
        # per source line
        agg, smp, src = collections.Counter(), collections.Counter(), {}
        cur, hdr = None, None
        for r in csv.reader(ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]).splitlines()):
            if not r:
                continue
            if r[0] == "File Path":
                cur, hdr = r[1].split("/")[-1], None
            elif r[0] == "Line No":
                hdr, ie, ns = r, r.index("Instructions Executed"), r.index("# Samples")
            elif hdr is not None and cur is not None and r[0].isdigit():
                try:
                    key = (cur, int(r[0]))
                    agg[key] += int(r[ie] or 0)
                    smp[key] += int(r[ns] or 0)
                    src[key] = r[1][:100]
                except ValueError:
                    pass
        tot, ts = max(sum(agg.values()), 1), max(sum(smp.values()), 1)
        out = ["# warp instructions executed and stall samples per source line, kernel %s (%d instructions, %d samples)" % (name, tot, ts)]
        byfile = collections.Counter()
        for (f, _), v in agg.items():
            byfile[f] += v
        out.append("# by file: " + ", ".join("%s %.1f%%" % (f, 100.0 * v / tot) for f, v in byfile.most_common()))
        for key, v in agg.most_common(45):
            out.append("%-22s%5d %5.1f%% inst %5.1f%% smp  %s" % (key[0], key[1], 100.0 * v / tot, 100.0 * smp[key] / ts, src[key]))
        open(os.path.join(here, "%s_%s_source_lines.txt" % (rnd, name)), "w").write("\n".join(out) + "\n")
    json.dump(traffic, open(tpath, "w"), indent=1)
    print(json.dumps(traffic))


if __name__ == "__main__":
    main()
