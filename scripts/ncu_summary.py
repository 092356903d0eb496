#!/usr/bin/env python
"""Turns gpurun_out ncu artefacts into the small text summaries committed under profiles/.
usage: ncu_summary.py launches <launches.csv> <out.csv> <note>
       ncu_summary.py kernel   <prof.ncu-rep>  <out.txt> <note>"""
import collections, csv, io, re, subprocess, sys

KEYS = re.compile(
    r"^(Kernel Name|Grid Size|Block Size)$|gpu__time_duration.sum|dram__bytes_(read|write).sum$|dram__throughput.avg.pct|"
    r"sm__throughput.avg.pct|sm__warps_active.avg.pct_of_peak|launch__registers_per_thread$|launch__occupancy_limit|"
    r"launch__shared_mem_per_block_dynamic|sm__inst_executed_pipe_fp64.avg.pct|"
    r"sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum|"
    r"l1tex__data_pipe_lsu_wavefronts_mem_shared.sum$|smsp__pcsamp_warps_issue_stalled_[a-z_]+$|lts__t_bytes.sum$|"
    r"lts__t_sector_hit_rate|sm__cycles_active.avg$|smsp__issue_active.avg.pct|smsp__inst_executed.sum$|"
    r"sm__inst_executed_pipe_(alu|fma|fmaheavy|lsu|fp64|xu|uniform|cbu|adu)\.sum$|l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum$")


def launches(src, dst, note):
    lines = [l for l in open(src) if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        k = re.sub(r"\(.*", "", r["Kernel Name"])
        agg[k][0] += 1
        agg[k][1] += float(r["Metric Value"].replace(",", "")) / 1e6
    tot = sum(v[1] for v in agg.values())
    out = ["# " + note, "# cold-cache, serialised per-launch times under ncu: compare SHARES, not absolutes",
           "kernel,launches,total_ms,share"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("%s,%d,%.3f,%.4f" % (k, v[0], v[1], v[1] / tot))
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


def kernel(src, dst, note):
    o = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(o)))
    with open(dst, "w") as f:
        f.write("# " + note + "\n")
        for row in rr[2:]:
            f.write("## launch\n")
            for h, u, v in zip(rr[0], rr[1], row):
                if KEYS.search(h) and not h.endswith("_not_issued"):
                    f.write("%s [%s] = %s\n" % (h, u, v))
    print(open(dst).read())


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel}[sys.argv[1]](*sys.argv[2:5])
