#!/usr/bin/env python
"""Writes the record bench.py's roofline.traffic is read from: DRAM bytes of ONE launch of the dominant kernel.

usage: ncu_traffic.py <summary.txt | raw.csv> <gates_per_launch> <out.json>
  summary.txt  a profiles/*_ncu.txt file written by scripts/ncu_summary.py ("metric [unit] = value" lines)
  raw.csv      `ncu -i X.ncu-rep --page raw --csv` of an `ncu --set full` capture (first kernel row is used)
The record holds dram__bytes_read.sum / dram__bytes_write.sum in bytes, the kernel name, the launch width it was captured at
and the file it came from; bench.py only reports it when kernel and width match the run (else traffic = null)."""
import csv, json, os, re, sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def from_summary(path):
    vals, kernel = {}, None
    for line in open(path):
        m = re.match(r"Kernel Name \[\] = (?:void )?(.*)", line)
        if m:
            kernel = m.group(1).strip()
        m = re.match(r"(dram__bytes_(?:read|write)\.sum) \[(\w+)\] = ([0-9.eE+-]+)", line)
        if m:
            vals[m.group(1)] = float(m.group(3)) * UNIT[m.group(2)]
    return kernel, vals


def from_raw_csv(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names, units, first = rows[hdr], rows[hdr + 1], rows[hdr + 2]
    col = {n: i for i, n in enumerate(names)}
    vals = {}
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        vals[k] = float(first[col[k]].replace(",", "")) * UNIT[units[col[k]]]
    return first[col["Kernel Name"]], vals


if __name__ == "__main__":
    src, width, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    kernel, vals = from_raw_csv(src) if src.endswith(".csv") else from_summary(src)
    kernel = re.sub(r"^void ", "", kernel or "")
    rec = {"kernel": kernel, "gates_per_launch": width, "dram_bytes_read": vals["dram__bytes_read.sum"],
           "dram_bytes_write": vals["dram__bytes_write.sum"], "source": os.path.basename(src) + ", ncu --set full --clock-control none, one launch"}
    json.dump(rec, open(out, "w"), indent=1)
    print(rec)
