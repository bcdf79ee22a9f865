"""Reads `ncu -i X.ncu-rep --page raw --csv` output and writes the per-launch DRAM traffic of a kernel into
profiles/ncu_traffic.json, which bench.py quotes as `roofline.traffic` (scaled to its batch size).

    python tools/ncu_extract.py gpurun_out/r2_conv_raw.csv conv_umma fp16c32 4096 profiles/r2_conv_ncu_summary.md
    python tools/ncu_extract.py gpurun_out/r2_emit_raw.csv emit_u8_flat u8 32768 profiles/r2_emit_ncu_summary.md
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu_traffic.json")


def to_bytes(value, unit):
    v = float(value.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def main():
    path, kernel, mode, n, source = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), sys.argv[5]
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    rd, wr, dur = [], [], []
    for r in rows[2:]:
        if kernel not in r[idx["Kernel Name"]]:
            continue
        rd.append(to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]]))
        wr.append(to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]]))
        dur.append(r[idx["gpu__time_duration.sum"]] + " " + units[idx["gpu__time_duration.sum"]])
    if not rd:
        raise SystemExit("no launch of %s in %s" % (kernel, path))
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    data["%s/%s" % (kernel, mode)] = {"positions_per_launch": n, "launches": len(rd), "dram_bytes_read": sum(rd) / len(rd),
                                      "dram_bytes_write": sum(wr) / len(wr), "bytes": (sum(rd) + sum(wr)) / len(rd),
                                      "gpu_time": dur, "source": source}
    with open(OUT, "w") as f:
        json.dump(data, f, indent=1, sort_keys=True)
    print(json.dumps(data["%s/%s" % (kernel, mode)]))


if __name__ == "__main__":
    main()
