"""DRAM traffic per launch of the kernels in an `ncu --set full` report -> JSON that bench.py reads for `roofline.traffic`.
usage: ncu -i rep --page raw --csv | python tools/ncu_traffic.py <units per launch> "<command that was profiled>" > profiles/r2_ncu_traffic.json"""
import csv
import json
import sys

rows = list(csv.reader(sys.stdin))
hdr, units, data = rows[0], rows[1], rows[2:]
ki = hdr.index("Kernel Name")
ri, wi, ti = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tscale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}
out = {"units_per_launch": int(sys.argv[1]), "command": sys.argv[2] if len(sys.argv) > 2 else "", "kernels": {}}
for r in data:
    name = r[ki].split("(")[0].replace("void ", "").strip()
    out["kernels"].setdefault(name, []).append({
        "dram_read_bytes": float(r[ri].replace(",", "")) * scale[units[ri]],
        "dram_write_bytes": float(r[wi].replace(",", "")) * scale[units[wi]],
        "duration_us": float(r[ti].replace(",", "")) * tscale.get(units[ti], 1.0)})
print(json.dumps(out, indent=1))
