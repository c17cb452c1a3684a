#!/usr/bin/env python
"""profiles/traffic.json from ncu captures: dram__bytes_read.sum + dram__bytes_write.sum of ONE launch per hot kernel
(`ncu --set full`, bench.py's default workload), stamped with the hash of the kernel sources they were taken on --
bench.py quotes `roofline.traffic` only when that hash matches the sources it runs.
usage: make_traffic.py <tag> <config> <targets>   (reads gpurun_out/<tag>_full10k_<kernel>.ncu-rep, writes profiles/traffic.json)"""
import csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import kernel_src_sha
tag, config, targets = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
names = {"k_merge_best2": "merge_best", "k_build_graph2": "build_graph", "k_normalize_warp": "normalize_trim"}
out = {"_source": f"ncu --set full, one launch each at the bench's default size ({targets} targets), profiles/{tag}_full10k_*_details.csv; "
                  "dram__bytes_read.sum + dram__bytes_write.sum per launch", "config": config, "targets": targets, "src_sha": kernel_src_sha()}
for k, key in names.items():
    rep = os.path.join(ROOT, "gpurun_out", f"{tag}_full10k_{k}.ncu-rep")
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    def get(name):
        i = hdr.index(name)
        v = float(vals[i].replace(",", ""))
        u = units[i].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}.get(u, 1)
    out[key] = get("dram__bytes_read.sum") + get("dram__bytes_write.sum")
json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(out)
