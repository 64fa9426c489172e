#!/usr/bin/env python3
"""profiles/r2_blur_traffic.json from the text summary of an `ncu --set full` capture of the octave-0
blurs (tools/ncu_summary.py output): DRAM bytes per launch, stamped with the hash of the kernel source
so that bench.py can tell a stale figure from a current one.
    python tools/make_blur_traffic.py profiles/r2_kernels_v9.txt"""
import hashlib, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KERNEL_SRC = "sift3d_b200/csrc/blur_zmarch.cuh"
src = sys.argv[1]
t = open(src).read()
per = []
for s in t.split("== ")[1:]:
    if not s.startswith("void blur_zmarch"):
        continue
    r = float(re.search(r"dram__bytes_read.sum\s+([\d.]+) Mbyte", s).group(1))
    w = float(re.search(r"dram__bytes_write.sum\s+([\d.]+) Mbyte", s).group(1))
    us = float(re.search(r"gpu__time_duration.sum\s+([\d.]+) ms", s).group(1)) * 1e3
    per.append({"kernel": s.split("(")[0].strip(), "read_MB": r, "write_MB": w, "us_under_ncu": round(us, 1)})
out = {
    "dram_bytes_per_blur": 1e6 * sum(p["read_MB"] + p["write_MB"] for p in per) / len(per),
    "per_launch": per,
    "algorithmic_bytes_per_blur": 134217728,
    "kernel_source": KERNEL_SRC,
    "kernel_source_sha256": hashlib.sha256(open(os.path.join(ROOT, KERNEL_SRC), "rb").read()).hexdigest(),
    "source": f"ncu --set full --clock-control none of tools/run_once.py C2_mri_256 (SIFT3D_CUDA_SERIAL=1); {src}; "
              "dram__bytes_read.sum + dram__bytes_write.sum, mean of the six octave-0 launches (writes partly "
              "still in the 126 MB L2 at kernel end)",
}
json.dump(out, open(os.path.join(ROOT, "profiles", "r2_blur_traffic.json"), "w"), indent=1)
print(out["dram_bytes_per_blur"], out["kernel_source_sha256"][:16])
