#!/usr/bin/env python3
"""profiles/traffic_per_launch.json from an ncu capture of the classification kernel
(`ncu --set full ... bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline`), stamped with the hash of
the kernel sources it was taken from: bench.py reports `roofline.traffic` only while that hash matches.

  python scripts/update_traffic.py gpurun_out/prof.ncu-rep [records] [tile]
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    rep = sys.argv[1]
    records = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
    tile = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {h: (float(v.replace(",", "")), u) for h, u, v in zip(hdr, units, vals) if v.replace(",", "").replace(".", "", 1).isdigit()}

    def to_bytes(key):
        v, u = d[key]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    rd, wr = to_bytes("dram__bytes_read.sum"), to_bytes("dram__bytes_write.sum")
    kernel = vals[hdr.index("Kernel Name")]
    assert f"{tile}" in kernel, kernel
    info = {"records": records, "tile": tile, "dram_bytes_per_launch": int(rd + wr), "dram_bytes_read": int(rd),
            "dram_bytes_write": int(wr), "gpu_time_us": d["gpu__time_duration.sum"][0],
            "kernel_source_sha256_16": bench.kernel_source_hash(), "kernel": kernel,
            "source": "ncu --set full --clock-control none of `python bench.py --steps 3 --warmup 3 --no-e2e "
                      "--no-cpu-baseline` (scripts/gpu_profile.sh); bench.py reports it only while the hash of "
                      "nxs_kernel.cuh + nxs_device.cuh matches"}
    json.dump(info, open(os.path.join(ROOT, "profiles", "traffic_per_launch.json"), "w"), indent=1)
    print(json.dumps(info))


if __name__ == "__main__":
    main()
