"""DRAM traffic per launch of the per-centre kernels, from `ncu --set full` captures (.ncu-rep): writes
profiles/<tag>_traffic.json, which bench.py reads for `roofline.traffic` (bytes per centre x centres per launch).

    python tools/ncu_traffic.py <tag> <centres_per_captured_launch> name=path.ncu-rep [name=path.ncu-rep ...]

`name` is the bench stage the capture belongs to (factor_solve, factor_solve_core, predict, score, syrk); each entry records
dram__bytes_read.sum + dram__bytes_write.sum of the captured launch, its duration, the capture's file name and sha256."""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def raw_metrics(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    names, units, vals = rows[0], rows[1], rows[2]
    return {n: (u, v) for n, u, v in zip(names, units, vals)}


def main():
    tag, centres = sys.argv[1], int(sys.argv[2])
    out = {"centres_per_captured_launch": centres, "kernels": {}}
    for arg in sys.argv[3:]:
        name, path = arg.split("=", 1)
        m = raw_metrics(path)
        total = 0.0
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            u, v = m[key]
            total += float(v.replace(",", "")) * UNIT[u]
        du, dv = m["gpu__time_duration.sum"]
        ms = float(dv.replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(du, 1.0)
        out["kernels"][name] = {"dram_bytes": total, "dram_bytes_per_centre": total / centres, "duration_ms": ms,
                                "capture": os.path.basename(path),
                                "sha256": hashlib.sha256(open(path, "rb").read()).hexdigest()}
    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", f"{tag}_traffic.json")
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
    print(dst, {k: round(v["dram_bytes_per_centre"] / 1e6, 2) for k, v in out["kernels"].items()}, "MB per centre")


if __name__ == "__main__":
    main()
