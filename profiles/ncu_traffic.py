"""Refresh profiles/NCU_TRAFFIC.json from `ncu --set full` captures of THIS round (bench.py reports these numbers as
`roofline.traffic`, so they must come from a capture, not from a hand-typed literal).

usage: python profiles/ncu_traffic.py <tag> <key>=<file.ncu-rep>:<kernel-name-substring> [...]
  e.g. python profiles/ncu_traffic.py r2a k_rollout_f64_65536x128=gpurun_out/roll.ncu-rep:k_rollout \
                                         wide_dominant_kernel=gpurun_out/wide.ncu-rep:k_gemm_x3
Per key: mean over the matching launches of dram__bytes_read.sum + dram__bytes_write.sum (bytes per launch)."""
import csv
import json
import os
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def launches(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r]
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[hi], rows[hi + 1]
    u = dict(zip(hdr, units))
    for r in rows[hi + 2:]:
        d = dict(zip(hdr, r))
        b = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            b += float(d[k].replace(",", "")) * UNIT.get(u.get(k, "byte"), 1.0)
        yield d["Kernel Name"], b, float(d["gpu__time_duration.sum"].replace(",", "")), u.get("gpu__time_duration.sum")


def main():
    tag, specs = sys.argv[1], sys.argv[2:]
    path = os.path.join(HERE, "NCU_TRAFFIC.json")
    tj = {"_doc": "bytes per launch = dram__bytes_read.sum + dram__bytes_write.sum, mean over the matching launches of an "
                  "`ncu --set full --clock-control none` capture; written by profiles/ncu_traffic.py, never by hand"}
    try:
        old = json.load(open(path))
        tj.update({k: v for k, v in old.items() if isinstance(v, dict)})
    except Exception:
        pass
    tj["_captured"] = f"{tag}, {time.strftime('%Y-%m-%d')}"
    for spec in specs:
        key, rest = spec.split("=", 1)
        rep, pat = rest.rsplit(":", 1)
        hits = [(n, b, t, tu) for n, b, t, tu in launches(rep) if pat in n]
        if not hits:
            print(f"{key}: no launch matches {pat!r} in {rep}", file=sys.stderr)
            continue
        tj[key] = {"bytes": int(sum(h[1] for h in hits) / len(hits)), "launches": len(hits),
                   "kernel": hits[0][0][:120], "mean_duration": sum(h[2] for h in hits) / len(hits), "duration_unit": hits[0][3],
                   "source": f"{os.path.basename(rep)} ({tag})"}
        print(key, tj[key])
    json.dump(tj, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
