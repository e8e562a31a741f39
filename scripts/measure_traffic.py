"""Re-measure roofline.traffic: DRAM bytes (read + write) of ONE lift_kernel launch per workload, from an ncu pass
(`--metrics dram__bytes_read.sum,dram__bytes_write.sum`), written to profiles/lift_traffic.json together with the launch
time ncu saw and the git revision. bench.py copies the figure of its workload into the JSON line; a stale file would
silently misreport, so scripts/gpu_round.sh runs this on every GPU visit that changes the kernel.

    python scripts/measure_traffic.py c2 c3 c4          (on the GPU box; needs ncu on PATH)
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def measure(workload: str):
    log = os.path.join(ROOT, "gpurun_out", f"traffic_{workload}.csv")
    cmd = ["ncu", "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum", "--clock-control", "none",
           "--profile-from-start", "off", "-k", "regex:lift_kernel", "-c", "1", "--csv", "--log-file", log,
           sys.executable, os.path.join(ROOT, "bench.py"), "--workload", workload, "--steps", "1", "--warmup", "3", "--profile"]
    subprocess.run(cmd, check=True, cwd=ROOT, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=900)
    rows = [r for r in csv.reader(open(log)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    name, unit, val = hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "usecond": 1, "msecond": 1e3, "nsecond": 1e-3, "ms": 1e3}
    got = {}
    for r in rows[1:]:
        got[r[name]] = float(r[val].replace(",", "")) * scale.get(r[unit], 1)
    return {"dram_bytes": int(got["dram__bytes_read.sum"] + got["dram__bytes_write.sum"]), "read": int(got["dram__bytes_read.sum"]),
            "write": int(got["dram__bytes_write.sum"]), "ncu_us": got["gpu__time_duration.sum"]}


def main():
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "profiles", "lift_traffic.json")
    data = json.load(open(path)) if os.path.exists(path) else {}
    detail = data.get("_detail", {})
    for wl in sys.argv[1:] or ["c2"]:
        m = measure(wl)
        data[wl] = m["dram_bytes"]
        detail[wl] = m
        print(wl, m)
    try:
        rev = subprocess.check_output(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, text=True).strip()
    except Exception:
        rev = os.environ.get("BFF_GIT_REV", "unknown (no .git on the GPU box)")
    data["_detail"] = detail
    data["_measured_at"] = rev
    json.dump(data, open(os.path.join(ROOT, "gpurun_out", "lift_traffic.json"), "w"), indent=1)
    print("wrote gpurun_out/lift_traffic.json (copy to profiles/)")


if __name__ == "__main__":
    main()
