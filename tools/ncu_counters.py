#!/usr/bin/env python3
"""ncu --csv (metrics pass over tools/frames_only.py) -> one entry of profiles/r2_counters.json.
usage: ncu_counters.py <csv> <scene> <frames> <w> <h> <bounces> <json to update>"""
import csv, json, os, sys
path, scene, frames, w, h, b, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]), sys.argv[7]
rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
hdr = next(r for r in rows if "Metric Name" in r)
ci = {n: i for i, n in enumerate(hdr)}
tot, per_kernel, launches = {}, {}, set()
for r in rows:
    if r is hdr or len(r) <= ci["Metric Value"] or r[ci["ID"]] == "ID":
        continue
    k = r[ci["Kernel Name"]]
    if not any(s in k for s in ("k_path", "k_extend", "k_wave")):
        continue
    name = r[ci["Metric Name"]]
    v = float(r[ci["Metric Value"]].replace(",", ""))
    unit = r[ci["Metric Unit"]].lower()
    v *= {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3,
          "ns": 1e-9, "nsecond": 1e-9}.get(unit, 1.0)
    tot[name] = tot.get(name, 0.0) + v
    short = k.split("(")[0].replace("void ptc::", "")
    d = per_kernel.setdefault(short, {})
    d[name] = d.get(name, 0.0) + v
    launches.add(r[ci["ID"]])
entry = {"warp_instructions_per_frame": tot["smsp__inst_executed.sum"] / frames,
         "dram_bytes_per_frame": (tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"]) / frames,
         "dram_read_bytes_per_frame": tot["dram__bytes_read.sum"] / frames,
         "kernel_seconds_per_frame_under_ncu": tot.get("gpu__time_duration.sum", 0.0) / frames,
         "launches_per_frame": len(launches) / frames,
         "per_kernel_per_frame": {k: {m: v / frames for m, v in d.items()} for k, d in per_kernel.items()},
         "source": f"profiles/{os.path.basename(path)} (ncu --metrics, {frames} frames of tools/frames_only.py)"}
table = json.load(open(out)) if os.path.exists(out) else {
    "_comment": "per-frame counters of the stage kernels from ncu metric passes (tools/ncu_counters.py); bench.py's "
                "issue bound and roofline.traffic read them -- committed captures, not measured inside bench.py"}
table[f"{scene} {w}x{h} b={b}"] = entry
json.dump(table, open(out, "w"), indent=1)
print(json.dumps({f"{scene} {w}x{h} b={b}": {k: v for k, v in entry.items() if k != "per_kernel_per_frame"}}))
