#!/usr/bin/env python3
"""Pipelines per frame (PTC_DUAL = 0 / 1) by frame size on the adopted schedule.  usage: sweep_dual.py"""
import json, os, subprocess, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
cases = [("indoor_tex", 1280, 720, 3), ("indoor_tex", 1920, 1080, 3), ("indoor_tex", 2720, 1530, 3), ("indoor_tex", 3840, 2160, 3),
         ("indoor_tex", 1920, 1080, 8), ("indoor_tex", 3840, 2160, 8), ("color_sample_plain", 1920, 1080, 4), ("synth2236", 1920, 1080, 4)]
for (sc, w, h, b) in cases:
    row = {"scene": sc, "w": w, "h": h, "b": b}
    for dual in ("0", "1"):
        env = dict(os.environ, PTC_DUAL=dual)
        n = "8" if sc.startswith("synth") else str(max(8, min(64, int(40e6 / (w * h)))))
        out = subprocess.run([sys.executable, os.path.join(REPO, "tools", "probe.py"), sc, n, str(w), str(h), str(b)],
                             env=env, capture_output=True, text=True).stdout.strip().splitlines()
        try:
            row["dual" + dual] = json.loads(out[-1])["ms_frame"]
        except Exception:
            row["dual" + dual] = None
    print(json.dumps(row), flush=True)
