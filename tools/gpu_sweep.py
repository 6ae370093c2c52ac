"""Tuning sweep (GPU only): runs tools/gpu_full.py in a subprocess per environment-knob configuration.
Prints total / traverse / shade ms per scene."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.environ.get("SWEEP_SCENES", "cornell,cornell_noredirect,demo1,demo2,pawn,bunny_volume").split(",")
CONFIGS = [dict(kv.split("=") for kv in a.split(",") if kv) for a in sys.argv[1:]] or [{}]
for cfg in CONFIGS:
    env = dict(os.environ, **cfg)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gpu_full.py")] + SCENES, env=env, capture_output=True, text=True, timeout=600)
    res = {}
    for line in out.stdout.splitlines():
        if line.startswith("{"):
            d = json.loads(line)
            res[d["scene"]] = d
    if out.returncode != 0:
        print("FAILED", cfg, out.stderr[-500:])
    tag = ",".join(f"{k.replace('RT_B200_', '')}={v}" for k, v in cfg.items()) or "default"
    print(f"{tag:36s} " + " ".join(f"{s[:5]}:{res[s]['ms']:6.1f}/{res[s]['trav']:5.1f}/{res[s]['shade']:5.1f}" if s in res else f"{s[:5]}:   fail" for s in SCENES), flush=True)
