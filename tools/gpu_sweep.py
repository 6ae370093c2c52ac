"""Tuning sweep (GPU only): runs tools/gpu_full.py in a subprocess per RtOptions configuration.
usage: python tools/gpu_sweep.py "" "bvh_reference=1" "node_burst=2,refill=16" ...      (SWEEP_SCENES=a,b,c picks the scenes)
Prints total / traverse / shade ms per scene."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.environ.get("SWEEP_SCENES", "cornell,cornell_noredirect,demo1,demo2,pawn,bunny_volume").split(",")
CONFIGS = sys.argv[1:] or [""]
for cfg in CONFIGS:
    cmd = [sys.executable, os.path.join(ROOT, "tools", "gpu_full.py")] + (["--opts", cfg] if cfg else []) + SCENES
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    res = {}
    for line in out.stdout.splitlines():
        if line.startswith("{"):
            d = json.loads(line)
            res[d["scene"]] = d
    if out.returncode != 0:
        print("FAILED", cfg, out.stderr[-500:])
    print(f"{cfg or 'default':36s} " + " ".join(f"{s[:5]}:{res[s]['ms']:6.1f}/{res[s]['trav']:5.1f}/{res[s]['shade']:5.1f}" if s in res else f"{s[:5]}:   fail" for s in SCENES), flush=True)
