"""north_star gates 1 + 2 at BASELINE.json's full sizes over ALL pixels, for the default (SAH) device tree and the
reference-shaped one, with every pixel that does not pass listed (GPU box only; the committed evidence is
profiles/r02*_parity_id_t_gates.txt).  usage: python tools/gpu_id_gate.py [--head GITHASH] [scene ...]"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from raytrace_b200 import scenes
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene, primary_gate

CASES = {"readme": dict(), "cornell": dict(spp=1), "demo1": dict(spp=1), "demo2": dict(width=800, spp=1),
         "pawn": dict(width=1920, aspect=16 / 9, spp=1), "bunny_volume": dict(spp=1)}
argv = sys.argv[1:]
head = ""
if argv and argv[0] == "--head":      # the GPU box has no .git: the caller passes `git rev-parse --short HEAD`
    head, argv = argv[1], argv[2:]
for name in (argv or list(CASES)):
    ctor, _ = scenes.CONFIGS[name]
    s, w = ctor(**CASES[name])
    ids5, ts5 = OracleScene(w).probe_tilted(s)
    for tree in ("sah", "reference"):
        ids, ts = CompiledScene(w, options=dict(bvh_reference=int(tree == "reference"))).probe_primary(s)
        g = primary_gate(ids, ts, ids5, ts5)
        bad = np.argwhere(~g.pop("passed_mask"))
        print(json.dumps(dict(scene=name, tree=tree, head=head, **g, gate="PASS" if g["pass_frac"] >= 0.9999 else "FAIL")), flush=True)
        if tree == "sah":
            for j, i in bad[:40]:
                print(f"    px ({i},{j}): gpu id {ids[j, i]} t {ts[j, i]:.7g} | oracle centre/tilts ids {ids5[:, j, i].tolist()} ts {[float(f'{x:.7g}') for x in ts5[:, j, i]]}")
