import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from raytrace_b200 import scenes
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene
s, w = scenes.demo2(width=800, spp=1)
ids, ts = CompiledScene(w).probe_primary(s)
oids, ots, stable = OracleScene(w).probe_stable(s)
both = stable & (ids == oids) & (oids != 0xFFFFFFFF)
rel = np.where(both, np.abs(ts - ots) / np.where(both, ots, 1), 0)
order = np.argsort(rel.ravel())[::-1][:15]
for k in order:
    j, i = divmod(int(k), 800)
    print(i, j, 'prim', ids[j, i], 'gpu t', ts[j, i], 'oracle t', ots[j, i], 'rel', rel[j, i])
for p in np.unique(ids[both]):
    m = both & (ids == p)
    if rel[m].max() > 2e-6: print('prim', p, 'n', m.sum(), 'max rel', rel[m].max(), 'mean', rel[m].mean())
