import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
from pyglasstools_b200 import _capi, synth, sharding
n = 2048
fr = synth.synth_frame(n, synth.SEEDS["cfg3"])
ctx = _capi.Context(0)
rc = 1.25
ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
ctx.set_cg("octic", 2.0, 5)
ctx.set_max_rcut(max(rc * float(fr["diameter"].max()), float(fr["diameter"].max())))
axis = synth.linspace_axis(float(fr["box"][0]), n)
ctx.set_grid_cartesian(axis, axis)
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
g0, g1 = sharding.grid_shards(n * n, world)[world // 2]
ctx.set_shard(g0, g1)
for rep in range(4):
    ctx.upload(fr["position"], fr["diameter"], fr["box"])
    ctx.build_cells(); ctx.build_bonds(); ctx.ik_compute(1); ctx.sync()
    print("rep", rep, {s: round(ctx.stage_ms(s), 4) for s in ("cells", "bonds", "ik")}, ctx.counts())
