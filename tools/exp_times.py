"""Experiment: per-env start/end times (build with -DSPARC_DBG_TIMES): are the warps of one scheduler served fairly?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sparc_b200 import params as P
from sparc_b200.engine import Engine
E = int(sys.argv[1]) if len(sys.argv) > 1 else 2368
prm = P.derive_params(P.EnvironmentConfig(workpiece_height=15.0, wire_diameter=0.25, target_cutting_distance=800.0), "position")
eng = Engine(E, prm, seed=1)
eng.set_actions(0.1, 80.0, 5, 3.0, 80.0)
for _ in range(5):
    eng.step(1000)
torch.cuda.synchronize()
tr = eng["mech_traj"].view(E, 1024)[:, 1016:1020].cpu().numpy()
t0, t1, sm, wid = tr[:, 0], tr[:, 1], tr[:, 2].astype(int), tr[:, 3].astype(int)
base = t0.min()
print("launch span us", (t1.max() - base) / 1e3, "start spread us", (t0.max() - base) / 1e3)
dur = (t1 - t0) / 1e3
print("env duration us: min %.1f median %.1f max %.1f" % (dur.min(), np.median(dur), dur.max()))
end = (t1 - base) / 1e3
print("end time us: p5 %.1f p25 %.1f p50 %.1f p75 %.1f p95 %.1f max %.1f" % tuple(np.percentile(end, [5, 25, 50, 75, 95, 100])))
# per scheduler: (sm, wid % 4)
import collections
g = collections.defaultdict(list)
for e in range(E):
    g[(sm[e], wid[e] % 4)].append((end[e], (t0[e] - base) / 1e3, wid[e]))
k = 0
for key in sorted(g)[:6]:
    print(key, ["start %.0f end %.0f w%d" % (b, a, w) for a, b, w in sorted(g[key])])
cnt = collections.Counter(len(v) for v in g.values())
print("warps per scheduler histogram", sorted(cnt.items()))
