"""Per-call host-side timing of the step paths (wall clock, 4096 default envs)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sparc_b200 import params as P
from sparc_b200.engine import Engine

cfg = dict(workpiece_height=15.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)
eng = Engine(4096, P.derive_params(P.EnvironmentConfig(**cfg), "position"), device=torch.device("cuda", 0), seed=1)
eng.set_actions(0.1, 80.0, 5, 3.0, 80.0)
ha = eng.host_actions
ha["servo"].fill_(0.1); ha["target_voltage"].fill_(80.0); ha["current_mode"].fill_(5); ha["on_time"].fill_(3.0); ha["off_time"].fill_(80.0)
def t(fn, n=200):
    eng.reset()                      # same simulated window for every variant
    for _ in range(10): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
print("step (device, async, warm L2)      %.1f us" % t(lambda: eng.step(1000)))
print("step + fused observe               %.1f us" % t(lambda: eng.step(1000, observe=True)))
print("step + observe kernel              %.1f us" % t(lambda: (eng.step(1000), eng.observe())))
print("step + sync each call              %.1f us" % t(lambda: (eng.step(1000), torch.cuda.synchronize())))
print("step_host                          %.1f us" % t(lambda: eng.step_host(1000)))
sv = ha["servo"].numpy()
def user():
    sv.fill(0.1)
    o = eng.step_host(1000).numpy()
    return float(o[3].sum())
print("step_host + numpy fill + obs sum   %.1f us" % t(user))
