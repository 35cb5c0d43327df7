"""Cost of the in-kernel signal log: ms per 1000-us control step at 4096 envs, plain vs sampled every k us."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sparc_b200 import params as P
from sparc_b200.engine import Engine
E = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
prm = P.derive_params(P.EnvironmentConfig(workpiece_height=15.0, wire_diameter=0.25, target_cutting_distance=800.0), "position")
eng = Engine(E, prm, seed=1)
eng.set_actions(0.1, 80.0, 5, 3.0, 80.0)
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
base = timed(lambda: eng.step(1000))
print(f"plain step            {base:.3f} ms")
for k in (1, 10, 100, 1000):
    t = timed(lambda: eng.step_sampled(1000, k))
    print(f"sampled every {k:4d} us {t:.3f} ms  ({t / base:.2f}x plain; includes the torch.zeros of the row buffer)")
