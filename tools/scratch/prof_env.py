import os, sys, time, cProfile, pstats, io
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub"), ROOT]
import torch, vmas
from vectorizedmultiagentsimulator_b200 import dropin
scenario = sys.argv[1] if len(sys.argv) > 1 else "navigation"
static_io = len(sys.argv) > 2 and sys.argv[2] == "1"
kw = dict(n_agents=3 if scenario == "simple_spread" else 4)
n = 32768
env = vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True, **kw)
dropin.enable(env, hooks=True, static_io=static_io)
acts = [[torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in env.agents] for _ in range(4)]
for i in range(20): env.step(acts[i % 4])
torch.cuda.synchronize(); t0 = time.perf_counter()
K = 300
for i in range(K): env.step(acts[i % 4])
torch.cuda.synchronize(); print(scenario, "static_io", static_io, "env.step %.3f ms" % ((time.perf_counter() - t0) / K * 1e3))
pr = cProfile.Profile(); pr.enable()
for i in range(K): env.step(acts[i % 4])
torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45); print(s.getvalue()[:9000])
