import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub"), ROOT]
import torch, vmas
from vectorizedmultiagentsimulator_b200 import dropin
HOOKS = len(sys.argv) > 1 and sys.argv[1] == "hooks"
for scenario, kw in (("navigation", dict(n_agents=4)), ("transport", dict(n_agents=4)), ("football", dict(ai_red_agents=False))):
    n = 32768
    env = vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True, **kw)
    dropin.enable(env, hooks=HOOKS)
    T = {}
    def wrap(obj, name, key):
        f = getattr(obj, name)
        def g(*a, **k):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            r = f(*a, **k)
            torch.cuda.synchronize(); T[key] = T.get(key, 0.0) + time.perf_counter() - t0
            return r
        setattr(obj, name, g)
    wrap(env, "_set_action", "set_action")
    wrap(env.scenario, "env_process_action", "process_action")
    wrap(env.world, "step", "world_step")
    wrap(env, "_get_from_scenario", "hooks")
    acts = [[torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in env.agents] for _ in range(4)]
    for i in range(5): env.step([a.clone() for a in acts[i % 4]])
    T.clear(); K = 30
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(K): env.step([a.clone() for a in acts[i % 4]])
    torch.cuda.synchronize(); tot = (time.perf_counter() - t0) / K * 1e3
    print("hooks" if HOOKS else "plain", scenario, "total %.3f ms" % tot, {k: round(v / K * 1e3, 3) for k, v in T.items()})
