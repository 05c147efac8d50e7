import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub")]
import vmas
from vectorizedmultiagentsimulator_b200 import dropin
from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment
scenario = sys.argv[1] if len(sys.argv) > 1 else "navigation"
n = 160
mk = lambda: vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True, max_steps=25, n_agents=4)
e1, e2 = mk(), mk()
dropin.enable(e1, static_io=True, hooks=True)
g = GraphedEnvironment(e2, warmup=2)
rs = type(e1).vmas_random_state
saved = list(rs); e1.reset(seed=1); rs[:] = saved; e2.reset(seed=1)
st = lambda e: torch.stack([torch.cat([x.state.pos, x.state.vel, x.state.rot, x.state.ang_vel], -1) for x in e.world.entities])
def leaves(x, pre=""):
    if isinstance(x, dict): return [t for k in sorted(x) for t in leaves(x[k], pre + "." + str(k))]
    if isinstance(x, (list, tuple)): return [t for i, v in enumerate(x) for t in leaves(v, pre + "[%d]" % i)]
    return [(pre, x)]
print("after reset state diff", float((st(e1) - st(e2)).abs().max()))
torch.manual_seed(7)
for t in range(6):
    acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in e1.agents]
    r1 = e1.step([a.clone() for a in acts]); r2 = g.step(acts)
    bad = [(k, float((a.float() - c.float()).abs().nan_to_num().max()), int((a != c).reshape(n, -1).any(1).sum())) for (k, a), (_, c) in zip(leaves(r1), leaves(r2)) if not torch.equal(a, c)]
    print(t, "state diff", float((st(e1) - st(e2)).abs().max()), "bad", bad[:10], "fixes", g.n_input_fixes)
    for i, (a1, a2) in enumerate(zip(e1.world.agents, e2.world.agents)):
        for k in ("pos_shaping", "distance_to_goal", "agent_collision_rew"):
            d = float((getattr(a1, k) - getattr(a2, k)).abs().max())
            if d: print("   agent", i, k, d)
        d = float((a1.goal.state.pos - a2.goal.state.pos).abs().max())
        if d: print("   agent", i, "goal", d)
