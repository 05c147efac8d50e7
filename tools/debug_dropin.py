import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub")]
import vmas
from vectorizedmultiagentsimulator_b200 import dropin
scenario = sys.argv[1] if len(sys.argv) > 1 else "joint_passage"
n = 96
mk = lambda: vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True)
ref, env = mk(), mk()
b = dropin.enable(env)
st = lambda e: torch.stack([torch.cat([x.state.pos, x.state.vel, x.state.rot, x.state.ang_vel], -1) for x in e.world.entities])
torch.manual_seed(7)
for t in range(26):
    if t == 20:
        torch.manual_seed(11); ref.reset_at(5); torch.manual_seed(11); env.reset_at(5); torch.manual_seed(12)
        d = (st(ref) - st(env)).abs()
        print("after reset_at: max state diff", float(d.max()), "copies so far", b.n_copies, "dirty", b.dirty)
    acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in ref.agents]
    ref.step([a.clone() for a in acts]); env.step([a.clone() for a in acts])
    d = (st(ref) - st(env)).abs()
    if t >= 18:
        per_ent = d.amax(dim=(1, 2)); per_env = d.amax(dim=(0, 2))
        bad_e = [(ref.world.entities[i].name, float(per_ent[i])) for i in torch.nonzero(per_ent > 1e-3).flatten().tolist()]
        bad_n = torch.nonzero(per_env > 1e-3).flatten().tolist()
        print(t, "max", float(d.max()), "entities", bad_e[:6], "envs", bad_n[:10], "copies", b.n_copies)
