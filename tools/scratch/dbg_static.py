import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub")]
import vmas
from vectorizedmultiagentsimulator_b200 import dropin
scenario = sys.argv[1] if len(sys.argv) > 1 else "joint_passage"
do_reset = (sys.argv[2] if len(sys.argv) > 2 else "1") == "1"
n = 96
mk = lambda: vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True)
e1, e2 = mk(), mk()
b1 = dropin.enable(e1); b2 = dropin.enable(e2, static_io=True)
st = lambda e: torch.stack([torch.cat([x.state.pos, x.state.vel, x.state.rot, x.state.ang_vel], -1) for x in e.world.entities])
def leaves(x):
    if isinstance(x, dict): return [t for k in sorted(x) for t in leaves(x[k])]
    if isinstance(x, (list, tuple)): return [t for v in x for t in leaves(v)]
    return [x]
torch.manual_seed(7)
for t in range(30):
    if t == 10:
        e1.reset(seed=3); e2.reset(seed=3)
    if t == 20 and do_reset:
        torch.manual_seed(11); e1.reset_at(5); torch.manual_seed(11); e2.reset_at(5); torch.manual_seed(12)
        d = (st(e1) - st(e2)).abs()
        print("after reset_at: max state diff", float(d.nan_to_num().max()), "copies", b1.n_copies, b2.n_copies, "desc same", b1.desc_json == b2.desc_json)
    acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in e1.agents]
    r1 = e1.step([a.clone() for a in acts]); r2 = e2.step([a.clone() for a in acts])
    d = (st(e1) - st(e2)).abs().nan_to_num()
    per_ent = d.amax(dim=(1, 2)); per_env = d.amax(dim=(0, 2))
    bad_e = [(e1.world.entities[i].name, float(per_ent[i])) for i in torch.nonzero(per_ent > 0).flatten().tolist()]
    bad_n = torch.nonzero(per_env > 0).flatten().tolist()
    badl = [(i, float((a.float() - c.float()).abs().nan_to_num().max()), torch.nonzero((a != c).reshape(n, -1).any(1)).flatten().tolist()[:6]) for i, (a, c) in enumerate(zip(leaves(r1), leaves(r2))) if not torch.equal(a.nan_to_num(), c.nan_to_num())]
    if bad_e or badl or t in (19, 20):
        print(t, "state max", float(d.max()), "entities", bad_e[:6], "envs", bad_n[:10], "leaves", badl[:8], "repairs", b1.plan.repairs() if hasattr(b1.plan, "repairs") else None)
