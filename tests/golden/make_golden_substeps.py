#!/usr/bin/env python
"""Per-SUBSTEP golden fixtures (tests/golden/substeps/<name>.npz) from the UNMODIFIED reference.

TEST INFRASTRUCTURE.  Run in the build container (needs oracle/_ref, torch CPU fp32):

    python tests/golden/make_golden_substeps.py            # all
    python tests/golden/make_golden_substeps.py waterfall  # one

One env-step of a stiff jointed world is 10 substeps of a spring with joint_force 900 on moments of inertia of ~5e-4:
a 1-ulp difference after substep 1 is ~1e-3 after substep 10, so a whole-step comparison cannot pin the joint
arithmetic at the stated fp32 bar.  Here the reference's `World.step` (core.py:1971-2014) is
instrumented INSIDE its substep loop, on the same seeded rollouts as tests/golden/make_golden.py:

* the state of every entity at the entry of every substep and after its `_integrate_state` (core.py:2861-2907),
* `agent.state.force / torque` at the entry of every substep (the clamped write-back of substep s is the input of s+1),
* `forces_dict / torques_dict` (core.py:1975-1992) as `_integrate_state` reads them: the total force and torque on every
  entity in that substep,
* `trig` / `jexp`: what this torch returns for the cos / sin / exp calls of the substep (utils.py:183-184,
  physics.py:299-301, 412; core.py:2852) on the substep's input rotations.  torch's unary CPU kernels are Intel MKL VML's
  (closed source, not correctly rounded: ~5 % of results are 1 ulp off libm's / CUDA's), and a stiff joint amplifies exactly
  that ulp; with these values handed in, the oracle reproduces every substep BIT for bit (tests/test_oracle_substeps.py),
* the four values `_vectorized_joint_constraints` (core.py:2200-2291) hands to `update_env_forces` for every joint:
  force / torque on entity_a and on entity_b - the constraint force (core.py:2804-2838) and torque (:2840-2857) directly.

Consumers start every substep from the reference's own substep input, so nothing is amplified between the comparison
points.  The collision masks of the substep are those of the main fixture (tests/golden/<name>.npz, masks[t][s]).
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as MG  # noqa: E402  (sets sys.path for oracle/_ref)

import vmas  # noqa: E402

from oracle.oracle import Oracle  # noqa: E402
from vectorizedmultiagentsimulator_b200.worlddesc import WorldDesc  # noqa: E402

# env-steps whose substeps are recorded
STEPS = [0, 1, 10, 25, 50]
NAMES = ["joint_passage", "asym_joint", "waterfall", "transport_s5", "football", "navigation", "pollock", "balance", "wheel"]


def run(name, cfg):
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    n, steps = cfg["n_envs"], cfg["steps"]
    env = vmas.make_env(cfg["scenario"], num_envs=n, device="cpu", seed=0, continuous_actions=True, **cfg["kw"])
    world = env.world
    if "substeps" in cfg:
        world._substeps = cfg["substeps"]
        world._sub_dt = world._dt / world._substeps
    env.reset(seed=0)
    desc = WorldDesc.from_world(world)
    E, S, J = len(desc.entities), desc.substeps, len(desc.joints)
    ents = list(world.entities)
    eidx = {id(e): i for i, e in enumerate(ents)}
    # joints in the order the oracle / the plan meet them (pair enumeration a<b, core.py:2111-2121)
    jkey = {}
    for a in range(E):
        for b in range(a + 1, E):
            j = world._joints.get(frozenset({ents[a].name, ents[b].name}), None)
            if j is not None:
                jkey[id(j)] = len(jkey)
    assert len(jkey) == J

    want = [t for t in STEPS if t < steps]
    rec = dict(t=-1, on=False, sub_in=[], sub_out=[], act_f=[], act_t=[], tot_f=[], tot_t=[], jf=[], jt=[], in_joints=False)
    orig_integrate = world._integrate_state
    orig_joints = world._vectorized_joint_constraints
    orig_update = world.update_env_forces
    orig_step = world.step
    orig_env_force = world._apply_vectorized_enviornment_force

    def env_force():
        # first call inside a substep after the per-entity forces: the state is still the substep's input
        if rec["on"]:
            rec["sub_in"].append(MG.state_of(world))
        return orig_env_force()

    def integrate(entity, substep):
        if rec["on"] and eidx[id(entity)] == 0:
            # agent.state.force / torque as this substep's _apply_action_force left them = what it read, clamped once more;
            # the value it READ is recorded by `step` (substep 0) or is the previous substep's write-back
            rec["tot_f"].append(np.stack([world.forces_dict[e].detach().clone().numpy() for e in ents], 0))
            rec["tot_t"].append(np.stack([world.torques_dict[e].detach().clone().numpy()[:, 0] for e in ents], 0))
            f, t = MG.agent_ft(world)
            rec["act_f"].append(f)
            rec["act_t"].append(t)
        r = orig_integrate(entity, substep)
        if rec["on"] and eidx[id(entity)] == E - 1:
            rec["sub_out"].append(MG.state_of(world))
        return r

    def joints(js):
        if rec["on"]:
            rec["cur_jf"] = np.zeros((J, 2, n, 2), np.float32)
            rec["cur_jt"] = np.zeros((J, 2, n), np.float32)
            rec["order"] = [jkey[id(j)] for j in js]
            rec["k"] = 0
            rec["in_joints"] = True
        r = orig_joints(js)
        if rec["on"]:
            rec["in_joints"] = False
            rec["jf"].append(rec["cur_jf"])
            rec["jt"].append(rec["cur_jt"])
        return r

    def update(ea, fa, ta, eb, fb, tb):
        if rec["on"] and rec["in_joints"]:
            j = rec["order"][rec["k"]]
            rec["k"] += 1
            rec["cur_jf"][j, 0], rec["cur_jf"][j, 1] = fa.detach().numpy(), fb.detach().numpy()
            rec["cur_jt"][j, 0], rec["cur_jt"][j, 1] = ta.detach().numpy()[:, 0], tb.detach().numpy()[:, 0]
        return orig_update(ea, fa, ta, eb, fb, tb)

    entry = dict(f=[], t=[])

    def step():
        rec["t"] += 1
        rec["on"] = rec["t"] in want
        if rec["on"]:
            f, t = MG.agent_ft(world)
            entry["f"].append(f)
            entry["t"].append(t)
        orig_step()
        rec["on"] = False

    world._integrate_state = integrate
    world._vectorized_joint_constraints = joints
    world.update_env_forces = update
    world._apply_vectorized_enviornment_force = env_force
    world.step = step

    for t in range(max(want) + 1):
        actions = [env.get_random_action(a) for a in env.agents]
        env.step(actions)

    K = len(want)
    sub_in = np.stack(rec["sub_in"], 0).reshape(K, S, E, n, 6).astype(np.float32)
    sub_out = np.stack(rec["sub_out"], 0).reshape(K, S, E, n, 6).astype(np.float32)
    assert (sub_in[:, 1:] == sub_out[:, :-1]).all()
    A = len(world.agents)
    act_after_f = np.stack(rec["act_f"], 0).reshape(K, S, A, n, 2).astype(np.float32)
    act_after_t = np.stack(rec["act_t"], 0).reshape(K, S, A, n).astype(np.float32)
    # input of substep s: the step's entry value for s = 0, else the write-back of s - 1
    act_in_f = np.concatenate([np.stack(entry["f"], 0)[:, None], act_after_f[:, :-1]], 1).astype(np.float32)
    act_in_t = np.concatenate([np.stack(entry["t"], 0)[:, None], act_after_t[:, :-1]], 1).astype(np.float32)
    orc = Oracle(desc)
    main = np.load(os.path.join(HERE, f"{name}.npz"))
    jrot_steps, jrot_vals = main["jrot_override_steps"], main["jrot_override_values"]
    trig = np.zeros((K, S, E, n, 4), np.float32)
    jexp = np.zeros((K, S, J, n), np.float32)
    for k, t in enumerate(want):
        jr = main["jrot0"]
        for st_, v in zip(jrot_steps, jrot_vals):
            if st_ <= t:
                jr = v
        for s_ in range(S):
            trig[k, s_], jexp[k, s_] = orc.torch_math(sub_in[k, s_], jr)
    cap = torch.backends.cpu.get_cpu_capability()
    data = dict(
        meta_json=np.array(json.dumps(dict(name=name, torch=torch.__version__, steps=want, substeps=S, n_envs=n,
                                           cpu_capability=cap, vec_width={"AVX512": 32, "AVX2": 16}.get(cap, 0)))),
        trig=trig, jexp=jexp,
        steps=np.array(want, np.int32),
        states=np.concatenate([sub_in, sub_out[:, -1:]], 1),    # [K, S+1, E, N, 6]: input of substep s, output of s-1
        act_force_in=act_in_f, act_torque_in=act_in_t,          # [K, S, A, N, 2] / [K, S, A, N]
        act_force_out=act_after_f, act_torque_out=act_after_t,
        total_force=np.stack(rec["tot_f"], 0).reshape(K, S, E, n, 2).astype(np.float32),
        total_torque=np.stack(rec["tot_t"], 0).reshape(K, S, E, n).astype(np.float32),
        joint_force=(np.stack(rec["jf"], 0).reshape(K, S, J, 2, n, 2) if J else np.zeros((K, S, 0, 2, n, 2), np.float32)),
        joint_torque=(np.stack(rec["jt"], 0).reshape(K, S, J, 2, n) if J else np.zeros((K, S, 0, 2, n), np.float32)),
    )
    os.makedirs(os.path.join(HERE, "substeps"), exist_ok=True)
    path = os.path.join(HERE, "substeps", f"{name}.npz")
    np.savez_compressed(path, **data)
    # the rollout must be the one of the main fixture
    assert (main["state0"] == data["states"][0, 0]).all(), "not the rollout of tests/golden/%s.npz" % name
    print(f"{name}: steps {want} S={S} E={E} J={J} -> {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    for nm in sys.argv[1:] or NAMES:
        run(nm, MG.CONFIGS[nm])
