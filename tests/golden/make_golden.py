#!/usr/bin/env python
"""Generates the golden parity fixtures under tests/golden/ by running the UNMODIFIED reference.

TEST INFRASTRUCTURE.  Run in the build container (needs oracle/_ref, made by oracle/make_ref.sh from
/root/reference; torch CPU fp32):

    python tests/golden/make_golden.py            # all fixtures
    python tests/golden/make_golden.py transport  # one

For every configuration the reference ``Environment`` (vmas/simulator/environment/environment.py:325) is
stepped with random continuous actions on CPU while ``World.step`` (core.py:1971) is instrumented:

* the state of every entity right before step 0 and right after selected steps,
* ``agent.state.force/torque`` as they enter ``World.step`` (after ``env_process_action``) for every step,
  and their clamped write-back (core.py:2017-2040) after selected steps,
* every ``World.collides`` verdict (core.py:2787-2802) issued from inside ``World.step`` - the per (substep,
  pair) "collision mask",
* ``Lidar._last_measurement`` (sensors.py:100-122) and ``scenario.done()`` after selected steps,
* any state the scenario rewrites between two ``World.step`` calls (goal scored in football, ...), so a
  replay can apply the same overrides.

Each fixture also embeds the JSON ``WorldDesc`` so that the oracle / CUDA tests need no reference install.
The torch version is recorded: the reference has no golden vectors of its own, so parity is pinned
operationally to this run.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub"), ROOT]

import vmas  # noqa: E402

from vectorizedmultiagentsimulator_b200.worlddesc import WorldDesc  # noqa: E402

SNAP = [1, 2, 10, 11, 25, 50, 51, 99, 100]

CONFIGS = {
    # BASELINE.json configs[0..4]
    "simple_spread": dict(scenario="simple_spread", kw=dict(n_agents=3), n_envs=32, steps=100),
    "navigation": dict(scenario="navigation", kw=dict(n_agents=4), n_envs=32, steps=100),
    "transport": dict(scenario="transport", kw=dict(n_agents=4), n_envs=32, steps=100),
    "transport_s5": dict(scenario="transport", kw=dict(n_agents=4), n_envs=32, steps=100, substeps=5),
    # dt / substeps and 1 - drag are formed in Python doubles before they meet fp32 tensors (core.py:1124, 2866): values whose
    # fp32-first evaluation would be 1 ulp off (dt 0.01 / 5 substeps, drag 0.6)
    "transport_dt": dict(scenario="transport", kw=dict(n_agents=4), n_envs=16, steps=30, substeps=5, dt=0.01, drag=0.6),
    "joint_passage": dict(scenario="joint_passage", kw=dict(), n_envs=16, steps=100),
    "football": dict(scenario="football", kw=dict(ai_red_agents=False), n_envs=32, steps=100),
    # coverage of the remaining pair types / force terms
    "pollock": dict(scenario="pollock", kw=dict(lidar=True), n_envs=12, steps=30),          # all six pair types + LIDAR (tests/test_lidar.py)
    "waterfall": dict(scenario="waterfall", kw=dict(), n_envs=16, steps=60),     # joints with rotate=False, box-box
    "balance": dict(scenario="balance", kw=dict(), n_envs=16, steps=60),          # world gravity, line-sphere, y bound
    "give_way": dict(scenario="give_way", kw=dict(), n_envs=16, steps=60),        # linear friction, f_range, v_range
    "ball_passage": dict(scenario="ball_passage", kw=dict(), n_envs=16, steps=60),
    "asym_joint": dict(scenario="asym_joint", kw=dict(), n_envs=16, steps=60),
    "wheel": dict(scenario="wheel", kw=dict(), n_envs=16, steps=60),
    "passage": dict(scenario="passage", kw=dict(), n_envs=16, steps=60),
    "reverse_transport": dict(scenario="reverse_transport", kw=dict(), n_envs=16, steps=60),
    "flocking": dict(scenario="flocking", kw=dict(), n_envs=16, steps=40),        # LIDAR with static sphere obstacles
    "discovery": dict(scenario="discovery", kw=dict(), n_envs=16, steps=40),
    "dropout": dict(scenario="dropout", kw=dict(), n_envs=16, steps=40),
    "simple_reference": dict(scenario="simple_reference", kw=dict(), n_envs=16, steps=40),  # dim_c = 10: communication actions / state
}


def state_of(world):
    out = []
    for e in world.entities:
        s = e.state
        out.append(torch.cat([s.pos, s.vel, s.rot, s.ang_vel], dim=-1).detach().clone())
    return torch.stack(out, 0).numpy()  # [E, N, 6]


def agent_ft(world):
    f = torch.stack([a.state.force.detach().clone() for a in world.agents], 0).numpy()  # [A,N,2]
    t = torch.stack([a.state.torque.detach().clone() for a in world.agents], 0).numpy()[..., 0]  # [A,N]
    return f, t


def joint_rots(world, desc, n):
    ents = list(world.entities)
    out = []
    wj = world._joints
    for a in range(len(ents)):
        for b in range(a + 1, len(ents)):
            j = wj.get(frozenset({ents[a].name, ents[b].name}), None)
            if j is None:
                continue
            fr = j.fixed_rotation
            if hasattr(fr, "numel"):
                out.append(fr.detach().reshape(-1).expand(n).clone().numpy() if fr.numel() == 1 else fr.detach().reshape(n).clone().numpy())
            else:
                out.append(np.full((n,), float(fr), dtype=np.float32))
    return np.stack(out, 0).astype(np.float32) if out else np.zeros((0, n), np.float32)


def run(name, cfg):
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    n, steps = cfg["n_envs"], cfg["steps"]
    env = vmas.make_env(
        cfg["scenario"], num_envs=n, device="cpu", seed=0, continuous_actions=True, **cfg["kw"]
    )
    world = env.world
    if "dt" in cfg:
        world._dt = cfg["dt"]
        world._sub_dt = world._dt / world._substeps
    if "drag" in cfg:
        world._drag = cfg["drag"]
    if "substeps" in cfg:  # BASELINE config 3: "transport ... substeps>1" (note i)
        world._substeps = cfg["substeps"]
        world._sub_dt = world._dt / world._substeps
    env.reset(seed=0)
    desc = WorldDesc.from_world(world)
    E, S = len(desc.entities), desc.substeps
    ents = list(world.entities)
    eidx = {id(e): i for i, e in enumerate(ents)}

    rec = dict(in_step=False, substep=-1, masks=[], forces=[], torques=[], overrides=[], jrot_over=[])
    orig_collides = world.collides
    orig_env_force = world._apply_vectorized_enviornment_force
    orig_step = world.step

    def collides(a, b):
        r = orig_collides(a, b)
        if rec["in_step"]:
            rec["cur_mask"][rec["substep"], eidx[id(a)], eidx[id(b)]] = bool(r)
        return r

    def env_force():
        rec["substep"] += 1
        return orig_env_force()

    last_exit = {"state": None, "jrot": None}

    def step():
        pre = state_of(world)
        if last_exit["state"] is not None:
            changed = np.nonzero((pre != last_exit["state"]).any(axis=(1, 2)))[0]
            for e in changed:
                rec["overrides"].append((len(rec["forces"]), int(e), pre[e].copy()))
            jr = joint_rots(world, desc, n)
            if jr.shape[0] and (jr != last_exit["jrot"]).any():
                rec["jrot_over"].append((len(rec["forces"]), jr.copy()))
        f, t = agent_ft(world)
        rec["forces"].append(f)
        rec["torques"].append(t)
        rec["cur_mask"] = np.zeros((S, E, E), dtype=bool)
        rec["in_step"], rec["substep"] = True, -1
        orig_step()
        rec["in_step"] = False
        rec["masks"].append(rec["cur_mask"])
        last_exit["state"] = state_of(world)
        last_exit["jrot"] = joint_rots(world, desc, n)

    world.collides = collides
    world._apply_vectorized_enviornment_force = env_force
    world.step = step

    state0 = state_of(world)
    jrot0 = joint_rots(world, desc, n)
    snaps = [s for s in SNAP if s <= steps]
    if steps not in snaps:
        snaps.append(steps)
    out_state, out_f, out_t, out_lidar, out_done = [], [], [], [], []
    sensors = []
    for ld in desc.lidars:
        sensors.append(ents[ld.entity].sensors[ld.agent_sensor_index])
    for t in range(1, steps + 1):
        actions = [env.get_random_action(a) for a in env.agents]
        env.step(actions)
        if t in snaps:
            out_state.append(last_exit["state"])
            f, tq = agent_ft(world)
            out_f.append(f)
            out_t.append(tq)
            out_done.append(env.scenario.done().detach().clone().numpy())
            if sensors:
                out_lidar.append(
                    [s._last_measurement.detach().clone().numpy().astype(np.float32) for s in sensors]
                )
    masks = np.stack(rec["masks"], 0)  # [T,S,E,E]
    data = dict(
        desc_json=np.array(desc.to_json()),
        meta_json=np.array(
            json.dumps(
                dict(name=name, scenario=cfg["scenario"], kwargs=cfg["kw"], n_envs=n, steps=steps, snaps=snaps,
                     torch=torch.__version__, substeps=S, names=[e.name for e in ents])
            )
        ),
        state0=state0.astype(np.float32),
        jrot0=jrot0,
        forces=np.stack(rec["forces"], 0).astype(np.float32),    # [T,A,N,2] as they enter World.step
        torques=np.stack(rec["torques"], 0).astype(np.float32),  # [T,A,N]
        snaps=np.array(snaps, dtype=np.int32),
        state_after=np.stack(out_state, 0).astype(np.float32),   # [K,E,N,6]
        force_after=np.stack(out_f, 0).astype(np.float32),       # [K,A,N,2] clamped write-back
        torque_after=np.stack(out_t, 0).astype(np.float32),
        done=np.stack(out_done, 0),
        masks_packed=np.packbits(masks.reshape(-1)),
        masks_shape=np.array(masks.shape, dtype=np.int32),
        override_steps=np.array([o[0] for o in rec["overrides"]], dtype=np.int32),
        override_entities=np.array([o[1] for o in rec["overrides"]], dtype=np.int32),
        override_states=(np.stack([o[2] for o in rec["overrides"]], 0).astype(np.float32)
                         if rec["overrides"] else np.zeros((0, n, 6), np.float32)),
        jrot_override_steps=np.array([o[0] for o in rec["jrot_over"]], dtype=np.int32),
        jrot_override_values=(np.stack([o[1] for o in rec["jrot_over"]], 0).astype(np.float32)
                              if rec["jrot_over"] else np.zeros((0,) + jrot0.shape, np.float32)),
    )
    for k, ld in enumerate(desc.lidars):
        data[f"lidar_{k}"] = np.stack([row[k] for row in out_lidar], 0)  # [K,N,R]
    path = os.path.join(HERE, f"{name}.npz")
    np.savez_compressed(path, **data)
    npair = int(np.asarray(desc.collide_mask).sum())
    print(f"{name}: E={E} S={S} pairs={npair} joints={len(desc.joints)} lidars={len(desc.lidars)} "
          f"overrides={len(rec['overrides'])} active_mask_mean={masks.mean():.4f} -> {os.path.getsize(path)/1024:.0f} KiB")


if __name__ == "__main__":
    names = sys.argv[1:] or list(CONFIGS)
    for nm in names:
        run(nm, CONFIGS[nm])
