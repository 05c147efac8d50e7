#!/usr/bin/env python
"""Golden vectors of the scenario hooks from the UNMODIFIED reference: what navigation.py:199-280 and
transport.py:130-193 return and leave behind (`reward`, `observation`, `done`, the attributes they set), together with
the inputs they read (entity state, goal positions, shaping terms of the previous step, LIDAR measurements).

TEST INFRASTRUCTURE.  Run in the build container (torch CPU):

    python tests/golden/make_golden_hooks.py

The reference Environment is stepped with random actions; in a few envs agents are placed on / next to their goals
and next to each other (and packages on their goal) before a step, so that on_goal, the final reward, the collision
penalties and the thin case `distance <= min_collision_distance while no env's centres are within r_a + r_b` occur.
Output: tests/golden/hooks/<scenario>.npz.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub"), ROOT, HERE]

import vmas  # noqa: E402

N, STEPS, KEEP = 96, 40, (0, 1, 5, 6, 7, 12, 13, 20, 21, 30, 39)


def t2n(t):
    return t.detach().clone().numpy()


def navigation(name="navigation", kw=None, tag=None):
    kw = dict(n_agents=4) if kw is None else kw
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    env = vmas.make_env("navigation", num_envs=N, device="cpu", seed=0, continuous_actions=True, **kw)
    sc, world = env.scenario, env.world
    agents = world.agents
    A = len(agents)
    env.reset(seed=0)
    g = torch.Generator().manual_seed(5)
    flags_log = []
    orig_collides = world.collides
    in_reward = {"on": False}

    def collides(a, b):
        r = orig_collides(a, b)
        if in_reward["on"]:
            flags_log[-1][agents.index(a), agents.index(b)] = bool(r)
        return r

    world.collides = collides
    rec = {k: [] for k in ("pos", "vel", "goal", "shaping_in", "lidar", "obs", "reward", "done", "pos_rew", "final_rew",
                           "coll_rew", "dist", "on_goal", "pos_rew_agent", "shaping_out", "all_goal", "flags")}
    orig_reward = sc.reward

    def reward(agent):
        if agent is agents[0]:
            rec["_pre"] = dict(pos=np.stack([t2n(a.state.pos) for a in agents]), vel=np.stack([t2n(a.state.vel) for a in agents]),
                               goal=np.stack([t2n(a.goal.state.pos) for a in agents]),
                               shaping_in=np.stack([t2n(a.pos_shaping) for a in agents]))
            flags_log.append(np.zeros((A, A), bool))
            in_reward["on"] = True
        r = orig_reward(agent)
        in_reward["on"] = False
        return r

    sc.reward = reward
    for t in range(STEPS):
        acts = [(torch.rand(N, a.action_size, generator=g) * 2 - 1) * a.u_range for a in agents]
        if t in (5, 12, 20):  # stage the rare events in a few envs BEFORE the step (the step then moves them a little)
            with torch.no_grad():
                for k in range(0, 8):
                    for a in agents[: 1 + k % A]:
                        a.state.pos[k] = a.goal.state.pos[k] + 0.01 * (torch.rand(2, generator=g) - 0.5)
                        a.state.vel[k] = 0.0
                for k in range(8, 12):  # all agents on their goals
                    for a in agents:
                        a.state.pos[k] = a.goal.state.pos[k]
                        a.state.vel[k] = 0.0
                for k in range(12, 24):  # agent pairs at contact distance: 2 r + [-0.01, 0.012]
                    i, j = k % A, (k + 1 + k // A) % A
                    if i != j:
                        d = 2 * agents[i].shape.radius + float(torch.rand(1, generator=g)) * 0.022 - 0.01
                        ang = float(torch.rand(1, generator=g)) * 6.28
                        agents[j].state.pos[k] = agents[i].state.pos[k] + d * torch.tensor([np.cos(ang), np.sin(ang)], dtype=torch.float32)
                        agents[i].state.vel[k] = 0.0
                        agents[j].state.vel[k] = 0.0
            acts = [a * 0 for a in acts]
        obs, rews, dones, infos = env.step(acts)
        if t in KEEP:
            pre = rec.pop("_pre")
            for k in ("pos", "vel", "goal", "shaping_in"):
                rec[k].append(pre[k])
            rec["lidar"].append(np.stack([t2n(a.sensors[0]._last_measurement) for a in agents]) if sc.collisions else np.zeros((A, N, 0), np.float32))
            rec["obs"].append(np.stack([t2n(o) for o in obs]))
            rec["reward"].append(np.stack([t2n(r) for r in rews]))
            rec["done"].append(t2n(dones))
            rec["pos_rew"].append(t2n(sc.pos_rew))
            rec["final_rew"].append(t2n(sc.final_rew))
            rec["coll_rew"].append(np.stack([t2n(a.agent_collision_rew) for a in agents]))
            rec["dist"].append(np.stack([t2n(a.distance_to_goal) for a in agents]))
            rec["on_goal"].append(np.stack([t2n(a.on_goal) for a in agents]))
            rec["pos_rew_agent"].append(np.stack([t2n(a.pos_rew) for a in agents]))
            rec["shaping_out"].append(np.stack([t2n(a.pos_shaping) for a in agents]))
            rec["all_goal"].append(t2n(sc.all_goal_reached))
            rec["flags"].append(flags_log[-1].copy())
        else:
            rec.pop("_pre", None)
    out = {k: np.stack(v) for k, v in rec.items()}
    out["radius"] = np.array([a.shape.radius for a in agents], np.float64)
    out["goal_radius"] = np.array([a.goal.shape.radius for a in agents], np.float64)
    out["params"] = np.array([sc.pos_shaping_factor, sc.final_reward, sc.agent_collision_penalty, sc.min_collision_distance,
                              float(sc.shared_rew), float(agents[0].sensors[0]._max_range) if sc.collisions else 0.0,
                              float(sc.observe_all_goals)], np.float64)
    os.makedirs(os.path.join(HERE, "hooks"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "hooks", f"{tag or name}.npz"), **out)
    print(tag or name, {k: v.shape for k, v in out.items()}, "events: on_goal", int(out["on_goal"].sum()), "all_goal",
          int(out["all_goal"].sum()), "collision penalties", int((out["coll_rew"] != 0).sum()), "flags", int(out["flags"].sum()),
          "done", int(out["done"].sum()))


def transport():
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    env = vmas.make_env("transport", num_envs=N, device="cpu", seed=0, continuous_actions=True, n_agents=4)
    sc, world = env.scenario, env.world
    agents, pk = world.agents, sc.packages
    env.reset(seed=0)
    g = torch.Generator().manual_seed(5)
    rec = {k: [] for k in ("agent_pos", "agent_vel", "pkg_state", "goal_pos", "shaping_in", "obs", "rew", "done", "on_goal",
                           "dist", "shaping_out", "color")}
    orig_reward = sc.reward

    def reward(agent):
        if agent is agents[0]:
            rec["_pre"] = dict(
                agent_pos=np.stack([t2n(a.state.pos) for a in agents]), agent_vel=np.stack([t2n(a.state.vel) for a in agents]),
                pkg_state=np.stack([t2n(torch.cat([p.state.pos, p.state.vel, p.state.rot, p.state.ang_vel], -1)) for p in pk]),
                goal_pos=np.stack([t2n(p.goal.state.pos) for p in pk]), shaping_in=np.stack([t2n(p.global_shaping) for p in pk]))
        return orig_reward(agent)

    sc.reward = reward
    for t in range(STEPS):
        acts = [(torch.rand(N, a.action_size, generator=g) * 2 - 1) * a.u_range for a in agents]
        if t in (5, 12, 20):
            with torch.no_grad():
                for p in pk:
                    L, W, r = p.shape.length, p.shape.width, p.goal.shape.radius
                    for k in range(0, 24):  # package centre from inside the goal to just outside the overlap distance
                        d = (k / 23.0) * (max(L, W) / 2 + r + 0.02)
                        ang = float(torch.rand(1, generator=g)) * 6.28
                        p.state.pos[k] = p.goal.state.pos[k] + d * torch.tensor([np.cos(ang), np.sin(ang)], dtype=torch.float32)
                        p.state.vel[k] = 0.0
                for a in agents:  # keep the agents away so that the step leaves the staged package alone
                    for k in range(0, 24):
                        a.state.pos[k] = torch.tensor([5.0, 5.0]) + torch.rand(2, generator=g)
        obs, rews, dones, infos = env.step(acts)
        if t in KEEP:
            pre = rec.pop("_pre")
            for k in ("agent_pos", "agent_vel", "pkg_state", "goal_pos", "shaping_in"):
                rec[k].append(pre[k])
            rec["obs"].append(np.stack([t2n(o) for o in obs]))
            rec["rew"].append(t2n(rews[0]))
            rec["done"].append(t2n(dones))
            rec["on_goal"].append(np.stack([t2n(p.on_goal) for p in pk]))
            rec["dist"].append(np.stack([t2n(p.dist_to_goal) for p in pk]))
            rec["shaping_out"].append(np.stack([t2n(p.global_shaping) for p in pk]))
            rec["color"].append(np.stack([t2n(p.color) for p in pk]))
        else:
            rec.pop("_pre", None)
    out = {k: np.stack(v) for k, v in rec.items()}
    out["pkg_dims"] = np.array([[0.0, p.shape.length, p.shape.width] for p in pk], np.float64)
    out["goal_radius"] = np.array([p.goal.shape.radius for p in pk], np.float64)
    out["params"] = np.array([sc.shaping_factor], np.float64)
    os.makedirs(os.path.join(HERE, "hooks"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "hooks", "transport.npz"), **out)
    print("transport", {k: v.shape for k, v in out.items()}, "on_goal", int(out["on_goal"].sum()), "done", int(out["done"].sum()))


def simple_spread():
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    env = vmas.make_env("simple_spread", num_envs=N, device="cpu", seed=0, continuous_actions=True, n_agents=3)
    sc, world = env.scenario, env.world
    agents, lms = world.agents, world.landmarks
    env.reset(seed=0)
    g = torch.Generator().manual_seed(5)
    rec = {k: [] for k in ("agent_pos", "agent_vel", "landmark_pos", "obs", "rew", "done")}
    for t in range(STEPS):
        acts = [(torch.rand(N, a.action_size, generator=g) * 2 - 1) * a.u_range for a in agents]
        if t in (5, 12, 20):  # overlapping agents in a few envs
            with torch.no_grad():
                for k in range(0, 24):
                    i, j = k % 3, (k + 1) % 3
                    d = 2 * agents[i].shape.radius * (0.8 + 0.4 * float(torch.rand(1, generator=g)))
                    agents[j].state.pos[k] = agents[i].state.pos[k] + torch.tensor([d, 0.0])
            acts = [a * 0 for a in acts]
        obs, rews, dones, infos = env.step(acts)
        if t in KEEP:
            rec["agent_pos"].append(np.stack([t2n(a.state.pos) for a in agents]))
            rec["agent_vel"].append(np.stack([t2n(a.state.vel) for a in agents]))
            rec["landmark_pos"].append(np.stack([t2n(l.state.pos) for l in lms]))
            rec["obs"].append(np.stack([t2n(o) for o in obs]))
            rec["rew"].append(t2n(rews[0]))
            rec["done"].append(t2n(dones))
    out = {k: np.stack(v) for k, v in rec.items()}
    out["radius"] = np.array([a.shape.radius for a in agents], np.float64)
    out["collide"] = np.array([bool(a.collide) for a in agents])
    out["params"] = np.array([float(sc.obs_agents)], np.float64)
    np.savez_compressed(os.path.join(HERE, "hooks", "simple_spread.npz"), **out)
    print("simple_spread", {k: v.shape for k, v in out.items()}, "rew range", float(out["rew"].min()), float(out["rew"].max()))


def _ent_state(e):
    return t2n(torch.cat([e.state.pos, e.state.vel, e.state.rot, e.state.ang_vel], -1))


def joint_passage(kw, tag):
    """joint_passage.py:387-563.  Samples are taken by calling the hooks on the world as it stands: after a step, and on
    states staged by hand (joint / agents beyond the passage, joint on its goal in EVERY env - done() couples envs)."""
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    env = vmas.make_env("joint_passage", num_envs=N, device="cpu", seed=0, continuous_actions=True, **kw)
    sc, world = env.scenario, env.world
    agents = world.agents
    bodies = list(agents) + ([sc.mass] if sc.asym_package else [])
    env.reset(seed=0)
    g = torch.Generator().manual_seed(5)
    names = ("pos_shaping_pre", "pos_shaping_post", "rot_shaping_pre", "rot_shaping_post")
    rec = {k: [] for k in ("bodies", "action_u", "joint", "goal", "passages", "walls", "shaping_in", "shaping_out", "obs", "rew",
                           "pos_rew", "rot_rew", "collision_rew", "energy_expenditure", "energy_rew", "all_passed", "done")}

    def sample():
        rec["bodies"].append(np.stack([_ent_state(e) for e in bodies]))
        rec["action_u"].append(np.stack([t2n(a.action.u) for a in agents]))
        rec["joint"].append(_ent_state(sc.joint.landmark))
        rec["goal"].append(_ent_state(sc.goal))
        rec["passages"].append(np.stack([_ent_state(p) for p in sc.passages]))
        rec["walls"].append(np.stack([_ent_state(w) for w in sc.walls]))
        rec["shaping_in"].append(np.stack([t2n(getattr(sc.joint, k)) for k in names]))
        rews = [sc.reward(a) for a in agents]
        rec["obs"].append(np.stack([t2n(sc.observation(a)) for a in agents]))
        rec["rew"].append(t2n(rews[0]))
        for k in ("pos_rew", "rot_rew", "collision_rew", "energy_expenditure", "energy_rew", "all_passed"):
            rec[k].append(t2n(getattr(sc, k)))
        rec["shaping_out"].append(np.stack([t2n(getattr(sc.joint, k)) for k in names]))
        rec["done"].append(t2n(sc.done()))

    for t in range(STEPS):
        acts = [(torch.rand(N, a.action_size, generator=g) * 2 - 1) * a.u_range for a in agents]
        env.step(acts)
        if t in KEEP:
            sample()
        if t in (6, 13, 21):
            with torch.no_grad():
                jl = sc.joint.landmark
                for k in range(0, 32):  # joint (and, in half of them, every agent) beyond the passage
                    jl.state.pos[k, 1] = 0.05 + 0.01 * k
                    if k % 2:
                        for a in agents:
                            a.state.pos[k, 1] = 0.11 + 0.01 * (k % 5)
                for k in range(32, 48):  # bodies next to walls / passages
                    for i, e in enumerate(bodies):
                        e.state.pos[k] = torch.tensor([0.95 + 0.004 * (k - 32), -0.5 + 0.1 * i]) if k % 2 else \
                            sc.passages[(k + i) % len(sc.passages)].state.pos[k] + torch.tensor([0.0, -0.13 - 0.002 * (k - 32)])
                for k in range(48, 56):  # rotations on both sides of the wrap
                    jl.state.rot[k] = float(torch.rand(1, generator=g)) * 12.0 - 6.0
            sample()
        if t == 30:  # joint on its goal everywhere (position), with the rotation matching in some envs: done() can be true
            with torch.no_grad():
                jl = sc.joint.landmark
                jl.state.pos[:] = sc.goal.state.pos + 0.004
                jl.state.rot[:] = sc.goal.state.rot + torch.linspace(-0.03, 0.03, N).unsqueeze(-1)
            sample()
    out = {k: np.stack(v) for k, v in rec.items()}
    out["body_radius"] = np.array([e.shape.radius for e in bodies], np.float64)
    out["passage_dims"] = np.array([[p.shape.length, p.shape.width] for p in sc.passages], np.float64)
    out["passage_collide"] = np.array([bool(p.collide) for p in sc.passages])
    out["wall_length"] = np.array([w.shape.length for w in sc.walls], np.float64)
    out["joint_length"] = np.array([sc.joint.landmark.shape.length], np.float64)
    out["f_range"] = np.array([a.f_range for a in agents], np.float64)
    out["params"] = np.array([sc.passage_width, sc.pos_shaping_factor, sc.rot_shaping_factor, sc.collision_reward,
                              sc.energy_reward_coeff, sc.min_collision_distance, sc.middle_angle, float(sc.all_passed_rot),
                              float(world.dim_p), float(sc.observe_joint_angle), float(len(agents))], np.float64)
    np.savez_compressed(os.path.join(HERE, "hooks", f"{tag}.npz"), **out)
    print(tag, {k: v.shape for k, v in out.items() if k in ("bodies", "obs", "rew")}, "all_passed", int(out["all_passed"].sum()),
          "done", int(out["done"].sum()), "collision_rew != 0", int((out["collision_rew"] != 0).sum()))


def football(kw, tag):
    """football.py:1118-1467 with teams of policy agents."""
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    env = vmas.make_env("football", num_envs=N, device="cpu", seed=0, continuous_actions=True, **kw)
    sc, world = env.scenario, env.world
    ball = sc.ball
    pol = env.agents
    env.reset(seed=0)
    g = torch.Generator().manual_seed(5)
    teams = (("blue", sc.blue_agents), ("red", sc.red_agents))
    rec = {k: [] for k in ("ball", "ball_force", "blue", "red", "blue_force", "red_force", "shaping_in", "obs", "rew_blue", "rew_red",
                           "done", "shaping_out", "dist_goal", "pos_rew", "min_dist", "pos_rew_agent", "sparse")}

    def sample():
        rec["ball"].append(_ent_state(ball))
        rec["ball_force"].append(t2n(ball.state.force))
        for name, team in teams:
            rec[name].append(np.stack([_ent_state(a) for a in team]))
            rec[name + "_force"].append(np.stack([t2n(a.state.force) for a in team]))
        zero = torch.zeros(N)
        rec["shaping_in"].append(np.stack([t2n(getattr(ball, k, zero)) for k in ("pos_shaping_blue", "pos_shaping_agent_blue",
                                                                                "pos_shaping_red", "pos_shaping_agent_red")]))
        rews = [sc.reward(a) for a in pol]
        rec["obs"].append(np.stack([t2n(sc.observation(a)) for a in pol]))
        rec["rew_blue"].append(t2n(rews[0]))
        rec["rew_red"].append(t2n(rews[-1]))
        rec["done"].append(t2n(sc.done()))
        rec["shaping_out"].append(np.stack([t2n(getattr(ball, k, zero)) for k in ("pos_shaping_blue", "pos_shaping_agent_blue",
                                                                                 "pos_shaping_red", "pos_shaping_agent_red")]))
        if sc.dense_reward:
            rec["dist_goal"].append(np.stack([t2n(ball.distance_to_goal_blue), t2n(ball.distance_to_goal_red)]))
            rec["pos_rew"].append(np.stack([t2n(ball.pos_rew_blue), t2n(ball.pos_rew_red)]))
            rec["min_dist"].append(np.stack([t2n(sc.min_agent_dist_to_ball_blue), t2n(sc.min_agent_dist_to_ball_red)]))
            rec["pos_rew_agent"].append(np.stack([t2n(ball.pos_rew_agent_blue), t2n(ball.pos_rew_agent_red)]))
        else:
            for k in ("dist_goal", "pos_rew", "min_dist", "pos_rew_agent"):
                rec[k].append(np.zeros((2, N), np.float32))
        rec["sparse"].append(np.stack([t2n(sc._sparse_reward_blue), t2n(sc._sparse_reward_red)]))

    for t in range(STEPS):
        acts = [(torch.rand(N, a.action_size, generator=g) * 2 - 1) * a.u_range for a in pol]
        env.step(acts)
        if t in KEEP:
            sample()
        if t in (6, 13, 21):
            with torch.no_grad():
                L, G = sc.pitch_length / 2 + sc.ball_size / 2, sc.goal_size / 2
                for k in range(0, 24):  # ball around the goal lines and the posts
                    side = 1.0 if k % 2 else -1.0
                    ball.state.pos[k] = torch.tensor([side * (L + 0.002 * (k % 5 - 2)), (G + 0.003 * (k % 7 - 3)) * (1 if k % 3 else -1)])
                for k in range(24, 48):  # ball at rest, agents around the trigger distance
                    ball.state.vel[k] = 0.0
                    for i, a in enumerate(sc.blue_agents + sc.red_agents):
                        d = sc.distance_to_ball_trigger + 0.02 * ((k + i) % 7 - 3)
                        ang = float(torch.rand(1, generator=g)) * 6.28
                        a.state.pos[k] = ball.state.pos[k] + d * torch.tensor([np.cos(ang), np.sin(ang)], dtype=torch.float32)
            sample()
    out = {k: np.stack(v) for k, v in rec.items()}
    out["goal_pos"] = np.stack([t2n(sc.right_goal_pos), t2n(sc.left_goal_pos)])
    out["params"] = np.array([sc.pitch_length, sc.ball_size, sc.goal_size, sc.scoring_reward, sc.pos_shaping_factor_ball_goal,
                              sc.pos_shaping_factor_agent_ball, sc.distance_to_ball_trigger, float(sc.dense_reward),
                              float(sc.observe_teammates), float(sc.observe_adversaries)], np.float64)
    np.savez_compressed(os.path.join(HERE, "hooks", f"{tag}.npz"), **out)
    print(tag, {k: v.shape for k, v in out.items() if k in ("ball", "obs", "rew_blue")}, "done", int(out["done"].sum()),
          "agent shaping rewards != 0", int((out["pos_rew_agent"] != 0).sum()))


if __name__ == "__main__":
    simple_spread()
    navigation()
    navigation(kw=dict(n_agents=3, shared_rew=False, observe_all_goals=True), tag="navigation_noshare")
    navigation(kw=dict(n_agents=2, collisions=False), tag="navigation_nocoll")
    transport()
    joint_passage(dict(), "joint_passage")
    joint_passage(dict(collision_reward=-0.1, energy_reward_coeff=0.01, observe_joint_angle=True, all_passed_rot=False), "joint_passage_terms")
    football(dict(ai_red_agents=False), "football")
    football(dict(ai_red_agents=False, n_blue_agents=2, n_red_agents=2, dense_reward=False, observe_teammates=False), "football_sparse")
