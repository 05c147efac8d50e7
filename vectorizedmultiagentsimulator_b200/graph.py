"""One CUDA graph per ``Environment.step``.

With the drop-in (``dropin.enable(env, static_io=True, hooks=True)``) a whole ``Environment.step`` of the reference
(environment.py:324-404: ``_set_action`` per agent, ``env_process_action``, ``World.step``, the scenario's reward /
observation / info / done) is ~10 launches and ~60 us of GPU work at 32 768 envs, but ~0.7 ms of Python.  Every one of
those launches goes to torch's current stream and none synchronises, so the whole step is capturable:
``GraphedEnvironment`` runs the reference's own ``Environment.step`` ONCE under ``torch.cuda.graph`` and replays it.

    genv = GraphedEnvironment(env)            # warm-up + capture; the env is left exactly where it was
    obs, rews, dones, infos = genv.step(actions)

The warm-up steps the capture needs (lazy allocations, NVRTC builds) really execute ``env.step``; everything they touch -
world state, ``env.steps``, the scenario's carried terms, agent forces, the environment's random state, the CUDA
generator - is snapshotted before and restored after, for the constructor and for a re-capture alike.

Semantics (the usual ones of a CUDA graph): the returned tensors are the SAME tensors every step, overwritten by the
next replay - clone what must outlive a step; actions are copied into static input tensors.  State lives in the
binding's persistent SoA buffers (``static_io``), the scenario's carried terms (``pos_shaping`` ...) are updated in
place.  ``env.reset()`` / ``reset_at()`` keep working: before a replay every tensor the captured launches READ is
compared with what the scenario currently holds; one a reset rebound is copied into the captured tensor and bound back
(the same discipline as ``WorldBinding.sync_in``).  A reset that changes the world's static description (collision
filters, joints) re-captures.  The reference's action assertions (NaN, out of range) are checked after the replay from
the flag word the launches wrote (``check_actions=False`` skips that one synchronisation).

Needs fused hooks for the scenario (hooks.py): the reference's torch hooks synchronise (``World.collides``,
boolean-mask indexing) and cannot be captured."""
from __future__ import annotations

from . import _lib, dropin


class GraphedEnvironment:
    def __init__(self, env, warmup: int = 3, check_actions: bool = True):
        import torch

        self.torch, self.env = torch, env
        self.b = b = dropin.enable(env, static_io=True, hooks=True)
        if not b.static_io:
            raise _lib.VmasB200Error("GraphedEnvironment: the world was enabled without static_io=True before")
        if b.hooks is None:
            raise _lib.VmasB200Error(
                "GraphedEnvironment: no fused hooks for this scenario (hooks.py REGISTRY); the reference's "
                "torch hooks synchronise with the host and cannot be captured in a CUDA graph")
        if b._env is not env:
            raise _lib.VmasB200Error("GraphedEnvironment needs the Environment the drop-in was enabled on")
        self.check = bool(check_actions)
        self.warmup = max(int(warmup), 1)
        self.actions = [torch.zeros(b.n, a.action_size, device=b.device) for a in env.agents]
        self.n_replays = 0
        self.n_captures = 0
        self.n_input_fixes = 0
        self._capture()

    # ------------------------------------------------------------------ capture
    def _capture(self):
        torch, env, b = self.torch, self.env, self.b
        if b.dirty:
            b.refresh()
        snap = self._snapshot()
        cur = torch.cuda.current_stream(b.device)
        side = torch.cuda.Stream(device=b.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):  # warm-up off the default stream (lazy allocations of the library happen here)
            for _ in range(self.warmup):
                env.step(self.actions)
        cur.wait_stream(side)
        torch.cuda.synchronize(b.device)
        b.check_actions()
        b.defer_action_check = True  # no .item() inside the capture; step() below reads the flag word after the replay
        self.graph = torch.cuda.CUDAGraph()
        # Python's cyclic collector may run at any allocation; if it finalises a StepPlan of an earlier world while the
        # stream is capturing, vmasb200_plan_destroy's cudaFree invalidates the capture (torch >= 2.9 no longer collects
        # on entering torch.cuda.graph): collect now, and keep the collector off until the capture has ended
        import gc

        gc.collect()
        gc_was_on = gc.isenabled()
        gc.disable()
        try:
            with torch.cuda.graph(self.graph):
                self.out = env.step(self.actions)
        except RuntimeError as e:
            # what stays the reference's own torch code (process_action, action scripts, info) must be capturable too:
            # football's ball_action_script builds device tensors from Python floats every step (football.py:1617-1677),
            # a host-to-device copy no CUDA graph can hold
            last = _lib.lib().vmasb200_last_error()
            raise _lib.VmasB200Error("GraphedEnvironment: this scenario's own code cannot be captured in a CUDA graph "
                                     f"(use dropin.enable(env, hooks=True) without the graph): {e}"
                                     f" [last libvmasb200 message: {last.decode() if last else ''}]") from e
        finally:
            if gc_was_on:
                gc.enable()
            b.defer_action_check = False
        b._act_pending = False
        self._plan = b.plan
        self._inputs = [(o, k, getattr(o, k)) for o, k in self._input_attrs()]
        self._restore(snap)
        self.n_captures += 1

    def _snapshot(self):
        """Everything the warm-up steps change: the SoA state buffer (entity state tensors are views of it under
        static_io), env.steps, every tensor the captured launches read (carried shaping terms, goal poses, agent force /
        torque), the Environment's random state (environment.py:30-46 local_seed swaps it in) and the CUDA generator."""
        torch, env, b = self.torch, self.env, self.b
        b.sync_in()  # the buffer holds the current state, the state tensors are its views
        vrs = type(env).vmas_random_state
        return dict(dyn=b.dyn.clone(), attrs=[(o, k, getattr(o, k), getattr(o, k).clone()) for o, k in self._input_attrs()],
                    vrs=[vrs[0].clone(), vrs[1], vrs[2]], cuda_rng=torch.cuda.get_rng_state(b.device),
                    act_flags=None if b._act_flags is None else b._act_flags.clone())

    def _restore(self, snap):
        torch, env, b = self.torch, self.env, self.b
        torch.cuda.synchronize(b.device)
        b.dyn.copy_(snap["dyn"])
        b.lidar_valid = False
        for o, k, t, saved in snap["attrs"]:
            cur = getattr(o, k)  # the tensor the capture reads (a warm-up step may have rebound the attribute to it)
            cur.copy_(saved)
        vrs = type(env).vmas_random_state
        vrs[0], vrs[1], vrs[2] = snap["vrs"]
        torch.cuda.set_rng_state(snap["cuda_rng"], b.device)
        if snap["act_flags"] is not None and b._act_flags is not None:
            b._act_flags.copy_(snap["act_flags"])

    def _input_attrs(self):
        """(object, attribute) of every tensor the captured launches read that a reset may REBIND (in-place writes need
        nothing).  Entity state staged in the binding's buffers is handled by WorldBinding.sync_in."""
        out = [(self.env, "steps")] + list(self.b.hooks.graph_inputs())
        for a in self.b.agents:  # World.step reads these through the pointers of the capture; World.reset rebinds them
            out += [(a.state, "force"), (a.state, "torque")]
        return [(o, k) for o, k in out if self.torch.is_tensor(getattr(o, k, None))]

    def _fix_inputs(self):
        for o, k, t in self._inputs:
            cur = getattr(o, k)
            if cur is not t:
                t.copy_(cur)
                setattr(o, k, t)
                self.n_input_fixes += 1

    # ------------------------------------------------------------------ step
    def step(self, actions):
        """``Environment.step(actions)`` as one graph launch; returns the capture's (static) result tensors."""
        b = self.b
        if b.dirty:  # a reset happened: collision filters / joints may have changed (dropin.WorldBinding.refresh)
            b.refresh()
        # (the reset itself usually refreshes already: its observations call Lidar.measure / World.cast_rays)
        if b.plan is not self._plan:
            self._capture()
        b.sync_in()
        self._fix_inputs()
        if isinstance(actions, dict):
            actions = [actions[a.name] for a in self.env.agents]
        for s, a in zip(self.actions, actions):
            s.copy_(a)
        self.graph.replay()
        self.n_replays += 1
        if self.check:
            b._act_pending = True
            b.check_actions()
        return self.out

    def reset(self, *args, **kwargs):
        return self.env.reset(*args, **kwargs)

    def reset_at(self, *args, **kwargs):
        return self.env.reset_at(*args, **kwargs)
