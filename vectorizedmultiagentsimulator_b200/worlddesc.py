"""Static description of a VMAS ``World`` as plain data (the ``vmasb200_world_desc`` of include/vmasb200.h).

``WorldDesc.from_world`` walks a live reference ``World`` (duck-typed: nothing from ``vmas`` is imported) and
records exactly what the reference's step reads from Python objects every substep:

* entity table in ``World.entities`` order (landmarks, then agents: core.py:1220) with the physical attributes
  of core.py:537-760 (Entity) and :829-1085 (Agent) and the shape data of core.py:102-202;
* the STATIC part of ``World.collides`` (core.py:2787-2795) and the joint exclusion of core.py:2115-2121,
  as an E x E boolean matrix;
* the ``JointConstraint`` list in the order ``_apply_vectorized_enviornment_force`` meets it (core.py:2111-2121),
  with anchor deltas from ``shape.get_delta_from_anchor`` (joints.py:185-201);
* the ``Lidar`` sensors of every agent (sensors.py:46-69) with their filtered target lists (core.py:1677-1690).

The description is JSON-serialisable so parity fixtures carry it without needing the reference installed.
"""
from __future__ import annotations

import ctypes as C
import json
import math
from dataclasses import asdict, dataclass, field
from typing import List, Optional

import numpy as np

SPHERE, BOX, LINE = 0, 1, 2

E_MOVABLE = 1 << 0
E_ROTATABLE = 1 << 1
E_AGENT = 1 << 2
E_HOLLOW = 1 << 3
E_MAX_SPEED = 1 << 4
E_V_RANGE = 1 << 5
E_DRAG = 1 << 6
E_LIN_FRICTION = 1 << 7
E_ANG_FRICTION = 1 << 8
E_GRAVITY = 1 << 9
E_MAX_F = 1 << 10
E_F_RANGE = 1 << 11
E_MAX_T = 1 << 12
E_T_RANGE = 1 << 13
E_COLLIDE = 1 << 14

ABI_VERSION = 3


class UnsupportedWorld(RuntimeError):
    """The World uses a feature the CUDA step does not implement (no silent fallback)."""


# ----------------------------------------------------------------------------------------------- ctypes mirrors
class CEntity(C.Structure):
    _fields_ = [
        ("shape", C.c_int32),
        ("flags", C.c_uint32),
        ("radius", C.c_float),
        ("length", C.c_float),
        ("width", C.c_float),
        ("mass", C.c_float),
        ("moment_of_inertia", C.c_float),
        ("max_speed", C.c_float),
        ("v_range", C.c_float),
        ("drag", C.c_double),
        ("linear_friction", C.c_float),
        ("angular_friction", C.c_float),
        ("gravity", C.c_float * 2),
        ("max_f", C.c_float),
        ("f_range", C.c_float),
        ("max_t", C.c_float),
        ("t_range", C.c_float),
        ("circumscribed_radius", C.c_double),
    ]


class CJoint(C.Structure):
    _fields_ = [
        ("entity_a", C.c_int32),
        ("entity_b", C.c_int32),
        ("delta_a", C.c_float * 2),
        ("delta_b", C.c_float * 2),
        ("dist", C.c_float),
        ("rotate", C.c_int32),
        ("fixed_rotation", C.c_float),
        ("fixed_rotation_per_env", C.c_int32),
    ]


class CLidar(C.Structure):
    _fields_ = [
        ("entity", C.c_int32),
        ("n_rays", C.c_int32),
        ("angles", C.POINTER(C.c_float)),
        ("max_range", C.c_float),
        ("n_targets", C.c_int32),
        ("targets", C.POINTER(C.c_int32)),
    ]


class CWorldDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("n_entities", C.c_int32),
        ("entities", C.POINTER(CEntity)),
        ("collide_mask", C.POINTER(C.c_uint8)),
        ("n_joints", C.c_int32),
        ("joints", C.POINTER(CJoint)),
        ("n_lidars", C.c_int32),
        ("lidars", C.POINTER(CLidar)),
        ("dt", C.c_double),
        ("substeps", C.c_int32),
        ("drag", C.c_double),
        ("linear_friction", C.c_float),
        ("angular_friction", C.c_float),
        ("has_x_semidim", C.c_int32),
        ("has_y_semidim", C.c_int32),
        ("x_semidim", C.c_float),
        ("y_semidim", C.c_float),
        ("collision_force", C.c_float),
        ("joint_force", C.c_float),
        ("torque_constraint_force", C.c_float),
        ("contact_margin", C.c_float),
        ("gravity", C.c_float * 2),
        ("envs_per_cta", C.c_int32),
    ]


class CPair(C.Structure):
    _fields_ = [("type", C.c_int32), ("a", C.c_int32), ("b", C.c_int32)]


class CLayout(C.Structure):
    _fields_ = [
        ("dyn_rows", C.c_int32),
        ("stat_rows", C.c_int32),
        ("act_rows", C.c_int32),
        ("lidar_rows", C.c_int32),
        ("n_pairs", C.c_int32),
        ("pair_words", C.c_int32),
        ("envs_per_cta", C.c_int32),
        ("n_dyn", C.c_int32),
        ("n_stat", C.c_int32),
        ("n_agents", C.c_int32),
        ("n_act_out", C.c_int32),
        ("alg_bytes_per_env", C.c_int64),
    ]


class CStepIO(C.Structure):
    _fields_ = [
        ("dyn_in", C.c_void_p),
        ("dyn_out", C.c_void_p),
        ("stat", C.c_void_p),
        ("force_ptrs", C.POINTER(C.c_void_p)),
        ("force_strides", C.POINTER(C.c_int64)),
        ("torque_ptrs", C.POINTER(C.c_void_p)),
        ("torque_strides", C.POINTER(C.c_int64)),
        ("act_out", C.c_void_p),
        ("lidar_out", C.c_void_p),
        ("pair_active_out", C.c_void_p),
        ("n_envs", C.c_int64),
        ("ld", C.c_int64),
        ("exact", C.c_int32),
        ("stream", C.c_void_p),
        ("substep_begin", C.c_int32),
        ("substep_end", C.c_int32),
        ("forces_are_substep_inputs", C.c_int32),
        ("trig_in", C.c_void_p),
        ("force_out", C.c_void_p),
        ("joint_out", C.c_void_p),
        ("jexp_in", C.c_void_p),
    ]


# ------------------------------------------------------------------------------------------------ plain data
@dataclass
class EntityDesc:
    name: str
    shape: int
    flags: int
    radius: float = 0.0
    length: float = 0.0
    width: float = 0.0
    mass: float = 1.0
    moment_of_inertia: float = 1.0
    max_speed: float = 0.0
    v_range: float = 0.0
    drag: float = 0.0
    linear_friction: float = 0.0
    angular_friction: float = 0.0
    gravity: List[float] = field(default_factory=lambda: [0.0, 0.0])
    max_f: float = 0.0
    f_range: float = 0.0
    max_t: float = 0.0
    t_range: float = 0.0
    circumscribed_radius: float = 0.0

    @property
    def movable(self):
        return bool(self.flags & E_MOVABLE)

    @property
    def rotatable(self):
        return bool(self.flags & E_ROTATABLE)

    @property
    def is_agent(self):
        return bool(self.flags & E_AGENT)

    @property
    def dynamic(self):
        return bool(self.flags & (E_MOVABLE | E_ROTATABLE))


@dataclass
class JointDesc:
    entity_a: int
    entity_b: int
    delta_a: List[float]
    delta_b: List[float]
    dist: float
    rotate: bool
    fixed_rotation: float = 0.0
    fixed_rotation_per_env: bool = False


@dataclass
class LidarDesc:
    entity: int
    angles: List[float]
    max_range: float
    targets: List[int]
    agent_sensor_index: int = 0  # position inside agent.sensors (host bookkeeping only)


@dataclass
class WorldDesc:
    entities: List[EntityDesc]
    collide_mask: List[List[int]]  # E x E, only a<b meaningful
    joints: List[JointDesc]
    lidars: List[LidarDesc]
    dt: float
    substeps: int
    drag: float
    linear_friction: float
    angular_friction: float
    x_semidim: Optional[float]
    y_semidim: Optional[float]
    collision_force: float
    joint_force: float
    torque_constraint_force: float
    contact_margin: float
    gravity: List[float]
    envs_per_cta: int = 0

    # ------------------------------------------------------------------ (de)serialisation
    def to_json(self) -> str:
        return json.dumps(asdict(self))

    @staticmethod
    def from_json(s: str) -> "WorldDesc":
        d = json.loads(s)
        d.pop("lanes_per_env", None)  # fixtures written for ABI 1
        d["entities"] = [EntityDesc(**e) for e in d["entities"]]
        d["joints"] = [JointDesc(**j) for j in d["joints"]]
        d["lidars"] = [LidarDesc(**l) for l in d["lidars"]]
        return WorldDesc(**d)

    @property
    def agent_indices(self) -> List[int]:
        return [i for i, e in enumerate(self.entities) if e.is_agent]

    # ------------------------------------------------------------------ ctypes view (keeps buffers alive)
    def to_ctypes(self):
        E = len(self.entities)
        ents = (CEntity * max(E, 1))()
        for i, e in enumerate(self.entities):
            c = ents[i]
            c.shape, c.flags = e.shape, e.flags
            c.radius, c.length, c.width = e.radius, e.length, e.width
            c.mass, c.moment_of_inertia = e.mass, e.moment_of_inertia
            c.max_speed, c.v_range, c.drag = e.max_speed, e.v_range, e.drag
            c.linear_friction, c.angular_friction = e.linear_friction, e.angular_friction
            c.gravity[0], c.gravity[1] = e.gravity
            c.max_f, c.f_range, c.max_t, c.t_range = e.max_f, e.f_range, e.max_t, e.t_range
            c.circumscribed_radius = e.circumscribed_radius
        mask = np.ascontiguousarray(np.asarray(self.collide_mask, dtype=np.uint8).reshape(E, E))
        J = len(self.joints)
        joints = (CJoint * max(J, 1))()
        for i, j in enumerate(self.joints):
            c = joints[i]
            c.entity_a, c.entity_b = j.entity_a, j.entity_b
            c.delta_a[0], c.delta_a[1] = j.delta_a
            c.delta_b[0], c.delta_b[1] = j.delta_b
            c.dist, c.rotate = j.dist, int(j.rotate)
            c.fixed_rotation, c.fixed_rotation_per_env = j.fixed_rotation, int(j.fixed_rotation_per_env)
        L = len(self.lidars)
        lidars = (CLidar * max(L, 1))()
        keep = [ents, mask, joints, lidars]
        for i, l in enumerate(self.lidars):
            ang = (C.c_float * max(len(l.angles), 1))(*l.angles)
            tg = (C.c_int32 * max(len(l.targets), 1))(*l.targets)
            keep += [ang, tg]
            c = lidars[i]
            c.entity, c.n_rays, c.max_range = l.entity, len(l.angles), l.max_range
            c.angles = C.cast(ang, C.POINTER(C.c_float))
            c.n_targets = len(l.targets)
            c.targets = C.cast(tg, C.POINTER(C.c_int32))
        d = CWorldDesc()
        d.abi_version = ABI_VERSION
        d.n_entities = E
        d.entities = C.cast(ents, C.POINTER(CEntity))
        d.collide_mask = mask.ctypes.data_as(C.POINTER(C.c_uint8))
        d.n_joints, d.joints = J, C.cast(joints, C.POINTER(CJoint))
        d.n_lidars, d.lidars = L, C.cast(lidars, C.POINTER(CLidar))
        d.dt, d.substeps, d.drag = self.dt, self.substeps, self.drag
        d.linear_friction, d.angular_friction = self.linear_friction, self.angular_friction
        d.has_x_semidim = int(self.x_semidim is not None)
        d.has_y_semidim = int(self.y_semidim is not None)
        d.x_semidim = self.x_semidim or 0.0
        d.y_semidim = self.y_semidim or 0.0
        d.collision_force, d.joint_force = self.collision_force, self.joint_force
        d.torque_constraint_force, d.contact_margin = self.torque_constraint_force, self.contact_margin
        d.gravity[0], d.gravity[1] = self.gravity
        d.envs_per_cta = self.envs_per_cta
        return d, keep

    # ------------------------------------------------------------------ extraction from a live reference World
    @staticmethod
    def from_world(world, with_lidars: bool = True) -> "WorldDesc":
        ents = list(world.entities)
        index = {id(e): i for i, e in enumerate(ents)}
        edescs = [_entity_desc(e) for e in ents]
        E = len(ents)
        mask = [[0] * E for _ in range(E)]
        joints: List[JointDesc] = []
        wjoints = getattr(world, "_joints", {})
        for a in range(E):
            for b in range(a + 1, E):
                ea, eb = ents[a], ents[b]
                joint = wjoints.get(frozenset({ea.name, eb.name}), None)
                if joint is not None:
                    joints.append(_joint_desc(joint, index))
                    if joint.dist == 0:
                        continue
                mask[a][b] = int(_static_collides(ea, eb))
        lidars: List[LidarDesc] = []
        if with_lidars:
            for ag in world.agents:
                for si, sensor in enumerate(getattr(ag, "sensors", []) or []):
                    ld = _lidar_desc(world, ag, sensor, si, ents, index)
                    if ld is not None:
                        lidars.append(ld)
        g = world._gravity.detach().cpu().reshape(-1).tolist()
        return WorldDesc(
            entities=edescs,
            collide_mask=mask,
            joints=joints,
            lidars=lidars,
            dt=float(world._dt),
            substeps=int(world._substeps),
            drag=float(world._drag),
            linear_friction=float(world._linear_friction),
            angular_friction=float(world._angular_friction),
            x_semidim=None if world._x_semidim is None else float(world._x_semidim),
            y_semidim=None if world._y_semidim is None else float(world._y_semidim),
            collision_force=float(world._collision_force),
            joint_force=float(world._joint_force),
            torque_constraint_force=float(world._torque_constraint_force),
            contact_margin=float(world._contact_margin),
            gravity=[float(g[0]), float(g[1])],
        )


def _shape_kind(shape) -> int:
    n = type(shape).__name__
    for cls in type(shape).__mro__:
        if cls.__name__ in ("Sphere", "Box", "Line"):
            n = cls.__name__
            break
    if n == "Sphere":
        return SPHERE
    if n == "Box":
        return BOX
    if n == "Line":
        return LINE
    raise UnsupportedWorld(f"unknown shape class {type(shape).__name__}")


def _entity_desc(e) -> EntityDesc:
    kind = _shape_kind(e.shape)
    flags = 0
    if e.movable:
        flags |= E_MOVABLE
    if e.rotatable:
        flags |= E_ROTATABLE
    if e.collide:
        flags |= E_COLLIDE
    d = EntityDesc(name=e.name, shape=kind, flags=0)
    if kind == SPHERE:
        d.radius = float(e.shape.radius)
    elif kind == BOX:
        d.length, d.width = float(e.shape.length), float(e.shape.width)
        if e.shape.hollow:
            flags |= E_HOLLOW
    else:
        d.length = float(e.shape.length)
    d.mass = float(e.mass)
    d.moment_of_inertia = float(e.moment_of_inertia)
    d.circumscribed_radius = float(e.shape.circumscribed_radius())
    if e.max_speed is not None:
        flags |= E_MAX_SPEED
        d.max_speed = float(e.max_speed)
    if e.v_range is not None:
        flags |= E_V_RANGE
        d.v_range = float(e.v_range)
    if e.drag is not None:
        flags |= E_DRAG
        d.drag = float(e.drag)
    if e.linear_friction is not None:
        flags |= E_LIN_FRICTION
        d.linear_friction = _scalar(e.linear_friction, "linear_friction")
    if e.angular_friction is not None:
        flags |= E_ANG_FRICTION
        d.angular_friction = _scalar(e.angular_friction, "angular_friction")
    if e.gravity is not None:
        g = e.gravity
        vals = g.detach().cpu().reshape(-1).tolist() if hasattr(g, "detach") else [float(g)]
        if len(vals) == 1:
            vals = [vals[0], vals[0]]  # a 0-d tensor broadcasts over (x, y): core.py:2049-2051
        if len(vals) != 2:
            raise UnsupportedWorld(f"entity {e.name}: per-env gravity tensors are not supported by the CUDA step")
        flags |= E_GRAVITY
        d.gravity = [float(vals[0]), float(vals[1])]
    if hasattr(e, "action"):  # Agent (core.py:829)
        flags |= E_AGENT
        if e.max_f is not None:
            flags |= E_MAX_F
            d.max_f = float(e.max_f)
        if e.f_range is not None:
            flags |= E_F_RANGE
            d.f_range = float(e.f_range)
        if e.max_t is not None:
            flags |= E_MAX_T
            d.max_t = float(e.max_t)
        if e.t_range is not None:
            flags |= E_T_RANGE
            d.t_range = float(e.t_range)
    d.flags = flags
    return d


def _scalar(v, what) -> float:
    if hasattr(v, "numel"):
        if v.numel() != 1:
            raise UnsupportedWorld(f"tensor-valued {what} is not supported by the CUDA step")
        return float(v.item())
    return float(v)


def _truthy(v) -> bool:
    # collision filters may return 0-d bool tensors (joint_passage.py:615-620)
    return bool(v.item()) if hasattr(v, "item") else bool(v)


def _static_collides(a, b) -> bool:
    """World.collides minus its batch-wide distance test (core.py:2787-2795)."""
    if not a.collide or not _truthy(a._collision_filter(b)):
        return False
    if not b.collide or not _truthy(b._collision_filter(a)):
        return False
    if a is b:
        return False
    if not a.movable and not a.rotatable and not b.movable and not b.rotatable:
        return False
    _shape_kind(a.shape), _shape_kind(b.shape)  # all six class pairs are collidable (core.py:1134-1141)
    return True


def _joint_desc(jc, index) -> JointDesc:
    ea, eb = jc.entity_a, jc.entity_b
    da = ea.shape.get_delta_from_anchor(jc.anchor_a)
    db = eb.shape.get_delta_from_anchor(jc.anchor_b)
    fr = jc.fixed_rotation
    per_env = hasattr(fr, "numel")
    return JointDesc(
        entity_a=index[id(ea)],
        entity_b=index[id(eb)],
        delta_a=[float(da[0]), float(da[1])],
        delta_b=[float(db[0]), float(db[1])],
        dist=float(jc.dist),
        rotate=bool(jc.rotate),
        fixed_rotation=0.0 if per_env or fr is None else float(fr),
        fixed_rotation_per_env=bool(per_env),
    )


def _lidar_desc(world, agent, sensor, sensor_index, ents, index) -> Optional[LidarDesc]:
    if type(sensor).__name__ != "Lidar" and not any(c.__name__ == "Lidar" for c in type(sensor).__mro__):
        return None
    angles = sensor._angles
    row0 = angles[0].detach().cpu().to(dtype=angles.dtype)
    targets = []
    for e in ents:
        if e is agent or not _truthy(sensor.entity_filter(e)):
            continue
        # core.py:1680: rays are only cast among mutually colliding entities
        if not (e.collides(agent) and agent.collides(e)):
            raise UnsupportedWorld(
                f"Lidar of {agent.name}: target {e.name} does not mutually collide (core.py:1680 asserts)"
            )
        targets.append(index[id(e)])
    return LidarDesc(
        entity=index[id(agent)],
        angles=[float(v) for v in row0.tolist()],
        max_range=float(sensor._max_range),
        targets=targets,
        agent_sensor_index=sensor_index,
    )


def f32(x: float) -> float:
    return float(np.float32(x))


def circumscribed_sum_f32(ra: float, rb: float) -> float:
    """Broad-phase threshold: Python double add, then the fp32 compare of core.py:2796-2800."""
    return float(np.float32(ra + rb))


assert math.isclose(4 / 6e2, 0.006666666666666667)
