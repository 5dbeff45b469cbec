"""Kinematic-tree model description and synthetic robots.

The reference gets its robots from URDF files through
``iDynTree::ModelLoader::loadReducedModelFromFile(urdf, joints_list)`` (e.g.
``utilities/balancing-position-control/script/blf-balancing-position-control.py:99-106``,
``bindings/python/IK/tests/test_QP_inverse_kinematics.py:59-71``).  No robot URDF
exists in the reference tree or in this image, so the robots here
keep the reference's *topology, joint names, joint order and frame names* and use
synthetic, humanoid-sized geometry and inertias.

A model is a flattened tree, structure-of-arrays, link 0 = floating base,
``parent[k] < k``; link k (k >= 1) is the child of the revolute joint with DoF
index ``dof[k]``::

    W_H_k = W_H_parent(k) * (Xr[k], Xp[k]) * Rot(axis[k], s[dof[k]])

``reduce_model`` restates the reduced-model semantics: joints that are not in
``joints_list`` are frozen at zero and their child links lumped into the parent
(mass, CoM and attached frames), and the DoF order is the order of
``joints_list``.
"""
from __future__ import annotations

import dataclasses
import math
from typing import Optional, Dict, List, Sequence, Tuple

import numpy as np


def rpy_to_rot(r: float, p: float, y: float) -> np.ndarray:
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return rz @ ry @ rx


@dataclasses.dataclass
class KinematicModel:
    """Flattened revolute kinematic tree (see module docstring)."""

    name: str
    joint_names: List[str]          # DoF order
    link_names: List[str]           # [n+1]
    parent: np.ndarray              # int32 [n+1]
    dof: np.ndarray                 # int32 [n+1]
    Xr: np.ndarray                  # [n+1,3,3]
    Xp: np.ndarray                  # [n+1,3]
    axis: np.ndarray                # [n+1,3]
    mass: np.ndarray                # [n+1]
    com: np.ndarray                 # [n+1,3]
    frame_names: List[str]
    frame_link: np.ndarray          # int32 [nf]
    Fr: np.ndarray                  # [nf,3,3]
    Fp: np.ndarray                  # [nf,3]
    q_lower: np.ndarray             # [n]
    q_upper: np.ndarray             # [n]
    q_nominal: np.ndarray           # [n]
    inertia: Optional[np.ndarray] = None   # [n+1,3,3] rotational inertia about the link CoM, link frame (AngularMomentumTask)

    def __post_init__(self):
        if self.inertia is None:
            self.inertia = np.zeros((len(self.joint_names) + 1, 3, 3))

    @property
    def n(self) -> int:
        return len(self.joint_names)

    @property
    def nv(self) -> int:
        return self.n + 6

    def frame_index(self, name: str) -> int:
        try:
            return self.frame_names.index(name)
        except ValueError:
            raise KeyError(f"frame '{name}' not in model '{self.name}'") from None

    def validate(self) -> None:
        n = self.n
        assert self.parent.shape == (n + 1,) and self.parent[0] == -1
        assert all(0 <= self.parent[k] < k for k in range(1, n + 1)), "links must be topologically ordered"
        assert sorted(self.dof[1:].tolist()) == list(range(n)), "dof must be a permutation"
        assert np.allclose(np.linalg.norm(self.axis[1:], axis=1), 1.0)
        for k in range(n + 1):
            assert np.allclose(self.Xr[k] @ self.Xr[k].T, np.eye(3), atol=1e-12)

    def depth(self) -> np.ndarray:
        d = np.zeros(self.n + 1, dtype=np.int32)
        for k in range(1, self.n + 1):
            d[k] = d[self.parent[k]] + 1
        return d

    # ---- plain-text exchange format read by the C++ host (ModelLoader) ----
    def to_text(self) -> str:
        out = [f"blfik_model 1", f"name {self.name}", f"n {self.n}", f"nf {len(self.frame_names)}"]
        for k in range(self.n + 1):
            jn = self.joint_names[self.dof[k]] if k > 0 else "-"
            vals = [*self.Xr[k].ravel(), *self.Xp[k], *self.axis[k], self.mass[k], *self.com[k]]
            out.append(f"link {self.link_names[k]} {self.parent[k]} {self.dof[k]} {jn} " + " ".join(repr(float(v)) for v in vals))
        for k in range(self.n + 1):
            I = self.inertia[k]
            vals = [I[0, 0], I[0, 1], I[0, 2], I[1, 1], I[1, 2], I[2, 2]]
            out.append(f"inertia {self.link_names[k]} {k} " + " ".join(repr(float(v)) for v in vals))
        for f, fn in enumerate(self.frame_names):
            vals = [*self.Fr[f].ravel(), *self.Fp[f]]
            out.append(f"frame {fn} {self.frame_link[f]} " + " ".join(repr(float(v)) for v in vals))
        for j in range(self.n):
            out.append(f"limits {self.joint_names[j]} {float(self.q_lower[j])!r} {float(self.q_upper[j])!r} {float(self.q_nominal[j])!r}")
        return "\n".join(out) + "\n"


# --------------------------------------------------------------------------
# full (unreduced) description -> reduced flattened model
# --------------------------------------------------------------------------
@dataclasses.dataclass
class JointDesc:
    name: str
    parent: str
    child: str
    xyz: Tuple[float, float, float]
    rpy: Tuple[float, float, float]
    axis: Tuple[float, float, float]
    lower: float = -1.0
    upper: float = 1.0
    nominal: float = 0.0


@dataclasses.dataclass
class LinkDesc:
    name: str
    mass: float
    com: Tuple[float, float, float]
    inertia: Optional[np.ndarray] = None   # 3x3 about the CoM in the link frame; None: default_inertia(mass, com)


def default_inertia(mass: float, com) -> np.ndarray:
    """Deterministic stand-in for a URDF <inertia>: a box with half extents 4 cm + |com| per axis, with mild products of
    inertia so that the full symmetric tensor is exercised (positive definite by diagonal dominance)."""
    h = 0.04 + np.abs(np.asarray(com, dtype=float))
    d = mass / 3.0 * np.array([h[1] ** 2 + h[2] ** 2, h[0] ** 2 + h[2] ** 2, h[0] ** 2 + h[1] ** 2])
    I = np.diag(d)
    for (a, b), f in (((0, 1), 0.10), ((0, 2), -0.07), ((1, 2), 0.05)):
        I[a, b] = I[b, a] = f * math.sqrt(d[a] * d[b])
    return I


@dataclasses.dataclass
class FrameDesc:
    name: str
    link: str
    xyz: Tuple[float, float, float]
    rpy: Tuple[float, float, float] = (0.0, 0.0, 0.0)


def reduce_model(name: str, base_link: str, links: Sequence[LinkDesc], joints: Sequence[JointDesc],
                 frames: Sequence[FrameDesc], joints_list: Sequence[str]) -> KinematicModel:
    """Reduced model with the DoFs of ``joints_list`` in that order; every other
    joint is frozen at 0 and its child link lumped into the parent."""
    link_by_name = {l.name: l for l in links}
    joint_by_name = {j.name: j for j in joints}
    for jn in joints_list:
        if jn not in joint_by_name:
            raise KeyError(f"joint '{jn}' not in the full model")
    children: Dict[str, List[JointDesc]] = {}
    for j in joints:
        children.setdefault(j.parent, []).append(j)

    # reduced link records: name -> dict(mass, mcom (mass*com), frames[(name,R,p)])
    rl_names: List[str] = []
    rl_parent: List[int] = []
    rl_dof: List[int] = []
    rl_Xr: List[np.ndarray] = []
    rl_Xp: List[np.ndarray] = []
    rl_axis: List[np.ndarray] = []
    rl_mass: List[float] = []
    rl_mcom: List[np.ndarray] = []
    rl_IO: List[np.ndarray] = []   # rotational inertia about the reduced link's frame origin (parallel-axis sums)
    fr_out: List[Tuple[str, int, np.ndarray, np.ndarray]] = []
    frames_by_link: Dict[str, List[FrameDesc]] = {}
    for f in frames:
        frames_by_link.setdefault(f.link, []).append(f)
    dof_index = {jn: i for i, jn in enumerate(joints_list)}

    def new_reduced_link(nm, parent, dof, Xr, Xp, axis):
        rl_names.append(nm), rl_parent.append(parent), rl_dof.append(dof)
        rl_Xr.append(Xr), rl_Xp.append(Xp), rl_axis.append(axis)
        rl_mass.append(0.0), rl_mcom.append(np.zeros(3)), rl_IO.append(np.zeros((3, 3)))
        return len(rl_names) - 1

    # BFS so that parent index < child index
    queue: List[Tuple[str, int, np.ndarray, np.ndarray]] = []
    root = new_reduced_link(base_link, -1, -1, np.eye(3), np.zeros(3), np.array([0.0, 0.0, 1.0]))
    queue.append((base_link, root, np.eye(3), np.zeros(3)))
    while queue:
        lname, ridx, R_l, p_l = queue.pop(0)  # pose of this full link inside reduced link ridx
        ld = link_by_name[lname]
        rl_mass[ridx] += ld.mass
        c_l = R_l @ np.asarray(ld.com) + p_l
        rl_mcom[ridx] = rl_mcom[ridx] + ld.mass * c_l
        I_l = default_inertia(ld.mass, ld.com) if ld.inertia is None else np.asarray(ld.inertia, dtype=float)
        rl_IO[ridx] = rl_IO[ridx] + R_l @ I_l @ R_l.T + ld.mass * (c_l @ c_l * np.eye(3) - np.outer(c_l, c_l))
        # every full link is also exposed as a frame (iDynTree: link frames are frames)
        fr_out.append((lname, ridx, R_l.copy(), p_l.copy()))
        for f in frames_by_link.get(lname, []):
            fr_out.append((f.name, ridx, R_l @ rpy_to_rot(*f.rpy), R_l @ np.asarray(f.xyz) + p_l))
        for j in children.get(lname, []):
            Rj = R_l @ rpy_to_rot(*j.rpy)
            pj = R_l @ np.asarray(j.xyz) + p_l
            if j.name in dof_index:
                a = np.asarray(j.axis, dtype=float)
                a = a / np.linalg.norm(a)
                cidx = new_reduced_link(j.child, ridx, dof_index[j.name], Rj, pj, a)
                queue.append((j.child, cidx, np.eye(3), np.zeros(3)))
            else:
                queue.append((j.child, ridx, Rj, pj))  # frozen at 0 -> lumped

    n = len(joints_list)
    assert len(rl_names) == n + 1, "joints_list must only name joints reachable from the base"
    mass = np.array(rl_mass)
    com = np.array([mc / m if m > 0 else np.zeros(3) for mc, m in zip(rl_mcom, rl_mass)])
    inertia = np.array([IO - m * (c @ c * np.eye(3) - np.outer(c, c)) for IO, m, c in zip(rl_IO, rl_mass, com)])
    model = KinematicModel(
        name=name, joint_names=list(joints_list), link_names=rl_names,
        parent=np.array(rl_parent, dtype=np.int32), dof=np.array(rl_dof, dtype=np.int32),
        Xr=np.array(rl_Xr), Xp=np.array(rl_Xp), axis=np.array(rl_axis), mass=mass, com=com,
        frame_names=[f[0] for f in fr_out], frame_link=np.array([f[1] for f in fr_out], dtype=np.int32),
        Fr=np.array([f[2] for f in fr_out]), Fp=np.array([f[3] for f in fr_out]),
        q_lower=np.array([joint_by_name[j].lower for j in joints_list]),
        q_upper=np.array([joint_by_name[j].upper for j in joints_list]),
        q_nominal=np.array([joint_by_name[j].nominal for j in joints_list]), inertia=inertia)
    model.validate()
    return model


# --------------------------------------------------------------------------
# synthetic humanoids with the reference's joint / frame names
# --------------------------------------------------------------------------
def _humanoid_description(scale: float, with_neck_and_wrists: bool):
    """iCub/ergoCub-like humanoid: x forward, y left, z up; base = root_link."""
    s = scale
    links: List[LinkDesc] = []
    joints: List[JointDesc] = []
    frames: List[FrameDesc] = []

    def L(name, mass, com):
        links.append(LinkDesc(name, mass, tuple(c * s for c in com)))

    def J(name, parent, child, xyz, rpy, axis, nominal=0.0, half_range=1.0):
        joints.append(JointDesc(name, parent, child, tuple(c * s for c in xyz), rpy, axis,
                                nominal - half_range, nominal + half_range, nominal))

    L("root_link", 6.5, (0.0, 0.0, 0.02))
    # torso: pitch -> roll -> yaw -> chest
    L("torso_1", 1.2, (0.0, 0.0, 0.01)); L("torso_2", 1.0, (0.0, 0.0, 0.01)); L("chest", 9.0, (-0.01, 0.0, 0.14))
    J("torso_pitch", "root_link", "torso_1", (0.0, 0.0, 0.09), (0.0, 0.0, 0.0), (0, 1, 0), 0.05)
    J("torso_roll", "torso_1", "torso_2", (0.0, 0.0, 0.035), (0.0, 0.0, 0.0), (1, 0, 0), 0.0)
    J("torso_yaw", "torso_2", "chest", (0.0, 0.0, 0.03), (0.0, 0.0, 0.0), (0, 0, 1), 0.0)
    # head
    L("neck_1", 0.4, (0, 0, 0.01)); L("neck_2", 0.3, (0, 0, 0.01)); L("head", 2.2, (0.01, 0, 0.07))
    J("neck_pitch", "chest", "neck_1", (0.0, 0.0, 0.30), (0, 0, 0), (0, 1, 0))
    J("neck_roll", "neck_1", "neck_2", (0.0, 0.0, 0.02), (0, 0, 0), (1, 0, 0))
    J("neck_yaw", "neck_2", "head", (0.0, 0.0, 0.03), (0, 0, 0), (0, 0, 1))
    frames.append(FrameDesc("head_imu", "head", (0.02 * s, 0.0, 0.10 * s)))
    for side, sg in (("l", 1.0), ("r", -1.0)):
        # arm: shoulder pitch/roll/yaw, elbow, (wrist prosup/pitch/yaw)
        L(f"{side}_shoulder_1", 0.9, (0, 0.01 * sg, 0)); L(f"{side}_shoulder_2", 0.6, (0, 0, -0.01))
        L(f"{side}_upper_arm", 1.6, (0.0, 0.0, -0.08)); L(f"{side}_forearm", 1.0, (0.0, 0.0, -0.07))
        L(f"{side}_wrist_1", 0.25, (0, 0, -0.01)); L(f"{side}_wrist_2", 0.15, (0, 0, -0.01))
        L(f"{side}_hand", 0.55, (0.0, 0.0, -0.04))
        J(f"{side}_shoulder_pitch", "chest", f"{side}_shoulder_1", (0.0, 0.11 * sg, 0.24), (0.26 * sg, 0, 0), (0, 1, 0), -0.15)
        J(f"{side}_shoulder_roll", f"{side}_shoulder_1", f"{side}_shoulder_2", (0.0, 0.05 * sg, 0.0), (0, 0, 0), (1, 0, 0), 0.35 * sg)
        J(f"{side}_shoulder_yaw", f"{side}_shoulder_2", f"{side}_upper_arm", (0.0, 0.0, -0.05), (0, 0, 0), (0, 0, 1), 0.1 * sg)
        J(f"{side}_elbow", f"{side}_upper_arm", f"{side}_forearm", (0.015, 0.0, -0.17), (0, 0, 0), (0, 1, 0), -0.55)
        J(f"{side}_wrist_prosup", f"{side}_forearm", f"{side}_wrist_1", (0.0, 0.0, -0.14), (0, 0, 0), (0, 0, 1))
        J(f"{side}_wrist_pitch", f"{side}_wrist_1", f"{side}_wrist_2", (0.0, 0.0, -0.02), (0, 0, 0), (0, 1, 0))
        J(f"{side}_wrist_yaw", f"{side}_wrist_2", f"{side}_hand", (0.0, 0.0, -0.02), (0, 0, 0), (1, 0, 0))
        frames.append(FrameDesc(f"{side}_hand_palm", f"{side}_hand", (0.0, 0.0, -0.06 * s), (0.0, 0.0, 0.5 * sg)))
        # leg: hip pitch/roll/yaw, knee, ankle pitch/roll
        L(f"{side}_hip_1", 1.1, (0, 0.01 * sg, 0)); L(f"{side}_hip_2", 0.9, (0, 0, -0.02))
        L(f"{side}_upper_leg", 3.2, (0.0, 0.0, -0.12)); L(f"{side}_lower_leg", 2.4, (0.0, 0.0, -0.13))
        L(f"{side}_ankle_1", 0.7, (0, 0, -0.01)); L(f"{side}_foot", 1.1, (0.025, 0.0, -0.035))
        J(f"{side}_hip_pitch", "root_link", f"{side}_hip_1", (0.0, 0.075 * sg, -0.05), (0, 0, 0), (0, 1, 0), -0.42)
        J(f"{side}_hip_roll", f"{side}_hip_1", f"{side}_hip_2", (0.0, 0.02 * sg, 0.0), (0, 0, 0), (1, 0, 0), 0.03 * sg)
        J(f"{side}_hip_yaw", f"{side}_hip_2", f"{side}_upper_leg", (0.0, 0.0, -0.07), (0, 0, 0), (0, 0, 1), 0.0)
        J(f"{side}_knee", f"{side}_upper_leg", f"{side}_lower_leg", (0.0, 0.0, -0.26), (0, 0, 0), (0, 1, 0), 0.90)
        J(f"{side}_ankle_pitch", f"{side}_lower_leg", f"{side}_ankle_1", (0.0, 0.0, -0.29), (0, 0, 0), (0, 1, 0), -0.48)
        J(f"{side}_ankle_roll", f"{side}_ankle_1", f"{side}_foot", (0.0, 0.0, -0.02), (0, 0, 0), (1, 0, 0), -0.03 * sg)
        frames.append(FrameDesc(f"{side}_sole", f"{side}_foot", (0.03 * s, 0.0, -0.055 * s)))
    frames.append(FrameDesc("chest_imu", "chest", (0.05 * s, 0.0, 0.12 * s)))
    return links, joints, frames


# 23-joint list of utilities/balancing-position-control/config/robots/ergoCubGazeboV1/
#   blf_balancing_position_control/robot_control.ini:3-7
ERGOCUB_JOINTS = [
    "torso_pitch", "torso_roll", "torso_yaw",
    "l_shoulder_pitch", "l_shoulder_roll", "l_shoulder_yaw", "l_elbow",
    "r_shoulder_pitch", "r_shoulder_roll", "r_shoulder_yaw", "r_elbow",
    "l_hip_pitch", "l_hip_roll", "l_hip_yaw", "l_knee", "l_ankle_pitch", "l_ankle_roll",
    "r_hip_pitch", "r_hip_roll", "r_hip_yaw", "r_knee", "r_ankle_pitch", "r_ankle_roll"]

# 32-joint list of bindings/python/IK/tests/test_QP_inverse_kinematics.py:303-307
ICUB32_JOINTS = [
    "l_hip_pitch", "l_hip_roll", "l_hip_yaw", "l_knee", "l_ankle_pitch", "l_ankle_roll",
    "r_hip_pitch", "r_hip_roll", "r_hip_yaw", "r_knee", "r_ankle_pitch", "r_ankle_roll",
    "torso_pitch", "torso_roll", "torso_yaw",
    "l_shoulder_pitch", "l_shoulder_roll", "l_shoulder_yaw", "l_elbow",
    "l_wrist_prosup", "l_wrist_pitch", "l_wrist_yaw",
    "neck_pitch", "neck_roll", "neck_yaw",
    "r_shoulder_pitch", "r_shoulder_roll", "r_shoulder_yaw", "r_elbow",
    "r_wrist_prosup", "r_wrist_pitch", "r_wrist_yaw"]

# 26-joint list: the 32 above without the six wrist pitch/yaw + ... (MANN-style control set)
ICUB26_JOINTS = [j for j in ICUB32_JOINTS if not (j.endswith("wrist_pitch") or j.endswith("wrist_yaw")
                                                  or j.endswith("wrist_prosup"))]


def ergocub23() -> KinematicModel:
    """ergoCub topology, 23 DoFs (nv = 29), synthetic geometry."""
    links, joints, frames = _humanoid_description(1.0, True)
    return reduce_model("ergoCub23_synthetic", "root_link", links, joints, frames, ERGOCUB_JOINTS)


def icub32() -> KinematicModel:
    """iCubGazeboV2_5 topology reduced to the 32 joints of the reference's Python IK test (nv = 38)."""
    links, joints, frames = _humanoid_description(0.68, True)
    return reduce_model("iCub32_synthetic", "root_link", links, joints, frames, ICUB32_JOINTS)


def icub26() -> KinematicModel:
    """iCub topology with 26 DoFs (nv = 32: exactly one warp lane per velocity variable)."""
    links, joints, frames = _humanoid_description(0.68, True)
    return reduce_model("iCub26_synthetic", "root_link", links, joints, frames, ICUB26_JOINTS)


def random_tree(n: int, seed: int = 2, extra_frames: int = 10) -> KinematicModel:
    """Random revolute tree in the spirit of ``iDynTree::getRandomModel`` /
    ``customGetRandomModelWithNoPrismaticJoints`` (src/IK/tests/QPInverseKinematicsTest.cpp:332-355):
    every new link hangs off a random existing link through a revolute joint with
    random placement and axis; ``extra_frames`` additional frames on random links."""
    rng = np.random.default_rng(seed)
    links = [LinkDesc("link0", float(rng.uniform(0.5, 3.0)), tuple(rng.uniform(-0.05, 0.05, 3)))]
    joints: List[JointDesc] = []
    for k in range(1, n + 1):
        par = int(rng.integers(0, k))
        links.append(LinkDesc(f"link{k}", float(rng.uniform(0.2, 3.0)), tuple(rng.uniform(-0.08, 0.08, 3))))
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        joints.append(JointDesc(f"joint{k - 1}", f"link{par}", f"link{k}",
                                tuple(rng.uniform(-0.25, 0.25, 3)), tuple(rng.uniform(-math.pi, math.pi, 3)),
                                tuple(axis), -1.5, 1.5, 0.0))
    frames = [FrameDesc(f"frame{i}", f"link{int(rng.integers(0, n + 1))}", tuple(rng.uniform(-0.2, 0.2, 3)),
                        tuple(rng.uniform(-math.pi, math.pi, 3))) for i in range(extra_frames)]
    return reduce_model(f"random{n}_seed{seed}", "link0", links, joints, frames, [f"joint{i}" for i in range(n)])


def model_from_text(text: str) -> KinematicModel:
    """Inverse of ``KinematicModel.to_text``."""
    lines = [l.split() for l in text.strip().splitlines()]
    assert lines[0][0] == "blfik_model"
    name = lines[1][1]
    n = int(lines[2][1])
    link_rows = [l for l in lines if l[0] == "link"]
    frame_rows = [l for l in lines if l[0] == "frame"]
    limit_rows = [l for l in lines if l[0] == "limits"]
    joint_names = [""] * n
    parent, dof, Xr, Xp, axis, mass, com, lnames = [], [], [], [], [], [], [], []
    for l in link_rows:
        lnames.append(l[1]); parent.append(int(l[2])); dof.append(int(l[3]))
        if int(l[3]) >= 0:
            joint_names[int(l[3])] = l[4]
        v = [float(x) for x in l[5:]]
        Xr.append(np.array(v[0:9]).reshape(3, 3)); Xp.append(v[9:12]); axis.append(v[12:15]); mass.append(v[15]); com.append(v[16:19])
    fn, fl, Fr, Fp = [], [], [], []
    for l in frame_rows:
        fn.append(l[1]); fl.append(int(l[2]))
        v = [float(x) for x in l[3:]]
        Fr.append(np.array(v[0:9]).reshape(3, 3)); Fp.append(v[9:12])
    lim = {l[1]: [float(x) for x in l[2:5]] for l in limit_rows}
    inertia = np.zeros((n + 1, 3, 3))
    for l in lines:
        if l[0] == "inertia":
            k, v = int(l[2]), [float(x) for x in l[3:9]]
            inertia[k] = [[v[0], v[1], v[2]], [v[1], v[3], v[4]], [v[2], v[4], v[5]]]
    m = KinematicModel(name, joint_names, lnames, np.array(parent, dtype=np.int32), np.array(dof, dtype=np.int32),
                       np.array(Xr), np.array(Xp, dtype=float), np.array(axis, dtype=float), np.array(mass), np.array(com, dtype=float),
                       fn, np.array(fl, dtype=np.int32), np.array(Fr), np.array(Fp, dtype=float),
                       np.array([lim[j][0] for j in joint_names]), np.array([lim[j][1] for j in joint_names]),
                       np.array([lim[j][2] for j in joint_names]), inertia)
    m.validate()
    return m
