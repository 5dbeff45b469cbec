"""URDF -> reduced kinematic model, the way the reference's callers obtain theirs.

Every reference caller builds its model with ``iDynTree::ModelLoader::loadReducedModelFromFile(urdf, joints_list)`` and may
re-root it with ``setFloatingBase`` / ``getFloatingBase`` (utilities/balancing-position-control/script/
blf-balancing-position-control.py:103, 197; src/IK/tests/QPInverseKinematicsTest.cpp builds iDynTree models the same way).
This module does the same on the host, in plain Python (xml.etree): the joints named in ``joints_list`` become the DoFs, in
that order; every other joint (fixed ones, and movable ones frozen at zero) is lumped into its parent, inertial parameters
included, exactly as ``models.reduce_model`` already does for the synthetic humanoids; every URDF link -- the mass-less "frame"
links attached by fixed joints too -- stays addressable as a frame.  The result is a ``models.KinematicModel``, i.e. a
``blfik_model_desc``.  Prismatic joints in ``joints_list`` are refused (the kernels are revolute-only); continuous joints get
+-pi limits.
"""
from __future__ import annotations

import math
import xml.etree.ElementTree as ET
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import models as M


def _vec(text: Optional[str], default=(0.0, 0.0, 0.0)) -> Tuple[float, ...]:
    return tuple(float(v) for v in text.split()) if text else tuple(default)


def parse_urdf(source: str):
    """``source``: a path or the XML itself.  Returns (robot name, links, joints, root link name); links / joints are
    ``models.LinkDesc`` / ``models.JointDesc`` lists, fixed joints carry ``kind == 'fixed'``."""
    root = ET.fromstring(source) if source.lstrip().startswith("<") else ET.parse(source).getroot()
    if root.tag != "robot":
        raise ValueError("not a URDF: the root element is not <robot>")
    links: List[M.LinkDesc] = []
    for le in root.findall("link"):
        mass, com, inertia = 0.0, (0.0, 0.0, 0.0), None
        ie = le.find("inertial")
        if ie is not None:
            me, oe, te = ie.find("mass"), ie.find("origin"), ie.find("inertia")
            mass = float(me.get("value")) if me is not None else 0.0
            com = _vec(oe.get("xyz")) if oe is not None else com
            if te is not None:
                g = lambda k: float(te.get(k, "0"))
                I = np.array([[g("ixx"), g("ixy"), g("ixz")], [g("ixy"), g("iyy"), g("iyz")], [g("ixz"), g("iyz"), g("izz")]])
                R = M.rpy_to_rot(*_vec(oe.get("rpy"))) if oe is not None else np.eye(3)
                inertia = R @ I @ R.T            # URDF gives it in the inertial frame; the model keeps it in the link frame
            else:
                inertia = np.zeros((3, 3))
        else:
            inertia = np.zeros((3, 3))
        links.append(M.LinkDesc(le.get("name"), mass, com, inertia))
    joints: List[M.JointDesc] = []
    kinds: Dict[str, str] = {}
    children = set()
    for je in root.findall("joint"):
        kind = je.get("type")
        oe, ae, lim = je.find("origin"), je.find("axis"), je.find("limit")
        xyz = _vec(oe.get("xyz")) if oe is not None else (0.0, 0.0, 0.0)
        rpy = _vec(oe.get("rpy")) if oe is not None else (0.0, 0.0, 0.0)
        axis = _vec(ae.get("xyz"), (1.0, 0.0, 0.0)) if ae is not None else (1.0, 0.0, 0.0)
        lo, hi = -math.pi, math.pi
        if kind == "revolute" and lim is not None:
            lo, hi = float(lim.get("lower", -math.pi)), float(lim.get("upper", math.pi))
        nominal = min(max(0.0, lo), hi)
        j = M.JointDesc(je.get("name"), je.find("parent").get("link"), je.find("child").get("link"), xyz, rpy, axis, lo, hi, nominal)
        joints.append(j)
        kinds[j.name] = kind
        children.add(j.child)
    roots = [l.name for l in links if l.name not in children]
    if len(roots) != 1:
        raise ValueError(f"a URDF tree has exactly one root link, found {roots}")
    return root.get("name", "robot"), links, joints, roots[0], kinds


def _reroot(links: List[M.LinkDesc], joints: List[M.JointDesc], old_root: str, new_root: str):
    """iDynTree's setFloatingBase: the same mechanism described from another link.  The joints on the path new root -> old root
    are traversed backwards.  URDF semantics put a joint's axis at the CHILD's frame origin; turning joint (p -> c, origin T, axis
    a) around therefore moves p's link frame onto the joint frame (everything attached to p is re-expressed through T^-1), makes
    c the parent with an identity origin, and flips the axis so that the joint coordinate keeps its meaning."""
    if new_root == old_root:
        return links, joints
    by_child = {j.child: j for j in joints}
    if new_root not in by_child and new_root != old_root:
        raise KeyError(f"link '{new_root}' not in the model")
    path = []
    k = new_root
    while k != old_root:
        path.append(by_child[k])
        k = by_child[k].parent
    links = {l.name: M.LinkDesc(l.name, l.mass, l.com, None if l.inertia is None else np.array(l.inertia)) for l in links}
    joints = [M.JointDesc(j.name, j.parent, j.child, j.xyz, j.rpy, j.axis, j.lower, j.upper, j.nominal) for j in joints]
    # pose of each link's NEW frame expressed in its OLD frame (identity unless the link's frame has to move)
    shift: Dict[str, Tuple[np.ndarray, np.ndarray]] = {}
    for j in path:            # from the new root towards the old one: j connects parent p (towards the old root) to child c
        p = j.parent
        R, t = M.rpy_to_rot(*j.rpy), np.asarray(j.xyz, dtype=float)     # joint frame (= c's frame at q = 0) in p's OLD frame
        shift[p] = (R, t)
    new_joints = []
    on_path = {j.name for j in path}
    for j in joints:
        Rp, tp = shift.get(j.parent, (np.eye(3), np.zeros(3)))
        R, t = M.rpy_to_rot(*j.rpy), np.asarray(j.xyz, dtype=float)
        if j.name in on_path:
            # reversed: parent = old child (its frame does not move unless it is itself shifted by the next joint of the path),
            # child = old parent whose new frame sits on this joint's frame: identity origin seen from the old child's OLD frame
            Rc, tc = shift.get(j.child, (np.eye(3), np.zeros(3)))
            Rn, tn = Rc.T, -Rc.T @ tc                                     # old child's OLD frame in its NEW frame
            a = -(np.asarray(j.axis, dtype=float))
            new_joints.append(_joint(j, parent=j.child, child=j.parent, R=Rn, t=tn, axis=a))
        else:
            Rn, tn = Rp.T @ R, Rp.T @ (t - tp)                            # origin re-expressed in the parent's NEW frame
            new_joints.append(_joint(j, parent=j.parent, child=j.child, R=Rn, t=tn, axis=np.asarray(j.axis, dtype=float)))
    for name, (Rs, ts) in shift.items():
        l = links[name]
        c = np.asarray(l.com, dtype=float)
        l.com = tuple(Rs.T @ (c - ts))
        if l.inertia is not None:
            l.inertia = Rs.T @ np.asarray(l.inertia) @ Rs
    # A link's own frame is intrinsic (iDynTree does not move it when the base changes): the shifted links get an internal
    # name, and their own name stays with the old frame, now an additional frame of theirs.
    moved = lambda nm: nm + "__base_side" if nm in shift else nm
    extra = [M.FrameDesc(name, moved(name), tuple(-Rs.T @ ts), _rot_to_rpy(Rs.T)) for name, (Rs, ts) in shift.items()]
    for l in links.values():
        l.name = moved(l.name)
    for j in new_joints:
        j.parent, j.child = moved(j.parent), moved(j.child)
    return list(links.values()), new_joints, extra


def _joint(j: M.JointDesc, parent: str, child: str, R: np.ndarray, t: np.ndarray, axis: np.ndarray) -> M.JointDesc:
    return M.JointDesc(j.name, parent, child, tuple(float(v) for v in t), _rot_to_rpy(R), tuple(float(v) for v in axis), j.lower, j.upper, j.nominal)


def _rot_to_rpy(R: np.ndarray) -> Tuple[float, float, float]:
    """inverse of models.rpy_to_rot (R = Rz(y) Ry(p) Rx(r))"""
    p = -math.asin(max(-1.0, min(1.0, float(R[2, 0]))))
    if abs(math.cos(p)) > 1e-12:
        return math.atan2(R[2, 1], R[2, 2]), p, math.atan2(R[1, 0], R[0, 0])
    return 0.0, p, math.atan2(-R[0, 1], R[1, 1])


def load_reduced_model_from_urdf(source: str, joints_list: Sequence[str], base_link: Optional[str] = None,
                                 name: Optional[str] = None) -> M.KinematicModel:
    """ModelLoader::loadReducedModelFromFile(urdf, joints_list) [+ setFloatingBase(base_link)].  ``source`` is a file path or the
    XML text.  Joints outside ``joints_list`` are frozen at zero and lumped; a prismatic joint inside it is an error."""
    robot, links, joints, root, kinds = parse_urdf(source)
    for jn in joints_list:
        if jn not in kinds:
            raise KeyError(f"joint '{jn}' not in the URDF")
        if kinds[jn] not in ("revolute", "continuous"):
            raise ValueError(f"joint '{jn}' is {kinds[jn]}: only revolute joints can be degrees of freedom of this model")
    frames: List[M.FrameDesc] = []
    if base_link is not None and base_link != root:
        links, joints, frames = _reroot(links, joints, root, base_link)
        root = base_link
    # massless links keep a zero inertia (reduce_model would otherwise give them its deterministic stand-in)
    links = [M.LinkDesc(l.name, l.mass, l.com, np.zeros((3, 3)) if l.inertia is None else l.inertia) for l in links]
    return M.reduce_model(name or robot, root, links, joints, frames, list(joints_list))


def to_urdf(name: str, links: Sequence[M.LinkDesc], joints: Sequence[M.JointDesc], frames: Sequence[M.FrameDesc] = ()) -> str:
    """A URDF for a link / joint description (tests use it to have a URDF at all: the image ships none).  Frames become
    mass-less links behind fixed joints, as iDynTree's exporter writes them."""
    out = [f'<robot name="{name}">']
    for l in links:
        I = (M.default_inertia(l.mass, l.com) if l.inertia is None else np.asarray(l.inertia)).astype(float).tolist()
        I = [[float(v) for v in row] for row in I]
        out.append(f'  <link name="{l.name}"><inertial><origin xyz="{float(l.com[0])!r} {float(l.com[1])!r} {float(l.com[2])!r}" rpy="0 0 0"/>'
                   f'<mass value="{float(l.mass)!r}"/><inertia ixx="{I[0][0]!r}" ixy="{I[0][1]!r}" ixz="{I[0][2]!r}" iyy="{I[1][1]!r}" '
                   f'iyz="{I[1][2]!r}" izz="{I[2][2]!r}"/></inertial></link>')
    for j in joints:
        out.append(f'  <joint name="{j.name}" type="revolute"><parent link="{j.parent}"/><child link="{j.child}"/>'
                   f'<origin xyz="{float(j.xyz[0])!r} {float(j.xyz[1])!r} {float(j.xyz[2])!r}" rpy="{float(j.rpy[0])!r} {float(j.rpy[1])!r} {float(j.rpy[2])!r}"/>'
                   f'<axis xyz="{float(j.axis[0])!r} {float(j.axis[1])!r} {float(j.axis[2])!r}"/>'
                   f'<limit lower="{float(j.lower)!r}" upper="{float(j.upper)!r}" effort="100" velocity="10"/></joint>')
    for f in frames:
        out.append(f'  <link name="{f.name}"/>')
        out.append(f'  <joint name="{f.name}_fixed_joint" type="fixed"><parent link="{f.link}"/><child link="{f.name}"/>'
                   f'<origin xyz="{float(f.xyz[0])!r} {float(f.xyz[1])!r} {float(f.xyz[2])!r}" rpy="{float(f.rpy[0])!r} {float(f.rpy[1])!r} {float(f.rpy[2])!r}"/></joint>')
    out.append("</robot>")
    return "\n".join(out) + "\n"
