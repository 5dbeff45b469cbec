"""URDF -> reduced model (blf_b200.urdf), the stand-in for iDynTree::ModelLoader::loadReducedModelFromFile + setFloatingBase that
every reference caller uses (blf-balancing-position-control.py:103, 197).  The image ships no URDF, so one is written from the
synthetic humanoid description first."""
import numpy as np
import pytest

import blf_b200  # noqa: F401
from blf_b200 import models as M, urdf as U, workloads as W


def _humanoid_urdf():
    links, joints, frames = M._humanoid_description(1.0, True)
    return U.to_urdf("humanoid", links, joints, frames), links, joints, frames


def test_urdf_round_trip_gives_the_same_reduced_model():
    xml, links, joints, frames = _humanoid_urdf()
    for joints_list in (M.ICUB32_JOINTS, M.ICUB26_JOINTS):
        ref = M.reduce_model("ref", "root_link", links, joints, frames, joints_list)
        m = U.load_reduced_model_from_urdf(xml, joints_list)
        assert m.joint_names == ref.joint_names and m.link_names == ref.link_names
        for a in ("parent", "dof"):
            assert np.array_equal(getattr(m, a), getattr(ref, a))
        for a in ("Xr", "Xp", "axis", "mass", "com", "q_lower", "q_upper"):
            assert np.allclose(getattr(m, a), getattr(ref, a), rtol=0, atol=1e-14), a
        assert np.allclose(m.inertia, ref.inertia, rtol=0, atol=1e-13)
        for f, name in enumerate(ref.frame_names):          # same frames, possibly in another order
            g = m.frame_names.index(name)
            assert m.frame_link[g] == ref.frame_link[f]
            assert np.allclose(m.Fr[g], ref.Fr[f], atol=1e-14) and np.allclose(m.Fp[g], ref.Fp[f], atol=1e-14)


def test_fixed_and_frozen_joints_are_lumped_and_prismatic_dofs_refused():
    xml, *_ = _humanoid_urdf()
    few = M.ICUB32_JOINTS[:6]
    m = U.load_reduced_model_from_urdf(xml, few)
    full = U.load_reduced_model_from_urdf(xml, M.ICUB32_JOINTS)
    assert m.n == 6 and abs(m.mass.sum() - full.mass.sum()) < 1e-12       # frozen subtrees are lumped, no mass is lost
    with pytest.raises(KeyError):
        U.load_reduced_model_from_urdf(xml, ["no_such_joint"])
    bad = xml.replace('name="torso_pitch" type="revolute"', 'name="torso_pitch" type="prismatic"')
    with pytest.raises(ValueError):
        U.load_reduced_model_from_urdf(bad, M.ICUB32_JOINTS)


@pytest.mark.parametrize("base", ["l_foot", "chest", "r_upper_arm"])
def test_set_floating_base_describes_the_same_mechanism(base):
    """Re-rooted at another link, the model must put every frame at the same place RELATIVE to any other frame for the same joint
    positions, keep its total mass and its centre of mass (seen from a frame)."""
    xml, *_ = _humanoid_urdf()
    a = U.load_reduced_model_from_urdf(xml, M.ICUB32_JOINTS)
    b = U.load_reduced_model_from_urdf(xml, M.ICUB32_JOINTS, base_link=base)
    assert b.link_names[0].startswith(base) or b.link_names[0] == base
    assert abs(a.mass.sum() - b.mass.sum()) < 1e-12 and b.joint_names == a.joint_names
    rng = np.random.default_rng(5)
    N = 6
    s = rng.uniform(-0.8, 0.8, (N, a.n))
    p, q = np.zeros((N, 3)), np.tile([0.0, 0.0, 0.0, 1.0], (N, 1))
    Ra, pa = W.fk_batch(a, p, q, s)
    Rb, pb = W.fk_batch(b, p, q, s)
    names = ["root_link", "chest", "l_sole", "r_sole", "l_hand_palm", "head", base]
    pose_a = {nm: W.frame_pose_batch(a, Ra, pa, a.frame_names.index(nm)) for nm in names}
    pose_b = {nm: W.frame_pose_batch(b, Rb, pb, b.frame_names.index(nm)) for nm in names}
    ca, cb = W.com_batch(a, Ra, pa), W.com_batch(b, Rb, pb)
    for x in names:
        Rxa, pxa = pose_a[x]
        Rxb, pxb = pose_b[x]
        for y in names:
            Rya, pya = pose_a[y]
            Ryb, pyb = pose_b[y]
            rel_a = np.einsum("nji,njk->nik", Rxa, Rya), np.einsum("nji,nj->ni", Rxa, pya - pxa)
            rel_b = np.einsum("nji,njk->nik", Rxb, Ryb), np.einsum("nji,nj->ni", Rxb, pyb - pxb)
            assert np.abs(rel_a[0] - rel_b[0]).max() < 1e-12 and np.abs(rel_a[1] - rel_b[1]).max() < 1e-12, (x, y)
        assert np.abs(np.einsum("nji,nj->ni", Rxa, ca - pxa) - np.einsum("nji,nj->ni", Rxb, cb - pxb)).max() < 1e-12
