"""A sheet folded onto itself with two seams and, optionally, the contact pairs those seams exclude (the reference's
filteredPairsPath): shared by the CPU test of the oracle and the GPU parity test."""
import numpy as np

from gaia_b200 import cloth

N = 14
SEAMS = np.array([[0, 13, 27], [100, 117, 131]], np.int32)
RATIOS = np.array([0.5, 0.25], np.float32)


def folded_sheet():
    v, f = cloth.grid_cloth(N, 1.0, origin=(0, 1.0, 0))
    x = v[:, 0].copy()
    fold = x > 0.1
    v[fold, 0] = 0.2 - x[fold]                       # fold the right part over the left
    v[fold, 1] += 0.024
    v += np.random.RandomState(11).uniform(-0.003, 0.003, v.shape).astype(np.float32)
    return v, f


def build(cls, kw, seams=True, vf_pairs=(), ee_pairs=()):
    v, f = folded_sheet()
    s = cls(cloth.default_cloth_physics_params(contactRadius=0.02, maxQueryDis=0.04), **kw)
    s.add_tri_mesh(v, f, cloth.default_cloth_object_params(), cloth.grid_cloth_coloring(N))
    if seams:
        s.set_sewing(0, SEAMS, RATIOS, 100.0)
    if len(vf_pairs) or len(ee_pairs):
        s.set_filtered_pairs(0, vf_pairs, ee_pairs)
    s.finalize()
    return s


def pick_pairs(vf, ee, every=7):
    """every `every`-th detected self contact becomes a filtered pair: (vertex, face) rows of vf_contacts(), (edge1, edge2) of ee_contacts()"""
    vf_pairs = np.ascontiguousarray(vf[::every][:, [1, 3]], np.int32)
    ee_pairs = np.ascontiguousarray(ee[::every][:, [1, 3]], np.int32)
    return vf_pairs, ee_pairs
