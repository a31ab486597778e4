"""CPU: the reference-held fixtures under tests/golden/ref_root (the meshes and the S01 scene BASELINE.json's configs 1 and
2 name) load through the product's host code and the oracle, unchanged, on any box (no /root/reference needed)."""
import os
import subprocess

import numpy as np
import pytest

from gaia_b200 import meshio
from gaia_b200.workloads import REF_MESHES, REF_ROOT, build_scene, ref_mesh, workload_ref_stack, workload_s01

HOST_BIN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gaia_b200", "host", "gaia_vbd_b200")


@pytest.mark.parametrize("name", sorted(REF_MESHES))
def test_shipped_colourings_are_valid(name):
    v, t, colors = ref_mesh(name)
    assert meshio.check_coloring(len(v), t, colors)
    assert sum(len(c) for c in colors) == len(v)


def test_fixtures_are_the_reference_files():
    if not meshio.reference_available():
        pytest.skip("/root/reference only exists in the build container")
    for rel in list(REF_MESHES.values()):
        for suffix in ("", ".vertexColoring.json"):
            with open(os.path.join(REF_ROOT, rel + suffix), "rb") as a, open(os.path.join("/root/reference", rel + suffix), "rb") as b:
                assert a.read() == b.read(), rel + suffix


def test_s01_scene_parses_unchanged_from_the_fixture_root():
    d = os.path.join(REF_ROOT, "S01")
    r = subprocess.run([HOST_BIN, d + "/Models.json", d + "/Parameters.json", "/tmp/unused", "-R", REF_ROOT, "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "32 models, 87339 vertices, 330335 tets, numFrames 720, numSubsteps 4, iterations 20, outputExt ply" in r.stdout
    wl = workload_s01()
    assert len(wl["meshes"]) == 32 and [m[2].frameToAppear for m in wl["meshes"][:3]] == [120, 126, 132]
    assert workload_s01(first_frame=1)["meshes"][31][2].frameToAppear == 1 + 6 * 31


def test_oracle_active_lists_and_relations_are_consistent():
    """row a15 on the oracle side: every contact with a closest face appears once in the group of its query vertex and in the
    groups of its three face vertices; every such contact leaves exactly three face-side relations."""
    from oracle.binding import OracleScene
    wl = workload_ref_stack(models=("teapot", "bunny_small"), substeps=2, iterations=4)
    o = build_scene(OracleScene, wl, threads=2)
    o.run_iterations(0)
    c = o.contacts()
    found = c[c["shortestPathFound"] == 1]
    assert len(found) > 5
    groups = o.parallel_groups()
    grp_of = [dict() for _ in range(2)]
    for gi, g in enumerate(groups):
        for m, v in g:
            grp_of[m][int(v)] = gi
    faces = [o.surface_faces(m) for m in range(2)]
    expect = [set() for _ in groups]
    rel_expect = [[] for _ in range(2)]
    for k in found:
        qm, qv, hm, hf = int(k["queryMesh"]), int(k["queryVertex"]), int(k["hitMesh"]), int(k["hitFace"])
        expect[grp_of[qm][qv]].add((qm, qv))
        for order, fv in enumerate(faces[hm][hf]):
            expect[grp_of[hm][int(fv)]].add((qm, qv))
            rel_expect[hm].append((int(fv), qm, qv, order))
    for gi in range(len(groups)):
        assert [tuple(x) for x in o.active_list(gi)] == sorted(expect[gi])
    for m in range(2):
        assert [tuple(x) for x in o.relations(m)] == sorted(rel_expect[m])
