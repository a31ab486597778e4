#!/usr/bin/env python
"""Generate tests/golden/ref_coloring.npz from the REFERENCE's own colouring code (OrderedGreedy::color, Mcs::color,
GraphColor::balanceColoredCategories: Simulator/Modules/GraphColoring/{gready,mcs,ColoringAlgorithms}.cpp compiled into
oracle/_ref/libgaia_ref.so by oracle/Makefile). Run in the build container only:

    python tests/golden/make_coloring_golden.py

Graphs: seeded random graphs and the vertex graphs (TetMeshVertexGraph: vertices sharing a tet) of two reference meshes.
balanceColoredCategories does not terminate on every graph, so each balanced colouring is computed in a child process with a
time limit; graphs where it ran out are recorded as such (key `<name>_balance_ok` = 0) and the test skips their balanced form.
"""
import ctypes as C
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from golden_inputs import coloring_graphs  # noqa: E402


def ref_color(n, edges, algorithm, balance, ratio, q=None):
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgaia_ref.so"))
    IP = C.POINTER(C.c_int)
    lib.ref_color_graph.argtypes = [C.c_int, C.c_longlong, IP, C.c_int, C.c_int, C.c_float, IP]
    e = np.ascontiguousarray(edges, np.int32)
    out = np.zeros(n, np.int32)
    devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1); os.dup2(devnull, 1)      # the reference prints progress
    try:
        k = lib.ref_color_graph(n, len(e), e.ctypes.data_as(IP), algorithm, balance, ratio, out.ctypes.data_as(IP))
    finally:
        os.dup2(saved, 1)
    if q is not None:
        q.put((k, out))
    return k, out


def main():
    out = {}
    for name, (n, edges) in coloring_graphs().items():
        for alg, tag in ((0, "greedy"), (1, "mcs")):
            k, c = ref_color(n, edges, alg, 0, 0.0)
            out[f"{name}_{tag}"] = c.astype(np.int16)
            q = mp.Queue()
            p = mp.Process(target=ref_color, args=(n, edges, alg, 1, 1.05, q))
            p.start()
            p.join(30)
            if p.is_alive():
                p.kill(); p.join()
                out[f"{name}_{tag}_balance_ok"] = np.zeros(1, np.int8)
                print(name, tag, "colours", k, "balance: no termination within 30 s")
            else:
                kb, cb = q.get()
                out[f"{name}_{tag}_balance_ok"] = np.ones(1, np.int8)
                out[f"{name}_{tag}_balanced"] = cb.astype(np.int16)
                print(name, tag, "colours", k, "balanced sizes", np.bincount(cb).tolist())
    path = os.path.join(ROOT, "tests", "golden", "ref_coloring.npz")
    np.savez_compressed(path, **out)
    print("wrote", path)


if __name__ == "__main__":
    main()
