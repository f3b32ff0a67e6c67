"""Generates tests/golden/mesh_<tutorial>.npz from the reference's own ASCII polyMeshes
(/root/reference/tutorials/CFD/<tutorial>/constant/polyMesh): 04unsteadyNS (11 644 cells, unstructured, a cylinder and two
`empty` patches), 01POD (10 000 hex cells, the lid-driven cavity block) and 03steadyNS (9 000 cells, backward-facing step).
Run in the build container only (the reference tree does not travel)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import polymesh as pm  # noqa: E402

if __name__ == "__main__":
    for tut in ("04unsteadyNS", "01POD", "03steadyNS"):
        out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mesh_%s.npz" % tut)
        pm.save_compact(out, "/root/reference/tutorials/CFD/" + tut)
        print(out, os.path.getsize(out))
