"""Generates tests/golden/mesh_04unsteadyNS.npz from the reference's own ASCII polyMesh
(/root/reference/tutorials/CFD/04unsteadyNS/constant/polyMesh: 11 644 cells, unstructured snappyHexMesh with a
cylinder and two `empty` patches). Run in the build container only (the reference tree does not travel)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import polymesh as pm  # noqa: E402

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mesh_04unsteadyNS.npz")
    pm.save_compact(out, "/root/reference/tutorials/CFD/04unsteadyNS")
    print(out, os.path.getsize(out))
