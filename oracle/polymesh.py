"""TEST INFRASTRUCTURE ONLY -- ASCII polyMesh reader and fixtures (the geometry and the synthetic mesh generators live in
bench_inputs/meshgen.py, a neutral input generator shared with the benchmark, and are re-exported here).

The product never imports this. In production OpenFOAM itself hands owner/neighbour/Sf/weights/... to
the C ABI (ithaca_b200_mesh_upload); here they are restated so that tests and the benchmark can fabricate
the same arrays without OpenFOAM (not installable in this image).

Parity status: UNPINNED at the OpenFOAM boundary. OpenFOAM (pinned by the reference to v2412,
/root/reference/README.md:19, .compileOF2412.sh:2-4) is a third-party dependency that is absent from
/root/reference; the formulas below restate its published algorithms:
  face centres / area vectors : primitiveMeshTools::faceCentresAndAreas (triangle fan about the vertex average)
  cell centres / volumes      : primitiveMeshTools::cellCentresAndVols (pyramids about the face-centre average)
  linear weights              : surfaceInterpolation::makeWeights      w = |Sf.(Cn-Cf)| / (|Sf.(Cf-Co)| + |Sf.(Cn-Cf)|)
  nonOrthDeltaCoeffs          : surfaceInterpolation::makeNonOrthDeltaCoeffs   1/max(nf.d, 0.05|d|), d = Cn-Co
  nonOrthCorrectionVectors    : surfaceInterpolation::makeNonOrthCorrectionVectors   nf - d*nonOrthDeltaCoeffs
  boundary deltaCoeffs        : fvPatch::delta() = nf*(nf.(Cf-Cc)) for non-coupled patches, coefficient 1/|delta|
Known answers used to check them: tests/test_fvc_oracle.py (uniform cavity V = 1e-8, closed cells sum Sf = 0,
Gauss theorem, exact gradients of linear fields on orthogonal meshes).
"""
import os
import re

import numpy as np


# ------------------------------------------------------------------------------------------------
# ASCII polyMesh reader (points, faces, owner, neighbour, boundary)
# ------------------------------------------------------------------------------------------------
def _body(path):
    txt = open(path).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//.*", "", txt)
    i = txt.find("}")  # end of FoamFile header
    return txt[i + 1:]


def _list_start(body):
    m = re.search(r"(\d+)\s*\(", body)
    return int(m.group(1)), m.end()


def read_polymesh(d):
    """Returns dict(points (P,3), faces list-of-arrays, owner, neighbour, patches [(name,type,nFaces,startFace)])."""
    b = _body(os.path.join(d, "points"))
    n, s = _list_start(b)
    nums = np.array(re.findall(r"[-+0-9.eE]+", b[s:]), dtype=np.float64)[:3 * n]
    points = nums.reshape(n, 3)
    b = _body(os.path.join(d, "faces"))
    n, s = _list_start(b)
    faces = [np.array(g.split(), dtype=np.int64) for g in re.findall(r"\d+\(([^)]*)\)", b[s:])][:n]
    out = {}
    for name in ("owner", "neighbour"):
        b = _body(os.path.join(d, name))
        n, s = _list_start(b)
        out[name] = np.array(b[s:].replace(")", " ").split()[:n], dtype=np.int64)
    b = _body(os.path.join(d, "boundary"))
    patches = []
    for m in re.finditer(r"(\w+)\s*\{([^}]*)\}", b):
        kv = dict(re.findall(r"(\w+)\s+([^;]+);", m.group(2)))
        if "nFaces" in kv:
            patches.append((m.group(1), kv.get("type", "patch").strip(), int(kv["nFaces"]), int(kv["startFace"])))
    return dict(points=points, faces=faces, owner=out["owner"], neighbour=out["neighbour"], patches=patches)


from bench_inputs.meshgen import FvMesh, cell_centres_volumes, face_centres_areas, faces_to_padded, synthetic_hex  # noqa: E402,F401


def load_reference_mesh(case_dir):
    pm = read_polymesh(os.path.join(case_dir, "constant", "polyMesh"))
    F, nv = faces_to_padded(pm["faces"])
    return FvMesh(pm["points"], F, nv, pm["owner"], pm["neighbour"], pm["patches"])


def save_compact(path, case_dir):
    """polyMesh -> compressed npz fixture (tests/golden/make_mesh_fixture.py)."""
    pm = read_polymesh(os.path.join(case_dir, "constant", "polyMesh"))
    F, nv = faces_to_padded(pm["faces"])
    np.savez_compressed(path, points=pm["points"], F=F.astype(np.int32), nv=nv.astype(np.int8),
                        owner=pm["owner"].astype(np.int32), neighbour=pm["neighbour"].astype(np.int32),
                        patch_names=np.array([p[0] for p in pm["patches"]]), patch_types=np.array([p[1] for p in pm["patches"]]),
                        patch_n=np.array([p[2] for p in pm["patches"]]), patch_start=np.array([p[3] for p in pm["patches"]]))


def load_compact(path):
    z = np.load(path)
    patches = [(str(a), str(b), int(c), int(d)) for a, b, c, d in zip(z["patch_names"], z["patch_types"], z["patch_n"], z["patch_start"])]
    return FvMesh(z["points"], z["F"].astype(np.int64), z["nv"].astype(np.int64), z["owner"].astype(np.int64),
                  z["neighbour"].astype(np.int64), patches)
