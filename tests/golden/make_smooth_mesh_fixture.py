"""Writes tests/golden/bunny_smooth_mesh.npz = the bunny after the product's OBJ ingest (join identical vertices + smooth normals,
host/MeshLoader.cpp): positions, normals, indices. bench.py's `--impl reference` arm assembles the reference scene from it with
the reference's own code (oracle/_ref: GLM scene assembly + BVH.cpp) WITHOUT loading libb200rt.so.
    python tests/golden/make_smooth_mesh_fixture.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import util  # noqa: E402
from _loader import load_package  # noqa: E402

rt = load_package()
pos, nrm, idx, _ = rt.Mesh(util.assets()["obj"]).arrays()
np.savez_compressed(os.path.join(HERE, "bunny_smooth_mesh.npz"), positions=pos, normals=nrm, indices=idx)
print(pos.shape, nrm.shape, idx.shape)
