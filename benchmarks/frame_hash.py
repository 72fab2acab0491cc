"""sha1 of two path-traced frames + final RNG states (pose B, 640x360, 16 spp, 4 bounces): equal hashes = bit-identical variants."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402

rt = load_package()
d = gpu_util.DeviceScene(rt, torch, 640, 360, "poseB", seed=1234, env=True)
for k in range(2):
    d.process(k, spp=16, bounces=4, write_last=1)
torch.cuda.synchronize()
h = hashlib.sha1()
h.update(d.surface.cpu().numpy().tobytes()); h.update(d.rng.cpu().numpy().tobytes())
print(h.hexdigest()[:16], d.rays())
