"""Counts behind tests/test_gpu_parity.py::test_rays_aimed_at_edges_and_vertices: product traversal vs the reference-order
kernel on rays aimed exactly at bunny vertices / edges (order-dependent last-ulp cases). One JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402
from test_gpu_parity import edge_aimed_rays  # noqa: E402

rt = load_package()
dev = gpu_util.DeviceScene(rt, torch, 320, 240, "default", seed=1234)
rays = edge_aimed_rays(dev.scene.arrays(), n=200000)
i0, t0 = dev.trace_rays(rays, True, 0)
i1, t1 = dev.trace_rays(rays, True, 1)
ulp = np.abs(t0.view(np.int32).astype(np.int64) - t1.view(np.int32).astype(np.int64))
print(json.dumps({"rays": len(rays), "hit": int((i0 != 0xFFFFFFFF).sum()), "t_differs": int((ulp > 0).sum()),
                  "product_closer": int((t0 < t1).sum()), "reference_order_closer": int((t1 < t0).sum()),
                  "same_t_other_id": int(((ulp == 0) & (i0 != i1)).sum()), "max_ulp": int(ulp.max()),
                  "max_rel": float((np.abs(t0 - t1) / np.maximum(t1, 1e-30)).max()),
                  "accel": {k: int(v) for k, v in dev.scene.accel_info().items()}}))
