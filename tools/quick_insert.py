#!/usr/bin/env python
"""Times hg_grid_insert_tsdf / hg_grid_insert_occupancy per scan against the CPU oracle (development tool)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import hectorgrapher_b200 as hg  # noqa: E402
from hectorgrapher_b200 import synth  # noqa: E402
from oracle import pyoracle  # noqa: E402

pyoracle.build()
sph = synth.spheres()
clouds = []
for k in range(8):
    origin, yaw = synth.corkscrew_pose(k, 8)
    c = synth.scan(origin, yaw, 64, 1600, sph, np.random.RandomState(k))
    q = synth.quat_from_yaw(yaw)
    clouds.append((origin.astype(np.float32), (synth.quat_rotate(q, c.astype(np.float64)) + origin).astype(np.float32)))
empty = np.zeros((0, 3), np.int32)
for free in (0, 3):
    opts = dict(epsilon=1.0, min_range=0.4, max_range=60.0, num_free_space_voxels=free)
    g = hg.Grid.tsdf(0.1, 2.5, 1000.0, empty, [], [])
    o = pyoracle.Grid.tsdf(0.1, 2.5, 1000.0)
    for k, (origin, pts) in enumerate(clouds):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = g.insert_tsdf(origin, pts, **opts)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        t2 = t1
        if k < 3:
            o.insert_tsdf(origin, pts, **opts)
            t2 = time.perf_counter()
        print("tsdf free=%d scan %d: %d returns, %d updates, gpu %.3f ms, cpu %.1f ms, bricks %d" %
              (free, k, len(pts), n, 1e3 * (t1 - t0), 1e3 * (t2 - t1), g.num_bricks), flush=True)
g = hg.Grid.prob(0.1, empty, [])
o = pyoracle.Grid.prob(0.1)
for k, (origin, pts) in enumerate(clouds):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g.insert_occupancy(origin, pts, 0.55, 0.49, 2)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    t2 = t1
    if k < 3:
        o.insert_occupancy(origin, pts, 0.55, 0.49, 2)
        t2 = time.perf_counter()
    print("occupancy scan %d: %d returns, gpu %.3f ms, cpu %.1f ms, bricks %d" %
          (k, len(pts), 1e3 * (t1 - t0), 1e3 * (t2 - t1), g.num_bricks), flush=True)
