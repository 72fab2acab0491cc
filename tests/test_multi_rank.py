"""World-size-2 test of the multi-GPU partitioning logic on CPU (gloo): each rank renders ITS samples of the same frame with
the plain-C oracle (oracle/rt_port.c) on ITS cuRAND subsequences into a sum buffer, one reduce(sum) brings them to rank 0, and
the result must equal the serial sum of both ranks' buffers; rank 0's stream must be the single-GPU stream."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import gpu_util
import util

W, H, SPP_TOTAL, BOUNCES, SEED = 40, 30, 5, 4, 99


def _render_rank(rank, world):
    """Sum buffer of `rank` (numpy [H, W, 4]) computed with the oracle port."""
    from _loader import load_package
    import port_oracle as po
    rt = load_package()
    mg = __import__("cuda_raytracing_b200.multigpu", fromlist=["x"])
    scene, _ = util.bunny_scene(rt)
    a = scene.arrays()
    p, ax, ay = util.POSES["poseB"]
    scene.set_camera(p, ax, ay, W, H)
    cam, _ = scene.camera()
    ps = po.PortScene({k: util.raw(a[k]) for k in ("vertices", "faces", "spheres", "materials")}, util.raw(a["nodes"]), a["face_indices"],
                      cam.view(np.uint8).ravel().copy(), util.cubemap_texels())
    n_states = gpu_util.rng_state_count(W, H)
    spp = mg.partition_spp(SPP_TOTAL, world)[rank]
    st = po.rng_init(n_states, SEED, mg.rank_first_subsequence(rank, n_states))
    acc = np.zeros((H, W, 4), np.float32)
    for n in mg.launches_for(spp, 2):
        ps.render(st, acc, None, W, H, 0, spp=n, bounces=BOUNCES, blend_mode=1)
    return acc, mg


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, util.ROOT); sys.path.insert(0, os.path.join(util.ROOT, "oracle")); sys.path.insert(0, os.path.join(util.ROOT, "tests"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    acc, mg = _render_rank(rank, world)
    t = torch.from_numpy(acc.copy())
    mg.reduce_frame(t, dist, dst=0)
    if rank == 0:
        q.put((acc, t.numpy().copy(), mg.finalize_mean(t).numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


def test_partition_helpers(rt):
    mg = __import__("cuda_raytracing_b200.multigpu", fromlist=["x"])
    assert mg.partition_spp(1024, 8) == [128] * 8 and mg.partition_spp(5, 2) == [3, 2] and mg.partition_spp(3, 4) == [1, 1, 1, 0]
    assert sum(mg.partition_spp(1000, 7)) == 1000
    assert mg.launches_for(10, 4) == [4, 4, 2] and mg.launches_for(0, 4) == []
    assert mg.rank_first_subsequence(3, 2073600) == 3 * 2073600
    with pytest.raises(ValueError):
        mg.partition_spp(4, 0)


def test_two_rank_spp_split_and_reduce_gloo(rt):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    acc0, reduced, mean = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    # serial restatement of what the two ranks did
    s0, _ = _render_rank(0, 2)
    s1, _ = _render_rank(1, 2)
    assert np.array_equal(acc0, s0)
    assert np.array_equal(reduced, s0 + s1)
    assert (reduced[..., 3] == SPP_TOTAL).all()
    assert np.allclose(mean[..., :3], (s0 + s1)[..., :3] / SPP_TOTAL, rtol=1e-6) and (mean[..., 3] == 1).all()
    # rank 0 of a 2-rank job == the first samples of the 1-rank job (same subsequences, same stream)
    one, _ = _render_rank(0, 1)
    assert not np.array_equal(one, s0)  # 5 spp vs 3 spp
    assert np.isfinite(one).all()
