"""N>1 path on CPU: world_size-2 `gloo` run of the sample-slice partition. Each rank renders its slice (with the CPU
oracle standing in for the device — this test is about the host-side partition and the reduction, not the kernels),
all-reduces the float4 accumulation buffer, and the result must equal the single-rank frame."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import henjou_renderer_b200 as hj

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, spp, out_dir):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle_binding import Oracle
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    scene.set_sky((0.8, 0.8, 0.8))
    xf, inv = scene.transforms(1 / 24)
    b, e = hj.sample_slice(spp, rank, world)
    fp = hj.frame_params(scene, 40, 30, spp, frame=1, max_depth=6, sample_begin=b, sample_end=e)
    res = Oracle(scene, xforms=xf, inv_xforms=inv).render(fp, threads=2)
    acc = torch.from_numpy(res["accum"].copy())
    dist.all_reduce(acc, op=dist.ReduceOp.SUM)          # the one exchange step of the path
    if rank == 0:
        np.save(os.path.join(out_dir, "acc.npy"), acc.numpy())
    dist.destroy_process_group()


def test_slices_cover_exactly():
    for spp in (0, 1, 7, 64, 1024):
        for world in (1, 2, 3, 8):
            got = [hj.sample_slice(spp, r, world) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == spp
            assert all(got[i][1] == got[i + 1][0] for i in range(world - 1))
            assert max(e - b for b, e in got) - min(e - b for b, e in got) <= 1
    assert hj.weak_slice(64, 3, 8) == (512, 192, 256)
    with pytest.raises(ValueError):
        hj.sample_slice(4, 2, 2)


def test_two_rank_gloo_equals_single_rank(tmp_path):
    spp, world = 5, 2
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, spp, str(tmp_path)), nprocs=world, join=True)
    acc = np.load(os.path.join(str(tmp_path), "acc.npy"))
    from oracle_binding import Oracle
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    scene.set_sky((0.8, 0.8, 0.8))
    xf, inv = scene.transforms(1 / 24)
    fp = hj.frame_params(scene, 40, 30, spp, frame=1, max_depth=6)
    full = Oracle(scene, xforms=xf, inv_xforms=inv).render(fp, threads=2)["accum"]
    assert np.all(acc[..., 3] == spp)
    np.testing.assert_allclose(acc, full, rtol=1e-6, atol=1e-6)   # float summation order differs across the slice boundary
