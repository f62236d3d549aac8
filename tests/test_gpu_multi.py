"""The multi-GPU reduction path on hardware: two processes, one B200 each, render the two sample slices of one frame,
hj_comm_init + hj_comm_reduce (ncclAllReduce of the float4 accumulation buffer in place, then finalize), and the result must be the
single-GPU frame: accumulation within float summation order (1e-6 relative, w == spp), colour and RGBA8 with it. Skipped on a box
with one GPU; the driver's scaling lane (2/4/8 GPUs) runs it."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _worker(rank, world, tmp, spp):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import time
    import henjou_renderer_b200 as hj
    from henjou_renderer_b200 import abi
    idfile = os.path.join(tmp, "nccl_id.bin")
    if rank == 0:
        uid = hj.Renderer.comm_unique_id()
        with open(idfile + ".tmp", "wb") as f:
            f.write(uid)
        os.replace(idfile + ".tmp", idfile)
    else:
        for _ in range(600):
            if os.path.exists(idfile):
                break
            time.sleep(0.1)
        uid = open(idfile, "rb").read()
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    scene.set_sky((0.8, 0.8, 0.8))
    xf, inv = scene.transforms(1 / 24)
    r = hj.Renderer(rank)
    r.upload(scene); r.build_accel(xf, inv)
    full = None
    if rank == 0:                                            # the single-GPU frame [0, spp)
        r.render(hj.frame_params(scene, 96, 64, spp, frame=1, max_depth=8, tonemap=abi.TONEMAP_ACES))
        full = [r.readback(a).copy() for a in (abi.AOV_ACCUM, abi.AOV_COLOR, abi.AOV_RGBA8)]
    r.comm_init(uid, rank, world)
    # hj_comm_scene_upload: the geometry crosses PCIe on rank 0 only; rank 1 offers poisoned arrays that must never be read
    import copy
    shared = copy.copy(scene)
    if rank != 0:
        shared.vertices = np.full_like(scene.vertices, np.nan); shared.normals = np.full_like(scene.normals, np.nan)
        shared.texcoords = np.full_like(scene.texcoords, np.nan); shared.indices = np.zeros_like(scene.indices); shared.material_ids = np.zeros_like(scene.material_ids)
    r.comm_upload(shared, 0); r.build_accel(xf, inv)
    b, e = hj.sample_slice(spp, rank, world)
    r.render(hj.frame_params(scene, 96, 64, spp, frame=1, max_depth=8, tonemap=abi.TONEMAP_ACES, sample_begin=b, sample_end=e), sync=False)
    r.comm_reduce()
    got = [r.readback(a).copy() for a in (abi.AOV_ACCUM, abi.AOV_COLOR, abi.AOV_RGBA8)]
    r.comm_reduce()                                          # a second call must not sum again
    again = r.readback(abi.AOV_ACCUM).copy()
    np.save(os.path.join(tmp, "acc_%d.npy" % rank), got[0])
    ok = np.array_equal(again.view(np.uint32), got[0].view(np.uint32))
    if rank == 0:
        ok = ok and np.all(got[0][..., 3] == spp)
        ok = ok and np.allclose(got[0], full[0], rtol=1e-6, atol=1e-6)           # float summation order differs at the slice boundary
        ok = ok and np.allclose(got[1], full[1], rtol=1e-5, atol=1e-6)
        ok = ok and (np.abs(got[2].astype(int) - full[2].astype(int)) <= 1).all() and (got[2] != full[2]).mean() < 1e-3
    open(os.path.join(tmp, "ok_%d" % rank), "w").write("1" if ok else "0")
    r.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs")
def test_two_gpu_reduce_equals_single_gpu_frame(tmp_path):
    import torch.multiprocessing as mp
    world, spp = 2, 6
    mp.spawn(_worker, args=(world, str(tmp_path), spp), nprocs=world, join=True)
    assert all(open(os.path.join(str(tmp_path), "ok_%d" % k)).read() == "1" for k in range(world))
    a0, a1 = (np.load(os.path.join(str(tmp_path), "acc_%d.npy" % k)) for k in range(world))
    assert np.array_equal(a0.view(np.uint32), a1.view(np.uint32))               # every rank holds the same reduced frame
