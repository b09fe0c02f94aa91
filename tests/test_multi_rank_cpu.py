"""Host-side logic of the multi-GPU path on CPU: tile ownership covers the frame
exactly once for every rank count, and the rank-0 frame handle travels to the other ranks
(world_size 2, gloo) the way bench.py ships the CUDA-IPC handle."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest


def owner(tx, ty, nranks):
    return (tx + 3 * ty) % nranks      # k_frame: rank tile (Tx,Ty) of 32x32 pixels


@pytest.mark.parametrize("nranks", [1, 2, 4, 8])
@pytest.mark.parametrize("wh", [(3840, 2160), (960, 540), (37, 23)])
def test_tile_ownership_is_a_balanced_partition(nranks, wh):
    w, h = wh
    ntx, nty = (w + 31) // 32, (h + 31) // 32
    own = np.array([[owner(tx, ty, nranks) for tx in range(ntx)] for ty in range(nty)])
    counts = np.bincount(own.ravel(), minlength=nranks)
    assert counts.sum() == ntx * nty
    if ntx * nty >= 8 * nranks:
        assert counts.max() - counts.min() <= max(2, 0.02 * counts.mean()), counts
    if nranks > 1 and ntx > 1:
        assert (own[:, 1:] != own[:, :-1]).all()      # neighbours in a row never share a rank
    if nranks > 3 and nty > 1:
        assert (own[1:, :] != own[:-1, :]).all()      # the 3-tile skew separates rows too


def test_handle_broadcast_world_size_2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent("""
        import os, sys
        import torch, torch.distributed as dist
        dist.init_process_group("gloo")
        rank = dist.get_rank()
        handle = torch.zeros(64, dtype=torch.uint8)
        if rank == 0:
            handle = torch.arange(64, dtype=torch.uint8) * 3 + 1      # stands for cudaIpcMemHandle_t bytes
        dist.broadcast(handle, 0)
        assert handle.tolist() == [(3 * i + 1) % 256 for i in range(64)]
        # max-over-ranks timing reduction and ray-count sum used by bench.py
        t = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == 11.0
        c = torch.tensor([5.0, 7.0]) * (rank + 1)
        dist.all_reduce(c)
        assert c.tolist() == [15.0, 21.0]
        dist.barrier()
        dist.destroy_process_group()
        print("rank", rank, "ok")
    """))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)], capture_output=True,
                       text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


def test_upload_group_barrier_and_failure_release(tmp_path):
    """rtb_render_multi's device threads meet at UploadGroup::barrier (csrc/scene_store.h) while device 0 uploads the
    scene for all of them; a thread that fails must release the others.  Host-only: compiled with g++ against the CUDA
    headers, no device call."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime_api.h")):
        pytest.skip("CUDA headers not found")
    exe = str(tmp_path / "upload_group_host")
    subprocess.run(["g++", "-std=c++17", "-O1", "-pthread", "-I" + cuda_inc, "-I" + os.path.join(root, "include"),
                    "-I" + os.path.join(root, "raytracing-gpu_b200", "csrc"), "-o", exe,
                    os.path.join(root, "tests", "native", "upload_group_host.cpp")], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout + r.stderr
