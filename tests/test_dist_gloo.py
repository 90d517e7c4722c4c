"""world_size = 2 over gloo on CPU: the host-side logic of the row-sharded path (partition layout,
cross-rank row gathering, all-reduce plumbing, host takeSample agreeing on every rank).  The CUDA
kernels themselves are covered by the -m gpu tests and tools/dist_check.py (torchrun, NCCL)."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
mp = pytest.importorskip("torch.multiprocessing")


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import hkm_b200
        from hkm_b200 import _runtime as RT, _sparkrand as SR
        from hkm_b200.ml._kmeans_init import PartitionLayout, _gather_global_rows, init_random
        rng = np.random.default_rng(0)
        full = rng.normal(size=(1000, 5))
        sizes_all = [600, 400]
        lo = sum(sizes_all[:rank]); X = torch.from_numpy(full[lo:lo + sizes_all[rank]].copy())
        sizes = RT.all_gather_object(int(X.shape[0]))
        assert sizes == sizes_all and RT.dist_info()[:2] == (rank, world)
        lay = PartitionLayout.build(sizes, rank, 8)
        assert lay.parts_per_rank == [4, 4] and lay.part_index0 == 4 * rank and lay.row_offset == lo
        assert lay.part_sizes == [150] * 4 + [100] * 4
        idx = [999, 0, 600, 599, 3]
        rows = _gather_global_rows(X, lay, idx)                       # same answer on every rank, in the requested order
        assert np.array_equal(rows, full[idx])
        C = init_random(X, lay, 7, seed=42)                           # host takeSample over the global layout + row gather
        want = SR.take_sample(lay.part_sizes, 7, SR.XORShiftRandom(42).next_int())
        assert np.array_equal(C, SR.distinct_rows(full[want]))
        t = torch.full((3,), float(rank + 1), dtype=torch.float64)
        RT.all_reduce_sum(t)
        assert t.tolist() == [3.0, 3.0, 3.0]
        q.put((rank, "ok", C.tolist()))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "fail: " + traceback.format_exc(), None))
    finally:
        dist.destroy_process_group()


def test_two_rank_host_logic():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res
    assert res[0][2] == res[1][2]                                      # both ranks derived identical initial centres
