"""CPU, world_size 2, gloo: the multi-GPU host logic (round-robin sharding of frames, padded-record all_gather,
re-ordering) that bench.py and a multi-GPU caller use.  The per-rank 'extraction' here is the CPU oracle, so the test
checks only the plumbing: the gathered, re-ordered results must equal a single-process run frame by frame."""
import os, socket, sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, nframes, cap, q):
    sys.path.insert(0, HERE); sys.path.insert(0, ROOT)
    import importlib
    import synth, oracle_lib as O
    sh = importlib.import_module("orb-slam2-noted_b200.sharding")
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = sh.shard_indices(nframes, rank, world)
    assert len(mine) == sh.local_count(nframes, rank, world)
    B = (nframes + world - 1) // world
    kps = torch.zeros((B, cap, 28), dtype=torch.uint8); desc = torch.zeros((B, cap, 32), dtype=torch.uint8)
    n = torch.zeros(B, dtype=torch.int32)
    p = O.params(300)
    for slot, f in enumerate(mine):
        k, d = O.extract(p, synth.frame(500 + f, 160, 120))
        n[slot] = len(k)
        kps[slot, :len(k)] = torch.from_numpy(k.view(np.uint8).reshape(-1, 28).copy())
        desc[slot, :len(k)] = torch.from_numpy(d)
    gk, gd, gn = sh.gather_frames(kps, desc, n, world)
    if rank == 0:
        src = sh.global_order((gk, gd, gn), nframes, world)
        q.put((gk[src].numpy(), gd[src].numpy(), gn[src].numpy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("nframes", [5, 6])
def test_round_robin_gather_world2(nframes):
    sys.path.insert(0, HERE)
    import synth, oracle_lib as O
    O.lib()
    world, cap = 2, 352
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nframes, cap, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    gk, gd, gn = q.get(timeout=120)
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    p = O.params(300)
    for f in range(nframes):
        k, d = O.extract(p, synth.frame(500 + f, 160, 120))
        assert gn[f] == len(k)
        assert np.array_equal(gk[f, :len(k)].reshape(-1), k.view(np.uint8).reshape(-1))
        assert np.array_equal(gd[f, :len(k)], d)


def test_shard_helpers():
    import importlib
    sys.path.insert(0, ROOT)
    sh = importlib.import_module("orb-slam2-noted_b200.sharding")
    for n in (0, 1, 7, 8, 512):
        for w in (1, 2, 4, 8):
            allidx = sorted(i for r in range(w) for i in sh.shard_indices(n, r, w))
            assert allidx == list(range(n))
            assert [sh.local_count(n, r, w) for r in range(w)] == [len(sh.shard_indices(n, r, w)) for r in range(w)]
