"""GPU: the final gather behind the C ABI (orbx_comm_*, orbx_gather_frames, orbx_gather_matches).
One rank: the device-side compaction and the record layout against numpy.  Two ranks (needs 2 GPUs, run under torchrun from
this test): what arrives at the root over NCCL equals what the root extracts itself from the other rank's frames."""
import importlib, os, subprocess, sys
import numpy as np
import pytest
import synth
from gpu_util import pkg

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def test_gather_single_rank_compacts_frames_and_match_lists():
    import torch
    P = pkg()
    sh = importlib.import_module("orb-slam2-noted_b200.sharding")
    dev = torch.device("cuda", 0)
    B, w, h, nf = 6, 320, 240, 500
    ex = P.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=B)
    cap = ex.max_keypoints()
    kinds = ["texture", "texture", "black", "gradient", "texture", "noise"]          # a frame without keypoints in the middle
    imgs = torch.from_numpy(np.stack([synth.frame(60 + i, w, h, kinds[i]) for i in range(B)])).to(dev)
    kps = torch.full((B, cap, 28), 0xAB, dtype=torch.uint8, device=dev); desc = torch.full((B, cap, 32), 0xCD, dtype=torch.uint8, device=dev)
    n = torch.zeros(B, dtype=torch.int32, device=dev)
    ex.extract_batch_device(imgs, kps, desc, n)
    comm = sh.Comm(ex)                                                              # one rank: no NCCL involved
    out = (torch.zeros(B, dtype=torch.int32, device=dev), torch.zeros((B * cap, 28), dtype=torch.uint8, device=dev),
           torch.zeros((B * cap, 32), dtype=torch.uint8, device=dev))
    for _ in range(2):                                                              # twice: the staging is reused
        total = comm.gather_frames(kps, desc, n, out)
        comm.sync()
        hn = n.cpu().numpy()
        assert total == int(hn.sum()) and hn[2] == 0 and np.array_equal(out[0].cpu().numpy(), hn)
        wk = np.concatenate([kps[f, :hn[f]].cpu().numpy() for f in range(B)]); wd = np.concatenate([desc[f, :hn[f]].cpu().numpy() for f in range(B)])
        assert np.array_equal(out[1][:total].cpu().numpy(), wk) and np.array_equal(out[2][:total].cpu().numpy(), wd)
    small = (out[0], out[1][:10], out[2][:10])
    with pytest.raises(RuntimeError, match="capacity"):
        comm.gather_frames(kps, desc, n, small)
    # match lists: (row, match[row]) records of the rows in use, ascending row, pair after pair
    rng = np.random.default_rng(3)
    pairs, mcap = 5, 3000
    rows = np.array([3000, 0, 1777, 256, 257], np.int32)
    match = np.where(rng.random((pairs, mcap)) < 0.3, rng.integers(0, 5000, (pairs, mcap)), -1).astype(np.int32)
    dm, dr = torch.from_numpy(match).to(dev), torch.from_numpy(rows).to(dev)
    mo = (torch.zeros(pairs, dtype=torch.int32, device=dev), torch.zeros((pairs * mcap, 2), dtype=torch.int32, device=dev))
    total = comm.gather_matches(dm, dr, mo)
    comm.sync()
    want = [np.stack([np.flatnonzero(match[p, :rows[p]] >= 0), match[p, :rows[p]][match[p, :rows[p]] >= 0]], 1) for p in range(pairs)]
    assert np.array_equal(mo[0].cpu().numpy(), [len(x) for x in want]) and total == sum(len(x) for x in want)
    assert np.array_equal(mo[1][:total].cpu().numpy(), np.concatenate(want))
    # dense extraction (ORBX_OUT_DENSE): the describe stage writes the records where the compaction would put them
    kd = torch.full((B, cap, 28), 0xAB, dtype=torch.uint8, device=dev); dd = torch.full((B, cap, 32), 0xCD, dtype=torch.uint8, device=dev)
    nd = torch.zeros(B, dtype=torch.int32, device=dev)
    ex.extract_batch_device(imgs, kd, dd, nd, dense=True)
    ex.sync()
    total = int(n.sum())
    assert torch.equal(nd, n) and torch.equal(kd.view(-1, 28)[:total], out[1][:total]) and torch.equal(dd.view(-1, 32)[:total], out[2][:total])
    assert bool((kd.view(-1, 28)[total:] == 0xAB).all()) and bool((dd.view(-1, 32)[total:] == 0xCD).all())      # nothing behind the records
    out_d = (torch.zeros_like(out[0]), torch.zeros_like(out[1]), torch.zeros_like(out[2]))
    assert comm.gather_frames(kd, dd, nd, out_d, dense=True) == total
    comm.sync()
    assert torch.equal(out_d[0], out[0]) and torch.equal(out_d[1][:total], out[1][:total]) and torch.equal(out_d[2][:total], out[2][:total])
    with pytest.raises(Exception):                                                  # dense output needs device images and device results
        ex.extract_batch_device(imgs.cpu().pin_memory(), kd, dd, nd, in_device=False, dense=True)
    # a peer window on one rank is just memory owned by the communicator: gathers into it give the same records
    win = comm.window(sh.Comm.carve_bytes(((B * cap, 28), torch.uint8), ((B * cap, 32), torch.uint8)))
    assert win is not None and win.is_cuda and win.data_ptr() % 256 == 0
    wk2, wd2 = sh.Comm.carve(win, ((B * cap, 28), torch.uint8), ((B * cap, 32), torch.uint8))
    total = comm.gather_frames(kps, desc, n, (out[0], wk2, wd2))
    comm.sync()
    assert torch.equal(wk2[:total], out[1][:total]) and torch.equal(wd2[:total], out[2][:total])
    with pytest.raises(RuntimeError, match="window"):
        comm.window(1024)                                                           # one per communicator
    comm.close()


WORKER = r'''
import importlib, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path[:0] = [os.environ["ORBX_ROOT"], os.path.join(os.environ["ORBX_ROOT"], "tests")]
import synth
P = importlib.import_module("orb-slam2-noted_b200")
sh = importlib.import_module("orb-slam2-noted_b200.sharding")
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("gloo")                     # only to hand the NCCL id round: the data path is orbx's own communicator
B, w, h, nf = 5, 320, 240, 500
ex = P.ORBextractor(nf, 1.2, 8, 20, 7, device=rank, max_w=w, max_h=h, max_batch=B * world)
cap = ex.max_keypoints()
frames = lambda r: np.stack([synth.frame(700 + f, w, h) for f in sh.shard_indices(B * world, r, world)])
def extract(imgs):
    k = torch.zeros((len(imgs), cap, 28), dtype=torch.uint8, device=dev); d = torch.zeros((len(imgs), cap, 32), dtype=torch.uint8, device=dev)
    n = torch.zeros(len(imgs), dtype=torch.int32, device=dev)
    ex.extract_batch_device(torch.from_numpy(imgs).to(dev), k, d, n)
    return k, d, n
k, d, n = extract(frames(rank))
comm = sh.Comm.from_torch(ex)
rows_out = world * B * cap if rank == 0 else 1       # only the root's buffers count (and must be judged alike on every rank)
out = (torch.zeros(world * B, dtype=torch.int32, device=dev), torch.zeros((rows_out, 28), dtype=torch.uint8, device=dev),
       torch.zeros((rows_out, 32), dtype=torch.uint8, device=dev))
try:                                                  # a root buffer that is too small: EVERY rank must get the error, nobody may hang
    comm.gather_frames(k, d, n, (out[0], out[1][:1] if rank == 0 else out[1], out[2][:1] if rank == 0 else out[2]), root=0)
    raise SystemExit("expected a capacity error")
except RuntimeError as e:
    assert "capacity" in str(e)
ticket = comm.mark()
ex.extract_batch_device(torch.from_numpy(frames(rank)).to(dev), *[torch.zeros_like(t) for t in (k, d, n)])   # later work on the context: must not delay or disturb the gather
total = comm.gather_frames(k, d, n, out, root=0, ticket=ticket)
match = torch.full((3, 64), -1, dtype=torch.int32, device=dev); match[:, rank::7] = rank + 100
mo = (torch.zeros(world * 3, dtype=torch.int32, device=dev), torch.zeros((world * 3 * 64 if rank == 0 else 1, 2), dtype=torch.int32, device=dev))
mt = comm.gather_matches(match, torch.tensor([64, 10, 0], dtype=torch.int32, device=dev), mo, root=0)
comm.sync()
if rank == 0:
    wn, wk, wd = [], [], []
    for r in range(world):
        kk, dd, nn = extract(frames(r)); ex.sync(); hn = nn.cpu().numpy(); wn.append(hn)
        wk += [kk[f, :hn[f]].cpu().numpy() for f in range(B)]; wd += [dd[f, :hn[f]].cpu().numpy() for f in range(B)]
    assert np.array_equal(out[0].cpu().numpy(), np.concatenate(wn)) and total == int(np.concatenate(wn).sum()) and total > 1000 * world
    assert np.array_equal(out[1][:total].cpu().numpy(), np.concatenate(wk)) and np.array_equal(out[2][:total].cpu().numpy(), np.concatenate(wd))
    cnt = mo[0].cpu().numpy().reshape(world, 3)
    for r in range(world):
        assert list(cnt[r]) == [len(range(r, 64, 7)), len(range(r, 10, 7)), 0]
    rec = mo[1][:mt].cpu().numpy()
    assert mt == int(cnt.sum()) and np.array_equal(rec[:cnt[0, 0], 0], np.arange(0, 64, 7)) and set(rec[:, 1]) == set(range(100, 100 + world))
    print("GATHER_OK", world, total, mt)
# a rank with nothing to send (every count zero): nothing is staged, sent or received for it, the others are unaffected
nz = n if rank == 0 else torch.zeros_like(n)
outz = tuple(torch.zeros_like(t) for t in out)
tz = comm.gather_frames(k, d, nz, outz, root=0)
comm.sync()
if rank == 0:
    mine = int(n.sum())
    assert tz == mine and torch.equal(outz[0][:B], n) and int(outz[0][B:].sum()) == 0
    assert torch.equal(outz[1][:mine], out[1][:mine]) and torch.equal(outz[2][:mine], out[2][:mine])
else:
    assert tz > 1000                                     # every rank learns the same total (rank 0's records)
# dense records (ORBX_OUT_DENSE) need no compaction: sent from where they lie (here still through ncclSend/ncclRecv)
kd, dd, nd = torch.zeros_like(k), torch.zeros_like(d), torch.zeros_like(n)
ex.extract_batch_device(torch.from_numpy(frames(rank)).to(dev), kd, dd, nd, dense=True)
ex.sync()
out3 = tuple(torch.zeros_like(t) for t in out)
assert comm.gather_frames(kd, dd, nd, out3, root=0, dense=True) == total
comm.sync()
if rank == 0:
    assert torch.equal(out3[0], out[0]) and torch.equal(out3[1][:total], out[1][:total]) and torch.equal(out3[2][:total], out[2][:total])
# the same gathers through the root's peer window: the other rank writes into rank 0's memory
assert comm.stats() == (0, 4)
shapes = (((rows_out, 28), torch.uint8), ((rows_out, 32), torch.uint8), ((world * 3 * 64, 2), torch.int32))
win = comm.window(sh.Comm.carve_bytes(*shapes) if rank == 0 else 0)
have = torch.tensor([0 if win is None else 1]); dist.broadcast(have, src=0)
if not int(have):
    if rank == 0:
        print("WINDOW_UNSUPPORTED", comm._L.orbx_comm_last_error(comm._h).decode())
else:
    if rank == 0:
        win.fill_(0xEE)
        k2, d2, p2 = sh.Comm.carve(win, *shapes)
        out2, mo2 = (torch.zeros_like(out[0]), k2, d2), (torch.zeros_like(mo[0]), p2)
        torch.cuda.synchronize()
    else:
        out2, mo2 = out, mo
    dist.barrier()
    for rep in range(2):                                 # padded records through the compaction, then dense records as they lie
        t2 = comm.gather_frames(kd, dd, nd, out2, root=0, dense=True) if rep else comm.gather_frames(k, d, n, out2, root=0)
        m2 = comm.gather_matches(match, torch.tensor([64, 10, 0], dtype=torch.int32, device=dev), mo2, root=0)
        comm.sync()
        assert (t2, m2) == (total, mt)
        if rank == 0:
            assert torch.equal(out2[0], out[0]) and torch.equal(out2[1][:total], out[1][:total]) and torch.equal(out2[2][:total], out[2][:total])
            assert torch.equal(mo2[0], mo[0]) and torch.equal(mo2[1][:mt], mo[1][:mt])
            assert bool((out2[1][total:total + 4] == 0xEE).all())               # nothing written past the records
            k2[:total].zero_(); torch.cuda.synchronize()
        dist.barrier()                                   # the root has looked before the next round overwrites
    assert comm.stats() == (4, 4), comm.stats()
    if rank == 0:
        print("WINDOW_OK", world)
dist.barrier()
comm.close()
dist.barrier()
dist.destroy_process_group()
'''


@pytest.mark.parametrize("window_form", ["dma", "sm_stores"])
def test_gather_two_ranks_over_nccl(tmp_path, window_form):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, ORBX_ROOT=ROOT, ORBX_GATHER_COPY="1" if window_form == "dma" else "0")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29617", str(script)], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and "GATHER_OK 2" in r.stdout, (r.stdout[-2000:], r.stderr[-3000:])
    assert "WINDOW_OK 2" in r.stdout or "WINDOW_UNSUPPORTED" in r.stdout, (r.stdout[-2000:], r.stderr[-3000:])
    if "WINDOW_UNSUPPORTED" in r.stdout:
        pytest.skip("peer windows are not available on this box: " + r.stdout[-300:])
