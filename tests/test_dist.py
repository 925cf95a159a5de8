"""N > 1 path.  CPU (gloo, world_size 2): the rendezvous plumbing, tree ownership, the packed histogram exchange of the
rows-sharded fit (layout from the library, sums by gloo) and the rule that only rank 0 runs/prints the reference arm.  GPU (needs >= 2 B200s): a tree-sharded fit equals the oracle forest."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from euromillioner_b200.dist import broadcast_bytes, env_rank, trees_of_rank
    uid = broadcast_bytes(dist, bytes(range(128)) if rank == 0 else None)
    q.put((rank, uid == bytes(range(128)), env_rank(), trees_of_rank(7, rank, world)))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_rendezvous_and_tree_ownership():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, 29533, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0] == (0, True, (0, 0, 2), [0, 2, 4, 6])
    assert got[1] == (1, True, (1, 1, 2), [1, 3, 5])


def _gloo_packed_worker(rank, world, port, q):
    """The rows-sharded exchange of one group on the CPU: every rank holds partial node histograms for its rows, lays them
    out with the library's packed slot layout (light nodes as 16-bit pairs), the ranks sum the buffers as 32-bit integers
    (gloo all-reduce standing in for NCCL's reduce-scatter: a rank then looks only at its own share) and every owner unpacks
    its nodes.  The result must be the plain sum of the ranks' histograms."""
    import ctypes as C

    import numpy as np
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from euromillioner_b200 import _lib
    L = _lib.lib()
    n_nodes, words = 37, 6 * 4 * 5 * 2                      # slot: f_sub x bins x classes counters (an even number)
    rng = np.random.default_rng(5)                          # the same on both ranks: global weights and the split of each cell
    weight = np.where(rng.random(n_nodes) < 0.7, rng.integers(1, 65536, n_nodes), rng.integers(65536, 400000, n_nodes)).astype(np.uint32)
    total = np.zeros((n_nodes, words), np.int64)
    for i in range(n_nodes):                                # cells of a node sum to its weight (so no cell exceeds it)
        total[i] = rng.multinomial(int(weight[i]), np.full(words, 1.0 / words))
    share0 = (total * rng.random((n_nodes, words))).astype(np.int64)
    mine = share0 if rank == 0 else total - share0          # this rank's partial histograms
    slot, owner = np.zeros(n_nodes, np.int32), np.zeros(n_nodes, np.int32)
    share = L.emrf_host_packed_slot_layout(weight.ctypes.data, n_nodes, world, slot.ctypes.data, owner.ctypes.data)
    half = words // 2
    buf = np.zeros(share * world * half, np.int64)
    for i in range(n_nodes):
        at = int(slot[i]) * half
        if weight[i] <= 65535:
            buf[at:at + half] += mine[i, 0::2] | (mine[i, 1::2] << 16)      # two 16-bit counters per word
        else:
            buf[at:at + words] += mine[i]
    t = torch.from_numpy(buf)
    dist.all_reduce(t)                                       # exact integers; every word stays below 2^32
    assert int(t.max()) < 2 ** 32
    ok = True
    for i in np.where(owner == rank)[0]:
        at = int(slot[i]) * half
        if weight[i] <= 65535:
            w = t[at:at + half].numpy()
            got = np.empty(words, np.int64)
            got[0::2], got[1::2] = w & 0xffff, w >> 16
        else:
            got = t[at:at + words].numpy()
        ok &= bool(np.array_equal(got, total[i]))
    q.put((rank, ok, int(np.sum(owner == rank))))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_packed_histogram_exchange_is_exact():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_packed_worker, args=(r, 2, 29541, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0][1] and got[1][1]
    assert got[0][2] + got[1][2] == 37 and min(got[0][2], got[1][2]) >= 17     # the nodes are dealt evenly


def test_reference_arm_runs_on_rank0_only(monkeypatch):
    """bench.py --impl reference under N ranks: rank 0 prints the line, the others exit 0 silently."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_reference_arm_uses_all_cores_under_torchrun():
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm must still use every core it may run on
    (round 1: the reference arm ran single-threaded at N > 1 and was killed by the driver's time limit)."""
    code = ("import sys; sys.path.insert(0, %r)\n"
            "import bench\n"
            "from oracle import binding as ob\n"
            "ob.lib().orc_set_num_threads(bench.host_threads())\n"
            "print('THREADS', ob.lib().orc_num_threads(), bench.host_threads())\n") % ROOT
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, OMP_NUM_THREADS="1"), capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 0, r.stderr[-1000:]
    used, avail = [int(v) for v in r.stdout.split("THREADS")[1].split()[:2]]
    assert used == avail and (avail == 1 or used > 1)


@pytest.mark.gpu
def test_tree_and_row_sharded_fits_match_oracle_on_2_gpus():
    import ctypes
    n = ctypes.c_int(0)
    cuda = ctypes.CDLL("libcuda.so.1")
    cuda.cuInit(0)
    cuda.cuDeviceGetCount(ctypes.byref(n))
    if n.value < 2:
        pytest.skip("needs 2 GPUs")
    # NCCL reduce-scatter path (light nodes packed to 16 bits), the same all 32-bit, then the fused NVLink-push path
    # and once with every rank exporting only its share of the trees (the default above 4 ranks)
    # (the second run also tiles the groups on the host from sizes read back, instead of on the device)
    for push, no_pack, shard in (("0", "0", "0"), ("0", "1", "1"), ("1", "0", "0")):
        env = dict(os.environ, EMRF_PEER_PUSH=push, EMRF_NO_RS_PACK=no_pack, EMRF_EXPORT_SHARD=shard, EMRF_RS_HOST_SIZES=no_pack)
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                            "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tests", "dist_fit_check.py")],
                           capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert "DIST_FIT_OK world=2" in r.stdout
        assert r.stdout.count("DIST_ROWS_OK world=2") == 3
