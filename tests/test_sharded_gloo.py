"""The N > 1 data path on CPU: two processes (gloo), each owning a contiguous document shard,
one all-reduce (SUM) of the packed statistics per EM iteration, every rank then forming the same
model - the protocol bench.py runs over NCCL and `lda est <#GPUs>` runs over peer memory
(reference: MPI sharding + Allreduce, lda-estimate.c:201-252).  The per-shard E-step here is
the oracle's (no GPU in this test); what is under test is the host-side protocol: shard
boundaries, packed layout [class_word | likelihood | alpha_ss | num_docs], identical models on
all ranks, and agreement with the single-process run."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"


def _worker(rank, world, port, case, q):
    sys.path.insert(0, str(ROOT / "oracle"))
    import torch
    import torch.distributed as dist
    import oracle_py
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    d = GOLD / case
    st = oracle_py.read_settings(d / "settings.txt")
    off, w, cn, tot, V = oracle_py.read_dat(d / "c.dat")
    alpha, lb = oracle_py.load_model(d / "000")
    K, D = lb.shape[0], len(off) - 1
    # contiguous shards balanced by entries (+1 per document), as split_by_nnz in csrc/est.cu
    want = (off[-1] + D) * np.arange(world + 1) / world
    bounds = np.searchsorted(off + np.arange(D + 1), want).clip(0, D)
    bounds[0], bounds[-1] = 0, D
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    soff = (off[lo:hi + 1] - off[lo]).astype(np.int64)
    sw, sc = w[off[lo]:off[hi]].copy(), cn[off[lo]:off[hi]].copy()
    liks = []
    L = oracle_py.lib()
    for _ in range(3):
        r = oracle_py.e_step(soff, sw, sc, V, alpha, lb, st.var_max_iter, st.var_converged)
        packed = np.concatenate([r["class_word"].T.ravel(), [r["lik"], r["alpha_ss"], float(r["num_docs"]), 0.0]])
        t = torch.from_numpy(packed)
        dist.all_reduce(t)                                   # the one exchange step per EM iteration
        cw = packed[:V * K].reshape(V, K).T.copy()
        lik, alpha_ss, ndocs = packed[V * K], packed[V * K + 1], int(round(packed[V * K + 2]))
        ct = cw.sum(axis=1)
        lb = np.zeros((K, V))
        L.oracle_lda_mle(K, V, cw, ct, lb)
        alpha = L.oracle_opt_alpha(alpha_ss, ndocs, K)
        liks.append(lik)
    q.put((rank, lo, hi, liks, alpha, lb))
    dist.destroy_process_group()


@pytest.mark.parametrize("case", ["k20"])
def test_two_rank_protocol_matches_single_process(oracle, case):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, case, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, lo0, hi0, liks0, a0, lb0), (r1, lo1, hi1, liks1, a1, lb1) = res
    assert lo0 == 0 and hi0 == lo1 and hi1 > lo1                     # shards tile the corpus
    assert liks0 == liks1 and a0 == a1 and np.array_equal(lb0, lb1)   # every rank holds the same model
    d = GOLD / case
    st = oracle.read_settings(d / "settings.txt")
    off, w, cn, tot, V = oracle.read_dat(d / "c.dat")
    alpha, lb = oracle.load_model(d / "000")
    st.em_max_iter = 2
    one = oracle.run_em(off, w, cn, V, alpha, lb, st)
    assert np.max(np.abs(np.array(liks0) - one["lik"][:3]) / np.abs(one["lik"][:3])) < 1e-12
    assert np.max(np.abs(lb0 - one["log_beta"]) / np.maximum(np.abs(one["log_beta"]), 1e-6)) < 1e-9
    assert abs(a0 - one["alpha"]) / one["alpha"] < 1e-9


def _worker_fused(rank, world, port, case, q):
    """The exchange libonilda runs over NCCL when a communicator is attached (csrc/em.cu mstep_group), restated on
    gloo: reduce-scatter of the word-major statistics by rows (rank r owns rows [r*ceil(V/P), (r+1)*ceil(V/P)), the
    table padded with zero rows), class_total partial sums of the own rows, ONE small all-reduce of
    [likelihood, alpha_ss, num_docs, class_total], lda_mle of the own rows, all-gather of the new log_beta rows."""
    sys.path.insert(0, str(ROOT / "oracle"))
    import torch
    import torch.distributed as dist
    import oracle_py
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    d = GOLD / case
    st = oracle_py.read_settings(d / "settings.txt")
    off, w, cn, tot, V = oracle_py.read_dat(d / "c.dat")
    alpha, lb = oracle_py.load_model(d / "000")
    K, D = lb.shape[0], len(off) - 1
    want = (off[-1] + D) * np.arange(world + 1) / world
    bounds = np.searchsorted(off + np.arange(D + 1), want).clip(0, D)
    bounds[0], bounds[-1] = 0, D
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    soff = (off[lo:hi + 1] - off[lo]).astype(np.int64)
    sw, sc = w[off[lo]:off[hi]].copy(), cn[off[lo]:off[hi]].copy()
    rows = (V + world - 1) // world                       # rows per rank; V need not divide
    liks = []
    L = oracle_py.lib()
    for _ in range(3):
        r = oracle_py.e_step(soff, sw, sc, V, alpha, lb, st.var_max_iter, st.var_converged)
        table = np.zeros((rows * world, K))
        table[:V] = r["class_word"].T                       # word-major, zero padding rows
        mine = torch.zeros(rows, K, dtype=torch.float64)
        for dst in range(world):                            # gloo has no reduce-scatter: one reduce per owner
            part = torch.from_numpy(table[dst * rows:(dst + 1) * rows].copy())
            dist.reduce(part, dst)
            if dst == rank:
                mine = part
        tail = torch.from_numpy(np.concatenate([[r["lik"], r["alpha_ss"], float(r["num_docs"]), 0.0], mine.numpy().sum(axis=0)]))
        dist.all_reduce(tail)
        lik, alpha_ss, ndocs = float(tail[0]), float(tail[1]), int(round(float(tail[2])))
        ct = tail[4:].numpy()
        cw = mine.numpy()
        own = np.where(cw > 0, np.log(np.where(cw > 0, cw, 1.0)) - np.log(ct)[None, :], -100.0)   # lda-model.c:35-43
        parts = [torch.zeros(rows, K, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(own))
        lb = np.ascontiguousarray(torch.cat(parts).numpy()[:V].T)
        alpha = L.oracle_opt_alpha(alpha_ss, ndocs, K)
        liks.append(lik)
    q.put((rank, rows, liks, alpha, lb))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_fused_exchange_protocol_matches_single_process(oracle, world):
    import torch.multiprocessing as mp
    case = "k20"                                           # V = 150: not a multiple of 3 (padding rows in play)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + os.getpid() % 300 + world
    procs = [ctx.Process(target=_worker_fused, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r in res[1:]:
        assert r[2] == res[0][2] and r[3] == res[0][3] and np.array_equal(r[4], res[0][4])    # same model everywhere
    d = GOLD / case
    st = oracle.read_settings(d / "settings.txt")
    off, w, cn, tot, V = oracle.read_dat(d / "c.dat")
    alpha, lb = oracle.load_model(d / "000")
    assert res[0][1] * world >= V > res[0][1] * (world - 1)
    st.em_max_iter = 2
    one = oracle.run_em(off, w, cn, V, alpha, lb, st)
    assert np.max(np.abs(np.array(res[0][2]) - one["lik"][:3]) / np.abs(one["lik"][:3])) < 1e-12
    assert np.max(np.abs(res[0][4] - one["log_beta"]) / np.maximum(np.abs(one["log_beta"]), 1e-6)) < 1e-9
    assert abs(res[0][3] - one["alpha"]) / one["alpha"] < 1e-9
