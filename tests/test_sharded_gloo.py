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
