"""How much lockstep work the bulk E-step launch wastes: documents run in bundles of 8 (4 lanes each), a bundle runs
max(len) x max(sweeps) word steps for each of its 8 slots.  Compares the current order (by length) with an order by
(previous iteration's sweeps, length).  python tools/bundle_waste.py [scale]"""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench

oni = importlib.import_module("open-network-insight_b200")
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
cfg = dict(bench.WORKLOADS[bench.DEFAULT_WORKLOAD])
cfg["num_docs"] = int(cfg["num_docs"] * scale)
off, w, c = bench.gen_corpus_torch(device=torch.device("cuda:0"), doc_seed=cfg["seed"], **cfg)
corpus = oni.Corpus.from_csr(off, w, c, num_terms=cfg["num_terms"])
m = oni.Model.new(20, cfg["num_terms"]); m.init_random(20161102); m.alpha = 2.5
em = oni.EM(corpus, m, oni.Settings.default())
lens = np.diff(off)
bulk = np.nonzero(lens < 1024)[0]


def cost(order, sweeps):
    o = order[: len(order) // 8 * 8].reshape(-1, 8)
    return float((lens[o].max(1) * sweeps[o].max(1)).sum() * 8)


prev = None
for it in range(21):
    em.iterate()
    sw = em.doc_sweeps().astype(np.int64)
    useful = float((lens[bulk] * sw[bulk]).sum())
    by_len = bulk[np.argsort(-lens[bulk], kind="stable")]
    line = "iter %2d: useful %.3g steps, by length x%.2f" % (it + 1, useful, cost(by_len, sw) / useful)
    if prev is not None:
        key = np.lexsort((-lens[bulk], -prev[bulk]))
        line += ", by (previous sweeps, length) x%.2f" % (cost(bulk[key], sw) / useful)
        key2 = np.lexsort((-lens[bulk], -sw[bulk]))
        line += ", oracle (this iteration's sweeps) x%.2f" % (cost(bulk[key2], sw) / useful)
    print(line + "  sweeps hist(1,2,3-5,6-19,20): %s" % [int(((sw[bulk] >= a) & (sw[bulk] <= b)).sum()) for a, b in ((1, 1), (2, 2), (3, 5), (6, 19), (20, 20))], flush=True)
    prev = sw
