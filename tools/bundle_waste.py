"""How many word steps the bulk E-step EXECUTES against how many it needs (cfg 2).

A warp of the bulk kernel sweeps 8 documents together until the slowest has converged, over the longest one's
words; the others idle on zero counts.  From the per-document sweep counts of successive EM iterations this
prints, for the documents below the long-document limit: useful = sum len*sweeps, executed under the current
bundling (by length only), and executed if the bundles were instead formed from the PREVIOUS iteration's sweep
counts (sort by previous sweeps, then length).  Run on the GPU box: python tools/bundle_waste.py [iters]"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
oni = importlib.import_module("open-network-insight_b200")
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 21
LONG = int(os.environ.get("ONILDA_LONG_LEN", "1024"))
DPW = 8
cfg = dict(bench.WORKLOADS[bench.DEFAULT_WORKLOAD])
K, V = cfg["num_topics"], cfg["num_terms"]
dev = torch.device("cuda", 0)
offsets, words, counts = bench.gen_corpus_torch(device=dev, doc_seed=cfg["seed"], **cfg)
corpus = oni.Corpus.from_csr(offsets, words, counts, num_terms=V)
m = oni.Model.new(K, V); m.init_random(20161102); m.alpha = float(np.float32(bench.ALPHA0))
s = oni.Settings.default(); s.em_max_iter = iters + 1
em = oni.EM(corpus, m, s, device=0)
lens = np.diff(offsets)
bulk = np.nonzero(lens < LONG)[0]

def executed(order):
    """order: bulk documents in bundle order; 8 per warp; warp runs max_len x max_sweeps x 8 slots"""
    n = len(order)
    pad = (-n) % DPW
    l = np.concatenate([lens[order], np.zeros(pad, np.int64)]).reshape(-1, DPW)
    w = np.concatenate([sw[order], np.zeros(pad, np.int64)]).reshape(-1, DPW)
    return int((l.max(1) * w.max(1)).sum()) * DPW, int((l.max(1) * w.max(1)).sum())

by_len = bulk[np.argsort(-lens[bulk], kind="stable")]


def windows(order, max_size=2048, ratio=1.1):
    """boundaries over the length-sorted order: at most max_size documents, longest / shortest <= ratio"""
    l = lens[order]
    b = [0]
    while b[-1] < len(order):
        lo = b[-1]
        hi = min(lo + max_size, len(order))
        # first position where the length falls below l[lo] / ratio
        cut = lo + int(np.searchsorted(-l[lo:hi], -l[lo] / ratio, side="right"))
        b.append(max(lo + 1, min(hi, cut)))
    return np.array(b)


win = windows(by_len)
win_id = np.repeat(np.arange(len(win) - 1), np.diff(win))
prev = None
for it in range(iters):
    em.estep(); em.mstep()
    sw = em.doc_sweeps().astype(np.int64)
    useful = int((lens[bulk] * sw[bulk]).sum())
    ex_len, warp_len = executed(by_len)
    line = f"it {it:2d} estep {em.stats()['estep_ms']:7.1f} ms  bulk useful {useful:.3e}  executed(by len) {ex_len:.3e} = x{ex_len / useful:.2f}  warp-steps {warp_len:.3e}"
    if prev is not None:
        # bundles from the previous iteration's sweeps: primary key previous sweeps (descending), then length
        o = bulk[np.lexsort((-lens[bulk], -prev[bulk]))]
        ex_p, warp_p = executed(o)
        # and with length as the primary key in coarse classes (powers of sqrt 2), previous sweeps inside a class
        cls = np.floor(2 * np.log2(np.maximum(lens[bulk], 1))).astype(np.int64)
        o2 = bulk[np.lexsort((-lens[bulk], -prev[bulk], -cls))]
        ex_c, warp_c = executed(o2)
        line += f"  | by prev sweeps x{ex_p / useful:.2f} (warp-steps {warp_p:.3e})  | len class, prev sweeps x{ex_c / useful:.2f} (warp-steps {warp_c:.3e})"
        o3 = by_len[np.lexsort((np.arange(len(by_len)), -prev[by_len], win_id))]      # stable sort by previous sweeps inside each window
        ex_w, warp_w = executed(o3)
        line += f"  | windows x{ex_w / useful:.2f} (warp-steps {warp_w:.3e})"
        line += f"  corr {np.corrcoef(prev[bulk], sw[bulk])[0, 1]:.3f}"
    print(line, flush=True)
    prev = sw
