"""Wall time of the drop-in program on cfg 2: model.dat -> `lda est 2.5 20 settings 1 model.dat random out/`
-> final.beta / final.gamma / likelihood.dat, as OniLDACWrapper.scala:83-112 runs it.
python tools/cli_time.py [scale]     (needs a GPU; ONILDA_IO_THREADS=1 shows the single-thread text path)"""
import importlib
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench

oni = importlib.import_module("open-network-insight_b200")
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
cfg = dict(bench.WORKLOADS[bench.DEFAULT_WORKLOAD])
cfg["num_docs"] = max(1000, int(cfg["num_docs"] * scale))
off, w, c = bench.gen_corpus_torch(device=torch.device("cuda:0"), doc_seed=cfg["seed"], **cfg)
corpus = oni.Corpus.from_csr(off, w, c, num_terms=cfg["num_terms"])
with tempfile.TemporaryDirectory() as td:
    dat = os.path.join(td, "model.dat")
    t = time.time(); corpus.save_dat(dat); print("wrote model.dat: %.1f MB in %.2f s" % (os.path.getsize(dat) / 1e6, time.time() - t), flush=True)
    open(os.path.join(td, "settings.txt"), "w").write(
        "var max iter 20\nvar convergence 1e-6\nem max iter 20\nem convergence 1e-6\nalpha estimate\n")
    for threads in (None, "1"):
        env = dict(os.environ, ONILDA_TIMING="1", ONILDA_SEED="20161102")
        if threads:
            env["ONILDA_IO_THREADS"] = threads
        out = os.path.join(td, "out" + (threads or ""))
        t = time.time()
        r = subprocess.run([str(oni.LDA_BIN), "est", "2.5", "20", os.path.join(td, "settings.txt"), "1", dat, "random", out],
                           env=env, capture_output=True, text=True)
        wall = time.time() - t
        assert r.returncode == 0, r.stderr[-2000:]
        lines = [l for l in r.stderr.splitlines() if l.startswith("[timing] read") or l.startswith("[timing] device")]
        iters = len(open(os.path.join(out, "likelihood.dat")).read().splitlines())
        print("lda est, io threads %s: %.2f s wall, %d EM iterations, D=%d nnz=%d tokens=%d\n  %s" % (
            threads or "default(%d)" % os.cpu_count(), wall, iters, corpus.num_docs, corpus.nnz, corpus.tokens, "\n  ".join(lines)), flush=True)
