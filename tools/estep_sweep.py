#!/usr/bin/env python3
"""Development tool: E-step time at one workload under several launch settings in one process
(the settings are environment variables read by onilda_em_create).
usage: python tools/estep_sweep.py [--workload flow-1m] [--iters 8] [--scale 1.0] "VAR=val VAR2=val" "..." """
import argparse
import importlib
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default=None)
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--num-docs", type=int, default=None)
    ap.add_argument("--hub-frac", type=float, default=None)
    ap.add_argument("configs", nargs="*", default=[""])
    args = ap.parse_args()
    import torch
    import bench
    oni = importlib.import_module("open-network-insight_b200")
    wl = args.workload or bench.DEFAULT_WORKLOAD
    cfg = dict(bench.WORKLOADS[wl])
    cfg["num_docs"] = max(1000, int(cfg["num_docs"] * args.scale))
    if args.num_docs is not None:
        cfg["num_docs"] = args.num_docs
    if args.hub_frac is not None:
        cfg["hub_frac"] = args.hub_frac
    K, V = cfg["num_topics"], cfg["num_terms"]
    dev = torch.device("cuda", 0)
    offsets, words, counts = bench.gen_corpus_torch(device=dev, doc_seed=cfg["seed"], **cfg)
    corpus = oni.Corpus.from_csr(offsets, words, counts, num_terms=V)
    del words, counts
    torch.cuda.empty_cache()
    print(f"corpus D={corpus.num_docs} N={corpus.nnz} T={corpus.tokens} max_len={corpus.max_len}", flush=True)
    model0 = oni.Model.new(K, V)
    model0.init_random(20161102)
    model0.alpha = float(np.float32(bench.ALPHA0))
    st = oni.Settings.default()
    st.em_max_iter = 100
    touched = set()
    for conf in args.configs:
        for k in touched:
            os.environ.pop(k, None)
        touched.clear()
        for kv in conf.split():
            k, v = kv.split("=", 1)
            os.environ[k] = v
            touched.add(k)
        try:
            em = oni.EM(corpus, model0, st, device=0)
            ms, liks = [], []
            for _ in range(args.iters):
                em.estep()
                lik = em.mstep()[0]
                ms.append(em.stats()["estep_ms"])
                liks.append(lik)
            rare = steps = None
            if os.environ.get("ONILDA_SWEEP_COUNTS"):
                rare, steps = em.log_space_words(), em.word_steps()
            em.close()
            print(f"[{conf or 'default'}] estep_ms " + " ".join(f"{x:.1f}" for x in ms) +
                  f" | mean(4..) {np.mean(ms[3:]):.2f} | lik_last {liks[-1]:.10e}" +
                  (f" | word steps {steps:.4g}, out-of-loop words {rare}" if rare is not None else ""), flush=True)
        except Exception as e:          # keep sweeping
            print(f"[{conf}] FAILED: {e}", flush=True)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
