#!/usr/bin/env python3
"""Development tool: the flow scoring kernel alone on synthetic theta / phi (Dirichlet rows) and Zipf-distributed
events, cfg-5 shape.  usage: python tools/score_time.py [--events 1000000000] [--docs 1000000] [--terms 100000]"""
import argparse
import importlib
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--events", type=int, default=1_000_000_000)
    ap.add_argument("--docs", type=int, default=1_000_000)
    ap.add_argument("--terms", type=int, default=100_000)
    ap.add_argument("--topics", type=int, default=20)
    args = ap.parse_args()
    import torch
    import bench
    oni = importlib.import_module("open-network-insight_b200")
    dev = torch.device("cuda", 0)
    D, V, K = args.docs, args.terms, args.topics
    torch.manual_seed(5)
    th = torch._standard_gamma(torch.full((D, K), 0.1, dtype=torch.float64, device=dev)).clamp_min(1e-300)
    th = th / th.sum(1, keepdim=True)
    ph = torch._standard_gamma(torch.full((K, V), 0.05, dtype=torch.float64, device=dev)).clamp_min(1e-300)
    ph = (ph / ph.sum(1, keepdim=True)).t().contiguous()
    pc = torch.cumsum(ph.t().contiguous(), 1)
    pc = pc / pc[:, -1:]
    flat = (pc + torch.arange(K, device=dev, dtype=torch.float64)[:, None]).reshape(-1)
    ev = bench.gen_flow_events_torch(args.events, th, flat, D, V, K, dev)
    sc = oni.Scorer(th.cpu().numpy(), ph.cpu().numpy())
    del th, ph, pc, flat
    torch.cuda.empty_cache()
    scores = torch.empty(args.events, dtype=torch.float64, device=dev)
    ks, ts = [], []
    for it in range(8):
        t0 = time.perf_counter()
        sc.score_flow_topk_device(ev[0].data_ptr(), ev[1].data_ptr(), ev[2].data_ptr(), ev[3].data_ptr(), args.events, 1.0, 1000,
                                  scores.data_ptr())
        ts.append(time.perf_counter() - t0)
        ks.append(sc.stats()["kernel_ms"])
    print(f"score kernel {np.mean(ks[3:]):.2f} ms, score + top-1000 {1e3 * np.mean(ts[3:]):.2f} ms for {args.events} flow events "
          f"({args.events / np.mean(ks[3:]) / 1e6:.2f} G events/s kernel)")


if __name__ == "__main__":
    main()
