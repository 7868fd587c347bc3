"""Timing of flow word creation with the columns resident on the device.
python tools/words_time.py [n_events]"""
import ctypes as C
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

oni = importlib.import_module("open-network-insight_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000_000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
hour = torch.randint(0, 24, (n,), device=dev, dtype=torch.int32, generator=g)
minute = torch.randint(0, 60, (n,), device=dev, dtype=torch.int32, generator=g)
second = torch.randint(0, 60, (n,), device=dev, dtype=torch.int32, generator=g)
ibyt = torch.empty(n, device=dev, dtype=torch.float32).normal_(6, 2.5, generator=g).exp_().to(torch.int64)
ipkt = (ibyt // torch.randint(40, 1500, (n,), device=dev, dtype=torch.int64, generator=g)).clamp_(min=1)
ports = torch.tensor([0, 22, 53, 80, 443, 1024, 1025, 8080, 50123, 65535], device=dev, dtype=torch.int32)
sport = ports[torch.randint(0, 10, (n,), device=dev, generator=g)]
dport = ports[torch.randint(0, 10, (n,), device=dev, generator=g)]
sc, dc = torch.empty(n, dtype=torch.int32, device=dev), torch.empty(n, dtype=torch.int32, device=dev)
cuts = oni.FlowCuts()
L = oni.lib()
st = torch.cuda.current_stream().cuda_stream


def run():
    oni._check(L.onilda_flow_words_device(hour.data_ptr(), minute.data_ptr(), second.data_ptr(), ibyt.data_ptr(), ipkt.data_ptr(),
                                          sport.data_ptr(), dport.data_ptr(), n, C.byref(cuts), sc.data_ptr(), dc.data_ptr(), st))


for _ in range(2):
    run()
torch.cuda.synchronize()
ts = []
for _ in range(5):
    t0 = time.perf_counter(); run(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
ms = 1e3 * float(np.mean(ts))
print("flow_words device: %.2f ms per %d events = %.2f G events/s (44 B per event algorithmic: %.0f GB/s)" % (ms, n, n / ms / 1e6, 44 * n / ms / 1e6))
print("ibyt cuts", list(cuts.ibyt)); print("ipkt cuts", list(cuts.ipkt)); print("time cuts", [round(x, 4) for x in cuts.time])
# cross-check the cuts against torch's own sort on a column
srt = torch.sort(ibyt).values
want = [float(srt[max(0, int(np.ceil(q * n)) - 1)]) for q in (0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0)]
print("ibyt cuts match sorted column:", want == list(cuts.ibyt))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); torch.sort(ibyt); e1.record(); torch.cuda.synchronize()
print("torch.sort of the ibyt column alone: %.2f ms" % e0.elapsed_time(e1))

# ---- the whole pre-LDA chain in one call, from host columns (pageable) to a host corpus
m = min(n, 100_000_000)
D = 1_000_000
sip = (torch.rand(m, device=dev, generator=g) ** 4 * D).to(torch.int32).clamp_(0, D - 1).cpu().numpy()
dip = (torch.rand(m, device=dev, generator=g) ** 4 * D).to(torch.int32).clamp_(0, D - 1).cpu().numpy()
cols = [x[:m].cpu().numpy() for x in (hour, minute, second, ibyt, ipkt, sport, dport)]
del srt, sc, dc
torch.cuda.empty_cache()
for _ in range(2):
    t0 = time.perf_counter()
    cuts2, corpus, vocab, sw, dw = oni.flow_prelda(*cols, sip, dip, num_docs=D)
    dt = time.perf_counter() - t0
print("flow_prelda (host columns -> host corpus): %.3f s per %d events = %.2f G events/s; corpus %d docs, %d entries, %d words" % (
    dt, m, m / dt / 1e9, corpus.num_docs, corpus.nnz, corpus.num_terms))
