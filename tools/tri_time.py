"""Timing of the triples -> corpus grouping: device-resident entry with CUDA events on the
launching stream, and the host-to-host entry by wall clock.  python tools/tri_time.py [n] [num_docs]"""
import ctypes as C
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

oni = importlib.import_module("open-network-insight_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 110_000_000
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
V = 25_000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
u = torch.rand(n, device=dev, generator=g, dtype=torch.float64)
doc = ((u ** 6) * D).to(torch.int32).clamp_(0, D - 1)            # heavy head, like the hub addresses
word = torch.randint(0, V, (n,), device=dev, dtype=torch.int32, generator=g)
cnt = torch.randint(1, 100, (n,), device=dev, dtype=torch.int32, generator=g)
off = torch.empty(D + 1, dtype=torch.int64, device=dev)
wo, co = torch.empty_like(word), torch.empty_like(cnt)
tot = torch.empty(D, dtype=torch.int32, device=dev)
ml = torch.empty(1, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
L = oni.lib()


def run():
    rc = L.onilda_triples_group_device(doc.data_ptr(), word.data_ptr(), cnt.data_ptr(), n, D, off.data_ptr(), wo.data_ptr(),
                                       co.data_ptr(), tot.data_ptr(), ml.data_ptr(), st)
    oni._check(rc)


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print("device: %.3f ms per %d triples = %.2f G triples/s" % (ms, n, n / ms / 1e6))
# check against torch's own stable sort
o = torch.sort(doc, stable=True).indices
assert torch.equal(wo, word[o]) and torch.equal(co, cnt[o])
assert torch.equal(off[1:], torch.cumsum(torch.bincount(doc, minlength=D), 0)) and int(off[0]) == 0
e0.record(); o = torch.sort(doc, stable=True).indices; w2 = word[o]; c2 = cnt[o]; e1.record(); torch.cuda.synchronize()
print("torch.sort(stable) + 2 gathers: %.3f ms" % e0.elapsed_time(e1))
hd, hw, hc = doc.cpu().numpy(), word.cpu().numpy(), cnt.cpu().numpy()
for _ in range(2):
    t = time.time(); c = oni.Corpus.from_triples(hd, hw, hc, num_docs=D, num_terms=V); print("host entry: %.3f s" % (time.time() - t), flush=True)
if n <= 200_000_000:
    t = time.time(); o = np.argsort(hd, kind="stable"); w = hw[o]; c = hc[o]; print("numpy stable argsort + gathers: %.3f s" % (time.time() - t))

# ---- word counts (reduceByKey over (doc, word) occurrences)
del c, o, w2, c2
od, ow, oc = torch.empty_like(doc), torch.empty_like(doc), torch.empty_like(doc)
m = torch.zeros(1, dtype=torch.int64, device=dev)


def run_pairs():
    oni._check(L.onilda_pairs_count_device(doc.data_ptr(), word.data_ptr(), n, D, V, od.data_ptr(), ow.data_ptr(), oc.data_ptr(),
                                           m.data_ptr(), st))


for _ in range(3):
    run_pairs()
torch.cuda.synchronize()
e0.record()
for _ in range(10):
    run_pairs()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print("pairs_count device: %.3f ms per %d pairs = %.2f G pairs/s, %d distinct" % (ms, n, n / ms / 1e6, int(m.item())))
key = doc.to(torch.int64) * V + word
e0.record(); u, cu = torch.unique(key, return_counts=True); e1.record(); torch.cuda.synchronize()
print("torch.unique(return_counts) on the 64-bit composite key: %.3f ms" % e0.elapsed_time(e1))
mm = int(m.item())
assert mm == len(u) and torch.equal(od[:mm].to(torch.int64) * V + ow[:mm], u) and torch.equal(oc[:mm].to(torch.int64), cu)
