"""Timing of DNS word creation with everything resident on the device.
python tools/dns_time.py [n_events]"""
import ctypes as C
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

oni = importlib.import_module("open-network-insight_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
L = 28                                  # "ssssssss.ssssss.dddddddd.com" shaped names, random letters
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(9)
names = torch.randint(97, 105, (n, L), device=dev, dtype=torch.uint8, generator=g)      # letters a-h: repeated characters
names[:, 8] = 46
names[:, 15] = 46
names[:, 24] = 46
names[:, 25:28] = torch.tensor([99, 111, 109], device=dev, dtype=torch.uint8)
off = torch.arange(n + 1, device=dev, dtype=torch.int64) * L
frame = torch.randint(60, 1500, (n,), device=dev, dtype=torch.int32, generator=g)
ts = 1_462_000_000 + torch.randint(0, 86400, (n,), device=dev, dtype=torch.int64, generator=g)
qt = torch.randint(1, 30, (n,), device=dev, dtype=torch.int32, generator=g)
rc = torch.randint(0, 4, (n,), device=dev, dtype=torch.int32, generator=g)
code = torch.empty(n, dtype=torch.int64, device=dev)
cuts = oni.DnsCuts()
L_ = oni.lib()
L_.onilda_dns_words_device.argtypes = [C.c_void_p] * 6 + [C.c_int64] + [C.c_void_p] * 3 + [C.c_int64] + [C.c_void_p] * 3
st = torch.cuda.current_stream().cuda_stream


def run():
    oni._check(L_.onilda_dns_words_device(names.data_ptr(), off.data_ptr(), frame.data_ptr(), ts.data_ptr(), qt.data_ptr(),
                                          rc.data_ptr(), n, None, None, None, 0, C.byref(cuts), code.data_ptr(), st))


for _ in range(2):
    run()
torch.cuda.synchronize()
t = []
for _ in range(4):
    t0 = time.perf_counter(); run(); torch.cuda.synchronize(); t.append(time.perf_counter() - t0)
ms = 1e3 * float(np.mean(t))
print("dns_words device: %.2f ms per %d events (%d-byte names) = %.2f G events/s" % (ms, n, L, n / ms / 1e6))
print("entropy cuts", list(cuts.entropy), "subdomain length cuts", list(cuts.subdomain_length))
print("word of event 0:", oni.dns_word_string(int(code[0])))
