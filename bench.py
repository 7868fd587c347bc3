#!/usr/bin/env python
"""bench.py - LDA EM tokens/s per EM iteration (BASELINE.json's metric) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one EM iteration (E-step + sufficient-statistics exchange + M-step) over the synthetic one-day
netflow corpus of BASELINE.json configs[1]: 1M IP documents, ~100M tokens, V = 100k flow words, K = 20, lda-c's
default settings.  At N > 1 (torchrun, one rank per GPU) the headline `value` keeps round 1's meaning - every rank
holds its own 1M-document corpus (weak scaling) - and the statistics go through the library's own NCCL exchange
(reduce-scatter -> lda_mle of the own rows -> all-gather, csrc/em.cu).  North_star's split - ONE corpus sharded over
the N GPUs - is measured beside it in `strong`, with the exchange timed by itself (bus GB/s), the longest document
and its lone-warp time; `strong_dns` does the same for configs[2]; `scoring.sharded` shards configs[4]'s events.

Prints ONE JSON line (contract in the task statement).  `value` is measured with the corpus resident in HBM;
`e2e` runs the same iterations through the C ABI with HOST buffers (corpus + model host->device and model + gamma
device->host inside every timed step).  `roofline` is for the E-step; `cpu_baseline` is the unmodified reference
lda-c (oracle/_ref/lda) on this box's host cores on a bounded document prefix, `parity` compares that very run's
likelihood.dat / final.gamma / final.beta / top-1000 events with the GPU path on the same documents.

`--impl reference` times the unmodified reference only: that process never loads libonilda (its inputs are written
with numpy); PyTorch generates the synthetic corpus there as here.

PyTorch is plumbing (synthetic data on the GPU, pinned host buffers, torch.distributed for rendezvous and for the
arithmetic of the max-over-ranks timing); every timed kernel is libonilda's.
"""
import argparse
import importlib
import importlib.util
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    # BASELINE.json configs[1]
    "netflow-1M-docs-100M-tokens-V100k-K20": dict(num_docs=1_000_000, mean_tokens=100, num_terms=100_000,
                                                  num_topics=20, seed=20161102),
    # BASELINE.json configs[2] and [3]: parity/scale cases, not the headline line (run with --workload, usually --scale < 1)
    "dns-2M-docs-200M-tokens-V50k-K20": dict(num_docs=2_000_000, mean_tokens=100, num_terms=50_000,
                                             num_topics=20, seed=20161103),
    "proxy-5M-docs-1B-tokens-V500k-K50": dict(num_docs=5_000_000, mean_tokens=200, num_terms=500_000,
                                              num_topics=50, seed=20161104),
}
DEFAULT_WORKLOAD = "netflow-1M-docs-100M-tokens-V100k-K20"
ALPHA0 = 2.5                     # OniLDACWrapper.scala:83


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ----------------------------------------------------------------- synthetic corpus (torch on the GPU: plumbing)

def gen_corpus_torch(num_docs, mean_tokens, num_terms, num_topics, seed, device, doc_seed=None, max_tokens=20_000,
                     hub_frac=0.001, hub_tokens=20_000, dup_frac=0.01, chunk_docs=50_000):
    """Same recipe as open-network-insight_b200/corpusgen.py (LDA generative model, lognormal
    document sizes, hub documents, a few duplicated word ids), run on the GPU so that the
    100M-token corpus takes seconds.  Returns host CSR arrays."""
    import torch
    cg = importlib.import_module("open-network-insight_b200.corpusgen")
    D, V, K = num_docs, num_terms, num_topics
    torch.manual_seed(seed)
    beta = torch._standard_gamma(torch.full((K, V), 0.05, dtype=torch.float64, device=device)).clamp_min(1e-300)
    cdf_w = torch.cumsum(beta / beta.sum(1, keepdim=True), 1)
    cdf_w = cdf_w / cdf_w[:, -1:]
    cdf_w_flat = (cdf_w + torch.arange(K, device=device, dtype=torch.float64)[:, None]).reshape(-1)
    torch.manual_seed(seed if doc_seed is None else doc_seed)
    sigma = 0.8
    mu = np.log(mean_tokens) - 0.5 * sigma * sigma
    n_tok = torch.exp(mu + sigma * torch.randn(D, device=device, dtype=torch.float64)).clamp(5, max_tokens).long()
    n_hub = int(round(hub_frac * D))
    if n_hub:
        hubs = torch.randperm(D, device=device)[:n_hub]
        n_tok[hubs] = torch.randint(hub_tokens // 4, hub_tokens + 1, (n_hub,), device=device)
    lens, words, counts = [], [], []
    for lo in range(0, D, chunk_docs):
        hi = min(D, lo + chunk_docs)
        nd = hi - lo
        theta = torch._standard_gamma(torch.full((nd, K), 0.1, dtype=torch.float64, device=device)).clamp_min(1e-300)
        cdf_t = torch.cumsum(theta / theta.sum(1, keepdim=True), 1)
        cdf_t = cdf_t / cdf_t[:, -1:]
        rows = torch.arange(nd, device=device, dtype=torch.float64)
        cdf_t_flat = (cdf_t + rows[:, None]).reshape(-1)
        doc = torch.repeat_interleave(torch.arange(nd, device=device), n_tok[lo:hi])
        T = doc.numel()
        z = torch.searchsorted(cdf_t_flat, torch.rand(T, device=device, dtype=torch.float64) + doc.double(), right=True)
        z = (z - doc * K).clamp_(0, K - 1)
        w = torch.searchsorted(cdf_w_flat, torch.rand(T, device=device, dtype=torch.float64) + z.double(), right=True)
        w = (w - z * V).clamp_(0, V - 1)
        key, cnt = torch.unique(doc * V + w, return_counts=True)
        lens.append(torch.bincount(key // V, minlength=nd).cpu().numpy())
        words.append((key % V).to(torch.int32).cpu().numpy())
        counts.append(cnt.to(torch.int32).cpu().numpy())
        del theta, cdf_t, cdf_t_flat, doc, z, w, key, cnt
    lens = np.concatenate(lens).astype(np.int64)
    words = np.concatenate(words)
    counts = np.concatenate(counts)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    rng = np.random.default_rng(seed + 17)
    offsets, words, counts = cg.add_duplicates(offsets, words, counts, rng, dup_frac)
    return offsets, words, counts


# ----------------------------------------------------------------- clocks

class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------- reference lda-c on the host cores

def bind_near_gpu(local_rank):
    """One process per GPU: run on the cores NVML lists for this rank's GPU, so that the host buffers this process
    allocates from here on (first touch) and the copies to and from them stay on the GPU's own NUMA node instead of all
    ranks sharing whichever node they were started on.  Returns the number of cores bound to (0: left as it was)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        index = local_rank
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        if vis and all(x.strip().isdigit() for x in vis.split(",")):
            index = int(vis.split(",")[local_rank])
        mask = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(index), (os.cpu_count() + 63) // 64)
        near = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (int(w) >> b) & 1}
        cpus = sorted(near & os.sched_getaffinity(0))
        if not cpus:
            return 0
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception as e:          # no NVML, no such call on this platform: leave the affinity alone
        log(f"[rank {local_rank}] no NUMA binding: {e}")
        return 0


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def start_model_numpy(K, V, seed):
    """A random start as random_initialize_ss + lda_mle make it (lda-model.c:104-116, 27-44): class_word =
    1/V + U[0,1), log-normalised per topic.  numpy's generator, so both arms can write it without libonilda."""
    rng = np.random.default_rng(seed)
    cw = 1.0 / V + rng.random((K, V))
    return ALPHA0, np.log(cw) - np.log(cw.sum(axis=1, keepdims=True))


def write_model_text(root, alpha, log_beta):
    """<root>.beta / <root>.other in save_lda_model's format (lda-model.c:193-217)."""
    K, V = log_beta.shape
    np.savetxt(str(root) + ".beta", log_beta, fmt=" %5.10f", delimiter="")
    Path(str(root) + ".other").write_text(f"num_topics {K}\nnum_terms {V}\nalpha {alpha:5.10f}\n")


def write_dat_numpy(path, offsets, words, counts, n_docs):
    """The first n_docs documents as a .dat file, built with numpy only.  read_data is token based and takes
    "%10d" fields (lda-data.c:22-59): numbers are written zero-padded to a fixed width so that the whole file is
    one vectorised byte fill: "\\n<7-digit length>" per document, " <7-digit word>:<6-digit count>" per entry."""
    off = np.asarray(offsets[:n_docs + 1], np.int64)
    n = int(off[-1])
    w = np.asarray(words[:n], np.int64)
    c = np.asarray(counts[:n], np.int64)
    assert n_docs == 0 or (np.diff(off).max() < 10 ** 7 and (n == 0 or (w.max() < 10 ** 7 and c.max() < 10 ** 6)))

    def digits(v, width):
        return ((v[:, None] // (10 ** np.arange(width - 1, -1, -1, dtype=np.int64))) % 10 + 48).astype(np.uint8)
    HW, EW = 8, 15
    head = np.empty((n_docs, HW), np.uint8)
    head[:, 0] = 10
    head[:, 1:] = digits(np.diff(off), 7)
    ent = np.empty((n, EW), np.uint8)
    ent[:, 0] = 32
    ent[:, 1:8] = digits(w, 7)
    ent[:, 8] = 58
    ent[:, 9:] = digits(c, 6)
    out = np.empty(n_docs * HW + n * EW + 1, np.uint8)
    doc_of = np.repeat(np.arange(n_docs, dtype=np.int64), np.diff(off))
    out[(np.arange(n_docs, dtype=np.int64) * HW + off[:-1] * EW)[:, None] + np.arange(HW)] = head
    out[(np.arange(n, dtype=np.int64) * EW + (doc_of + 1) * HW)[:, None] + np.arange(EW)] = ent
    out[-1] = 10
    out.tofile(str(path))


def run_reference(workdir, offsets, words, counts, n_docs, K, V, n_iters, nproc, start):
    """Runs the UNMODIFIED reference (oracle/_ref/lda; fork-based MPI stand-in with `nproc` ranks) for n_iters EM
    iterations on the first n_docs documents from the start model `start` = (alpha, log_beta).  Inputs are written
    with numpy.  Returns (seconds per iteration read off its own progress lines, output directory) or None."""
    ref_bin = ROOT / "oracle" / "_ref" / "lda"
    if not ref_bin.exists():
        return None
    work = Path(workdir)
    write_dat_numpy(work / "sample.dat", offsets, words, counts, n_docs)
    write_model_text(work / "000", *start)
    (work / "settings.txt").write_text(
        f"var max iter 20\nvar convergence 1e-6\nem max iter {n_iters - 1}\nem convergence 1e-6\nalpha estimate\n")
    env = dict(os.environ, ONI_MPI_STUB_NP=str(nproc), ONI_MPI_STUB_MAXCOUNT=str(max(V, 1 << 16)))
    if nproc <= 1:
        env.pop("ONI_MPI_STUB_NP")
    cmd = ["stdbuf", "-oL", str(ref_bin), "est", str(ALPHA0), str(K), "settings.txt", str(nproc), "sample.dat", "000", "out"]
    starts, ends = [], []
    p = subprocess.Popen(cmd, cwd=work, env=env, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    for line in p.stdout:
        now = time.perf_counter()
        if line.startswith("**** em iteration"):
            starts.append(now)
        elif line.startswith("**** reducing likelihood"):
            ends.append(now)
    p.wait()
    if p.returncode != 0 or not starts or len(ends) < len(starts):
        return None
    return [e - s for s, e in zip(starts, ends)], work / "out"


def reference_sample_docs(corpus_offsets, cores, target_seconds=8.0, nnz_per_core_second=120_000):
    """Document prefix whose nnz gives roughly target_seconds per EM iteration on `cores` cores."""
    want = int(cores * nnz_per_core_second * target_seconds)
    n = int(np.searchsorted(corpus_offsets, want))
    return max(min(n, len(corpus_offsets) - 1), 1)


# ----------------------------------------------------------------- scoring leg (BASELINE.json configs[4])

def gen_flow_events_torch(n, theta_t, phi_cdf_flat, D, V, K, device, chunk=1 << 24, unknown_frac=0.001):
    """Flow events (src_doc, src_word, dst_doc, dst_word) int32 on the device: documents Zipf(1.1)
    over D, words drawn from the document's theta^T phi, a few unknown (-1) ids."""
    import torch
    rank = torch.arange(1, D + 1, device=device, dtype=torch.float64) ** -1.1
    doc_cdf = torch.cumsum(rank, 0)
    doc_cdf = doc_cdf / doc_cdf[-1]
    theta_cdf = torch.cumsum(theta_t, 1)
    theta_cdf = (theta_cdf / theta_cdf[:, -1:].clamp_min(1e-300)).float()
    out = [torch.empty(n, dtype=torch.int32, device=device) for _ in range(4)]
    for side in range(2):
        for lo in range(0, n, chunk):
            m = min(chunk, n - lo)
            doc = torch.searchsorted(doc_cdf, torch.rand(m, device=device, dtype=torch.float64)).clamp_(0, D - 1)
            z = (torch.rand(m, 1, device=device) > theta_cdf[doc]).sum(1).clamp_(0, K - 1)
            w = torch.searchsorted(phi_cdf_flat, torch.rand(m, device=device, dtype=torch.float64) + z.double(), right=True)
            w = (w - z * V).clamp_(0, V - 1)
            doc = doc.to(torch.int32)
            w = w.to(torch.int32)
            doc[torch.rand(m, device=device) < unknown_frac] = -1
            w[torch.rand(m, device=device) < unknown_frac] = -1
            out[2 * side][lo:lo + m] = doc
            out[2 * side + 1][lo:lo + m] = w
            del doc, z, w
    return out


def scoring_roofline(n_events, K, kernel_s, hbm_peak):
    """The scorer's rows come from L2, not HBM (ncu: 84 % L2 hits, 58 B/event of DRAM traffic for 672 B/event of row
    gathers), so its ceiling is what L2 delivers to a gather of 160-byte rows: measured by tools/l2_gather_bench.cu
    (profiles/l2_gather_r2.json) as 14.3 TB/s for a 16 MB table (phi) and 7.4 TB/s for a 160 MB one under uniform
    ids (theta; Zipf ids do better).  `frac` is against the L2-resident figure; the HBM view of SURVEY 8(d) is kept
    beside it, labelled."""
    b_alg = n_events * 2 * (16 + 16 * K)
    rows_bytes = n_events * 4 * 8 * K
    l2 = {"16MB": 14323.2, "160MB": 7394.8}
    src = "profiles/l2_gather_r2.json (defaults compiled into bench.py if the file is absent)"
    p = ROOT / "profiles" / "l2_gather_r2.json"
    if p.exists():
        try:
            l2 = json.loads(p.read_text())["rows_gbs"]
        except (ValueError, KeyError):
            pass
    traffic = lts = l1tex = None
    t = ROOT / "profiles" / "score_traffic.json"
    if t.exists():
        try:
            tj = json.loads(t.read_text())
            traffic = tj.get("dram_bytes_per_event", 0) * n_events or None
            lts, l1tex = tj.get("lts_throughput_pct"), tj.get("l1tex_throughput_pct")
        except ValueError:
            pass
    achieved = rows_bytes / kernel_s / 1e9
    blended_s = (rows_bytes / 2) / (l2["16MB"] * 1e9) + (rows_bytes / 2) / (l2["160MB"] * 1e9)
    return {"bound": "l2", "kernel": "score_pairs_kernel<true, 10>", "achieved": achieved, "peak": l2["16MB"], "unit": "GB/s",
            "frac": achieved / l2["16MB"], "traffic": traffic, "peak_source": src,
            "row_bytes_per_launch": rows_bytes, "lts_throughput_pct_of_peak": lts, "l1tex_throughput_pct_of_peak": l1tex,
            "binding_unit": "L1TEX (ncu: 89 % of its peak; L2 33 %, DRAM 10 %): 16-byte pieces of 160-byte rows straddle 128-byte lines, "
                            "and the products' transpose goes through shared memory in the same unit; the microbenchmark that sets "
                            "`peak` gathers with the same shape, so it meets the same unit",
            "gather_floor_ms": 1e3 * blended_s,
            "gather_floor_note": "time the microbenchmark needs for the same rows: half from a 16 MB table (phi), half from a 160 MB "
                                 "table under uniform ids (theta); the kernel sits at that floor, i.e. at what L2 + HBM deliver to "
                                 "this access shape",
            "hbm_view": {"algorithmic_bytes_per_launch": b_alg, "achieved_gbs": b_alg / kernel_s / 1e9, "hbm_peak_gbs": hbm_peak,
                         "frac_of_hbm": b_alg / kernel_s / 1e9 / hbm_peak,
                         "note": "SURVEY 8(d)'s figure (672 B per flow event, every row gather charged to HBM): algorithmic, L2-served, "
                                 "which is why it exceeds 1"},
            "compulsory_bytes_per_launch": 24 * n_events, "compulsory_gbs": 24 * n_events / kernel_s / 1e9}


def scoring_leg(oni, torch, device, gamma, log_beta, n_events, n_e2e, steps, peak, cpu_cores=0):
    """Events scored per second: device-resident (score kernel + top-1000 select) and end to end
    from pinned host buffers through onilda_score_flow_min."""
    D, K = gamma.shape
    V = log_beta.shape[1]
    theta = oni.theta_from_gamma(gamma)
    phi = oni.phi_from_log_beta(log_beta)
    theta_t = torch.from_numpy(theta).to(device)
    phi_t = torch.from_numpy(phi).to(device)                      # [V][K]
    pc = torch.cumsum(phi_t.t().contiguous(), 1)
    pc = pc / pc[:, -1:]
    phi_cdf_flat = (pc + torch.arange(K, device=device, dtype=torch.float64)[:, None]).reshape(-1)
    ev = gen_flow_events_torch(n_events, theta_t, phi_cdf_flat, D, V, K, device)
    del theta_t, phi_t, pc, phi_cdf_flat
    torch.cuda.empty_cache()
    sc = oni.Scorer(theta, phi)
    scores = torch.empty(n_events, dtype=torch.float64, device=device)
    torch.cuda.synchronize()
    kms, tot = [], []
    for it in range(3 + steps):
        t0 = time.perf_counter()
        idx, val = sc.score_flow_topk_device(ev[0].data_ptr(), ev[1].data_ptr(), ev[2].data_ptr(), ev[3].data_ptr(),
                                             n_events, 1.0, 1000, scores.data_ptr())
        dt = time.perf_counter() - t0
        if it >= 3:
            kms.append(sc.stats()["kernel_ms"])
            tot.append(dt)
    kernel_s = float(np.mean(kms)) * 1e-3
    b_alg = n_events * 2 * (16 + 16 * K)
    res = {"events": n_events, "events_per_sec": n_events / float(np.mean(tot)), "unit": "flow events/s",
           "score_kernel_ms": 1e3 * kernel_s, "score_plus_top1000_ms": 1e3 * float(np.mean(tot)),
           "most_suspicious_score": float(val[0]) if len(val) else None,
           "roofline": scoring_roofline(n_events, K, kernel_s, peak)}
    # end to end from pinned host memory
    n2 = min(n_e2e, n_events)
    host = [torch.empty(n2, dtype=torch.int32, pin_memory=True) for _ in range(4)]
    for h, d in zip(host, ev):
        h.copy_(d[:n2])
    out = torch.empty(n2, dtype=torch.float64, pin_memory=True)
    torch.cuda.synchronize()
    ts = []
    for it in range(3 + steps):
        t0 = time.perf_counter()
        sc.score_raw(host[0].data_ptr(), host[1].data_ptr(), host[2].data_ptr(), host[3].data_ptr(), n2, out.data_ptr())
        if it >= 3:
            ts.append(time.perf_counter() - t0)
    check = bool(torch.equal(out.to(device), scores[:n2]))
    res["e2e"] = {"value": n2 / float(np.mean(ts)), "unit": "flow events/s", "events": n2,
                  "h2d_bytes_per_step": 16 * n2, "d2h_bytes_per_step": 8 * n2, "matches_device_leg": check}
    res["gpu_launches"] = sc.stats()["launches"]
    sc.close()
    if cpu_cores > 0:
        # CPU baseline: the C restatement of SuspiciousConnectsScoreFunction + flow min (oracle/score_oracle.c;
        # the Scala/Spark original cannot run here: no JVM), all host threads, on a bounded sample
        sys.path.insert(0, str(ROOT / "oracle"))
        import oracle_py
        m = min(n2, 1 << 25)
        ids = [h.numpy()[:m] for h in host]
        oracle_py.score_flow_min(*[x[:1 << 16] for x in ids], theta, phi, threads=cpu_cores)      # warm
        t0 = time.perf_counter()
        ref = oracle_py.score_flow_min(*ids, theta, phi, threads=cpu_cores)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": m / dt, "unit": "flow events/s", "cores": cpu_cores, "kind": "port",
                               "sample": f"first {m} events, oracle/score_oracle.c with {cpu_cores} pthreads",
                               "bit_identical_to_gpu": bool(np.array_equal(ref, out.numpy()[:m]))}
    return res


def grouping_leg(oni, torch, device, offsets, words, counts, steps, peak, cpu=True):
    """(document, word, count) triples in arrival order -> CSR corpus
    (OniLDACWrapper.createModel).  The workload's own nnz entries, shuffled."""
    D, n = len(offsets) - 1, int(offsets[-1])
    lens = torch.from_numpy(np.diff(offsets)).to(device)
    doc = torch.repeat_interleave(torch.arange(D, device=device, dtype=torch.int32), lens)
    perm = torch.randperm(n, device=device, generator=torch.Generator(device=device).manual_seed(7))
    doc = doc[perm].contiguous()
    word = torch.from_numpy(np.asarray(words[:n])).to(device)[perm].contiguous()
    cnt = torch.from_numpy(np.asarray(counts[:n])).to(device)[perm].contiguous()
    del perm, lens
    off = torch.empty(D + 1, dtype=torch.int64, device=device)
    wo, co = torch.empty_like(word), torch.empty_like(cnt)
    tot = torch.empty(D, dtype=torch.int32, device=device)
    ml = torch.empty(1, dtype=torch.int32, device=device)
    st = torch.cuda.current_stream().cuda_stream
    L = oni.lib()

    def run():
        oni._check(L.onilda_triples_group_device(doc.data_ptr(), word.data_ptr(), cnt.data_ptr(), n, D, off.data_ptr(),
                                                 wo.data_ptr(), co.data_ptr(), tot.data_ptr(), ml.data_ptr(), st))
    for _ in range(3):
        run()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(steps):
        run()
    ev[1].record()
    torch.cuda.synchronize()
    sec = ev[0].elapsed_time(ev[1]) * 1e-3 / steps
    order = torch.sort(doc, stable=True).indices
    ok = bool(torch.equal(wo, word[order]) and torch.equal(co, cnt[order]) and
              torch.equal(off.cpu(), torch.from_numpy(np.asarray(offsets, np.int64))) and int(ml.item()) == int(np.diff(offsets).max()))
    del order
    b_alg = 20 * n + 12 * D
    bits = max(1, int(D - 1).bit_length())
    passes = (bits + 7) // 8
    res = {"triples": n, "num_docs": D, "triples_per_sec": n / sec, "unit": "triples/s", "ms": 1e3 * sec,
           "matches_torch_stable_sort": ok, "gpu_launches_per_call": 5 * passes + 2,
           "roofline": {"bound": "hbm", "kernel": "radix_scatter_kernel (x%d passes) + histogram/scan/totals/offsets" % passes,
                        "achieved": b_alg / sec / 1e9, "peak": peak, "unit": "GB/s", "frac": b_alg / sec / 1e9 / peak,
                        "algorithmic_bytes_per_launch": b_alg, "pass_bytes": (28 * passes + 8) * n, "traffic": None,
                        "note": "algorithmic = triples in (12 B) + word,count out (8 B) + offsets,totals; the LSD sort moves "
                                "28 B per triple per pass (4 B histogram read, 12 B in, 12 B out), pass_bytes in total"}}
    hd, hw, hc = doc.cpu().numpy(), word.cpu().numpy(), cnt.cpu().numpy()
    del doc, word, cnt, wo, co
    ts = []
    for it in range(3):
        t0 = time.perf_counter()
        c = oni.Corpus.from_triples(hd, hw, hc, num_docs=D)
        if it >= 1:
            ts.append(time.perf_counter() - t0)
    res["e2e"] = {"value": n / float(np.mean(ts)), "unit": "triples/s", "h2d_bytes_per_step": 12 * n,
                  "d2h_bytes_per_step": 8 * n + 12 * D, "note": "pageable host arrays in, host corpus out (onilda_corpus_from_triples)"}
    if cpu:
        m = min(n, 20_000_000)
        t0 = time.perf_counter()
        o = np.argsort(hd[:m], kind="stable")
        w2, c2 = hw[:m][o], hc[:m][o]
        lens2 = np.bincount(hd[:m], minlength=D)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": m / dt, "unit": "triples/s", "cores": 1, "kind": "port",
                               "sample": f"first {m} triples, numpy stable argsort + gathers + bincount (oracle_py.triples_to_csr's "
                                         "steps; the Spark groupByKey original cannot run here: no JVM)"}
        del o, w2, c2, lens2
    return res


# ----------------------------------------------------------------- parity against the reference on the same documents

def _rel(a, b, floor):
    return np.abs(a - b) / np.maximum(np.abs(b), floor)


def theta_phi_numpy(gamma, log_beta):
    """OniLDACWrapper.scala:121-144, 170-195 in numpy (fp64): theta rows = gamma / sum (zeros if the sum is not
    positive), phi[w][k] = exp(log_beta[k][w]) / sum_w exp(log_beta[k][w])."""
    s = gamma.sum(axis=1, keepdims=True)
    theta = np.where(s > 0, gamma / np.where(s > 0, s, 1.0), 0.0)
    p = np.exp(log_beta)
    return theta, np.ascontiguousarray((p / p.sum(axis=1, keepdims=True)).T)


def flow_min_numpy(ev, theta, phi):
    """score = min over the two sides of sum_k theta[ip][k] phi[word][k] (SuspiciousConnectsScoreFunction.scala:11-23,
    FlowPostLDA.scala:99-125): unknown word -> uniform 1/K; unknown ip -> the inner join drops the event (:94), here +inf."""
    K = theta.shape[1]
    out = np.full(len(ev[0]), np.inf)
    for doc, word in ((ev[0], ev[1]), (ev[2], ev[3])):
        known_d = (doc >= 0) & (doc < theta.shape[0])
        known_w = (word >= 0) & (word < phi.shape[0])
        t = theta[np.where(known_d, doc, 0)]
        p = np.where(known_w[:, None], phi[np.where(known_w, word, 0)], 1.0 / K)
        s = np.zeros(len(doc))
        for k in range(K):                      # left to right, as Scala's zip/map/sum
            s += t[:, k] * p[:, k]
        s[~known_d] = np.inf
        out = s if doc is ev[0] else np.where(np.isinf(out) | np.isinf(s), np.inf, np.minimum(out, s))
    return out


def parity_leg(oni, torch, device, offsets, words, counts, K, V, cores, local_rank, n_iters=3, target_seconds=6.0):
    """The unmodified reference on a document prefix (all host cores), the GPU path on the SAME documents from the
    SAME 000 files, and what the two wrote compared with north_star's bars: likelihood.dat per iteration (1e-6),
    final.gamma / final.beta (1e-4 max relative), the 1000 most suspicious of 2M flow events (99 % overlap).
    Also the timing baselines: the reference on all cores and as a single process, the GPU on the same prefix."""
    n_docs = reference_sample_docs(offsets, cores, target_seconds=target_seconds)
    start = start_model_numpy(K, V, 20161102)
    out = {"cpu_baseline": None, "parity": None}
    with tempfile.TemporaryDirectory() as td:
        td = Path(td)
        ref = run_reference(td, offsets, words, counts, n_docs, K, V, n_iters, cores, start)
        nnz = int(offsets[n_docs])
        tok = int(np.asarray(counts[:nnz], np.int64).sum())
        if ref is None:
            out["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": cores, "kind": "reference",
                                   "sample": "oracle/_ref/lda missing or failed on this box"}
            return out
        secs, ref_out = ref
        sample = (f"unmodified lda-c (oracle/_ref/lda, {cores} forked ranks), first {n_docs} docs ({nnz} nnz, {tok} tokens), "
                  f"EM iterations 1-{n_iters} from a 000 model both sides load")
        cb = {"value": tok / float(np.mean(secs)), "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": sample,
              "nnz_per_sec": nnz / float(np.mean(secs)), "sec_per_iter": [round(x, 3) for x in secs]}
        # the GPU path on the same documents, same files, same iterations
        c = oni.Corpus.from_csr(np.asarray(offsets[:n_docs + 1]), np.asarray(words[:nnz]), np.asarray(counts[:nnz]), num_terms=V)
        m = oni.Model.load(td / "000")
        st = oni.Settings.default()
        st.em_max_iter = n_iters - 1
        em = oni.EM(c, m, st, device=local_rank)
        liks, ms, sweeps, keep = [], [], [], True
        while keep:
            t0 = time.perf_counter()
            lik, _, keep = em.iterate()
            ms.append(1e3 * (time.perf_counter() - t0))
            liks.append(lik)
            sweeps.append(int(em.stats(with_sweeps=True)["sweeps"]))
        gamma, model = em.gamma(), em.model()
        em.close()
        cb["gpu_same_sample"] = {"value": tok / (np.mean(ms) * 1e-3), "unit": "tokens/s", "ms_per_iter": [round(x, 3) for x in ms],
                                 "variational_sweeps_per_iter": sweeps,
                                 "note": "same documents, same start files, same EM iterations as the reference run above; "
                                         "wall time of onilda_em_iterate, corpus resident"}
        out["cpu_baseline"] = cb
        ref_lik = np.loadtxt(ref_out / "likelihood.dat", ndmin=2)[:, 0]
        g_ref = np.loadtxt(ref_out / "final.gamma", ndmin=2)
        b_ref = np.loadtxt(ref_out / "final.beta", ndmin=2)
        a_ref = float((ref_out / "final.other").read_text().split()[-1])
        par = {"documents": n_docs, "iterations_reference": int(len(ref_lik)), "iterations_gpu": len(liks),
               "iterations_equal": len(ref_lik) == len(liks)}
        if par["iterations_equal"]:
            lik_err = np.abs(np.array(liks) - ref_lik) / np.abs(ref_lik)
            g_err = _rel(gamma, g_ref, 1e-3).max(axis=1)
            par.update({"lik_rel_err_per_iter": [float(x) for x in lik_err], "lik_rel_err_max": float(lik_err.max()),
                        "gamma_max_rel": float(g_err.max()), "gamma_docs_beyond_1e-4": int((g_err >= 1e-4).sum()),
                        "beta_max_rel": float(_rel(np.array(model.log_beta), b_ref, 1e-2).max()),
                        "alpha_rel": abs(model.alpha - a_ref) / a_ref,
                        "floors": "relative error with an absolute floor of 1e-3 (gamma) / 1e-2 (log beta), as in tests/test_gpu_parity.py; "
                                  "the reference's values carry the 10 decimals of its %5.10f files"})
            # top-1000 most suspicious flow events under each side's own theta / phi
            th_r, ph_r = theta_phi_numpy(g_ref, b_ref)
            th_g, ph_g = oni.theta_from_gamma(gamma, device=local_rank), oni.phi_from_log_beta(np.array(model.log_beta), device=local_rank)
            th_t = torch.from_numpy(th_r).to(device)
            pc = torch.cumsum(torch.from_numpy(ph_r).to(device).t().contiguous(), 1)
            pc = pc / pc[:, -1:]
            flat = (pc + torch.arange(K, device=device, dtype=torch.float64)[:, None]).reshape(-1)
            ev = [e.cpu().numpy() for e in gen_flow_events_torch(2_000_000, th_t, flat, n_docs, V, K, device)]
            del th_t, pc, flat
            want = flow_min_numpy(ev, th_r, ph_r)
            sc = oni.Scorer(th_g, ph_g, device=local_rank)
            mine = sc.score_flow_min(*ev)
            idx, _ = sc.topk(mine, 1.0, 1000)
            sc.close()
            order = np.lexsort((np.arange(len(want)), want))[:1000]
            cut = want[order[-1]] * (1 + 1e-9)
            finite = np.isfinite(want)
            par.update({"top1000_overlap": float(np.mean(want[idx] <= cut)),
                        "score_max_rel": float(np.max(np.abs(mine[finite] - want[finite]) / np.maximum(want[finite], 1e-12))),
                        "top1000_note": "tie-aware: share of the GPU's 1000 that the reference's own scores place at or below its "
                                        "1000th score; 2M flow events drawn from the reference's theta/phi"})
            par["within_bars"] = bool(par["lik_rel_err_max"] < 1e-6 and par["gamma_max_rel"] < 1e-4 and par["beta_max_rel"] < 1e-4
                                      and par["top1000_overlap"] >= 0.99)
        out["parity"] = par
    # north_star: "single process and mpiexec across all host cores": the single-process rate on a smaller prefix
    n1 = reference_sample_docs(offsets, 1, target_seconds=4.0)
    with tempfile.TemporaryDirectory() as td:
        one = run_reference(td, offsets, words, counts, n1, K, V, 2, 1, start)
    if one is not None:
        tok1 = int(np.asarray(counts[:int(offsets[n1])], np.int64).sum())
        out["cpu_baseline"]["single_process"] = {"value": tok1 / float(np.mean(one[0])), "unit": "tokens/s", "cores": 1,
                                                 "sample": f"first {n1} docs ({int(offsets[n1])} nnz, {tok1} tokens), EM iterations 1-2",
                                                 "sec_per_iter": [round(x, 3) for x in one[0]]}
    return out


def cfg1_leg(oni, local_rank):
    """BASELINE.json configs[0] (tests/golden/cfg1: 10k documents, V = 5k, K = 20, all 21 iterations, files written by
    the unmodified reference): the drop-in `lda est` on the GPU, wall time and the likelihood.dat it writes."""
    import gzip
    import shutil
    gold = ROOT / "tests" / "golden" / "cfg1"
    if not (gold / "c.dat.gz").exists():
        return None
    with tempfile.TemporaryDirectory() as td:
        td = Path(td)
        for name in ("c.dat", "000.beta"):
            with gzip.open(gold / (name + ".gz"), "rb") as f, open(td / name, "wb") as g:
                shutil.copyfileobj(f, g)
        shutil.copy(gold / "000.other", td / "000.other")
        shutil.copy(gold / "settings.txt", td / "settings.txt")
        env = dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", str(local_rank)))
        t0 = time.perf_counter()
        r = subprocess.run([str(oni.LDA_BIN), "est", "2.5", "20", "settings.txt", "1", "c.dat", "000", "out"], cwd=td, env=env,
                           capture_output=True, text=True)
        wall = time.perf_counter() - t0
        if r.returncode != 0:
            return {"error": r.stderr[-300:]}
        mine = np.loadtxt(td / "out" / "likelihood.dat", ndmin=2)[:, 0]
    ref = np.loadtxt(gold / "likelihood.dat", ndmin=2)[:, 0]
    res = {"workload": "10k docs, 1.1M tokens, V=5k, K=20, settings.txt defaults, from the fixture's 000", "iterations": int(len(mine)),
           "lda_est_wall_s": wall, "iterations_equal": len(mine) == len(ref),
           "reference_recorded": (gold / "README").read_text().strip()}
    if len(mine) == len(ref):
        res["lik_rel_err_max"] = float(np.max(np.abs(mine - ref) / np.abs(ref)))
    return res


# ----------------------------------------------------------------- north_star's split: one corpus over N GPUs

def strong_leg(oni, torch, dist, device, local_rank, rank, world, workload, steps, warmup, comm_id_fn, scale=1.0):
    """ONE corpus (the same on every rank: same seeds), contiguous document ranges balanced by entries
    (csrc/est.cu split_by_nnz, lda-estimate.c:201), E-step on the shard, the library's NCCL exchange + fused M-step.
    Timed on the device per iteration, max over ranks."""
    cfg = dict(WORKLOADS[workload])
    cfg["num_docs"] = max(1000, int(cfg["num_docs"] * scale))
    K, V = cfg["num_topics"], cfg["num_terms"]
    offsets, words, counts = gen_corpus_torch(device=device, doc_seed=cfg["seed"], **cfg)
    torch.cuda.empty_cache()
    corpus = oni.Corpus.from_csr(offsets, words, counts, num_terms=V)
    D, N, T = corpus.num_docs, corpus.nnz, corpus.tokens
    want = (N + D) * np.arange(world + 1) / world
    bounds = np.searchsorted(offsets + np.arange(D + 1), want).clip(0, D)
    bounds[0], bounds[-1] = 0, D
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    model0 = oni.Model.new(K, V)
    model0.init_random(20161102)
    model0.alpha = float(np.float32(ALPHA0))
    st = oni.Settings.default()
    st.em_max_iter = steps + warmup + 8
    em = oni.EM(corpus, model0, st, device=local_rank, doc_begin=lo, doc_end=hi)
    oni.comm_attach(em, comm_id_fn(), world, rank)
    liks, ms, est, rs, ag = [], [], [], [], []
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream = torch.cuda.ExternalStream(em.stream(), device=device)
    for it in range(warmup + steps):
        dist.barrier()
        torch.cuda.synchronize()
        ev0.record(stream)
        em.estep()
        lik = em.mstep()[0]
        ev1.record(stream)
        torch.cuda.synchronize()
        if it >= warmup:
            ms.append(ev0.elapsed_time(ev1))
            est.append(em.stats()["estep_ms"])
            cs = oni.comm_stats(em)
            rs.append(cs["reduce_scatter_ms"])
            ag.append(cs["all_gather_ms"])
        liks.append(lik)
    t = torch.tensor([float(np.mean(ms)), float(np.mean(est))], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, est_max = (float(x) for x in t.tolist())
    coll = oni.collective_time([em], repeats=10)
    # the longest document of the corpus, by itself on one warp (its serial word chain bounds any split)
    lens = np.diff(offsets)
    dmax = int(np.argmax(lens))
    lone = None
    if rank == 0:
        one = oni.EM(corpus, em.model(), st, device=local_rank, doc_begin=dmax, doc_end=dmax + 1)
        for _ in range(3):                 # the same E-step three times (the model is the corpus's, not refitted to one document)
            one.estep()
        one.mstep()                        # closes the timing of the last E-step
        lone = one.stats()["estep_ms"]
        lone_sweeps = int(one.stats(with_sweeps=True)["sweeps"])
        one.close()
    em.close()
    del corpus
    if rank != 0:
        return None
    P = world
    nb = coll["table_bytes"]
    b_em = algorithmic_bytes(N, D, K)
    return {"workload": workload, "num_docs": D, "nnz": N, "tokens": T, "num_terms": V, "num_topics": K, "n_gpus": P,
            "split": "contiguous document ranges balanced by entries (+1 per document)", "steps": steps, "warmup": warmup,
            "ms_per_iter": ms_max, "estep_ms_max_over_ranks": est_max, "tokens_per_sec": T / (ms_max * 1e-3),
            "nnz_per_sec": N / (ms_max * 1e-3), "likelihood_last": liks[-1],
            "roofline_frac": b_em / (ms_max * 1e-3) / 1e9 / (P * peak_hbm()[0]),
            "in_step_exchange_ms": {"reduce_scatter": float(np.mean(rs)), "all_gather_x3": float(np.mean(ag)),
                                    "note": "rank 0, CUDA events on the EM stream; includes waiting for the slowest rank's E-step"},
            "exchange_alone": {"reduce_scatter_ms": coll["reduce_scatter_ms"], "all_gather_x3_ms": coll["all_gather_ms"],
                               "table_bytes": nb,
                               "reduce_scatter_bus_gbs": (P - 1) / P * nb / (coll["reduce_scatter_ms"] * 1e-3) / 1e9,
                               "all_gather_bus_gbs": (P - 1) / P * (2 * nb + nb // (corpus_kpad(K))) / (coll["all_gather_ms"] * 1e-3) / 1e9,
                               "allreduce_equivalent_bus_gbs": 2 * (P - 1) / P * nb / ((coll["reduce_scatter_ms"] + coll["all_gather_ms"] / 2) * 1e-3) / 1e9,
                               "note": "NCCL over NVLink/NVSwitch, 10 back-to-back calls behind a barrier; bus bandwidth = (P-1)/P x bytes / time "
                                       "(reduce-scatter of the K x V fp64 statistics; all-gather of log_beta + beta' + row maxima); the "
                                       "all-reduce equivalent pairs the reduce-scatter with ONE table's all-gather"},
            "longest_document": {"entries": int(lens[dmax]), "sweeps": lone_sweeps, "lone_warp_estep_ms": lone,
                                 "note": "the document alone in an E-step: one warp, nothing else on the GPU; no split of the corpus "
                                         "over more GPUs ends an iteration earlier than this"}}


def corpus_kpad(K):
    g = 4 if K <= 32 else 8 if K <= 64 else 16 if K <= 128 else 32
    return g * ((K + g - 1) // g)


def peak_hbm():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text())["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs"
    return 6650.0, "fallback (B200_PROFILING.md)"


def sharded_scoring_leg(oni, torch, dist, device, local_rank, rank, world, gamma0, log_beta0, n_events, steps):
    """configs[4] at N GPUs: theta / phi replicated (rank 0's documents and the common model), the events sharded
    contiguously, each rank's 1000 most suspicious merged on rank 0 (no collective on the data path)."""
    D, K = gamma0.shape
    V = log_beta0.shape[1]
    theta = oni.theta_from_gamma(gamma0, device=local_rank)
    phi = oni.phi_from_log_beta(log_beta0, device=local_rank)
    per = n_events // world
    theta_t = torch.from_numpy(theta).to(device)
    pc = torch.cumsum(torch.from_numpy(phi).to(device).t().contiguous(), 1)
    pc = pc / pc[:, -1:]
    flat = (pc + torch.arange(K, device=device, dtype=torch.float64)[:, None]).reshape(-1)
    torch.manual_seed(20161105 + rank)
    ev = gen_flow_events_torch(per, theta_t, flat, D, V, K, device)
    del theta_t, pc, flat
    torch.cuda.empty_cache()
    sc = oni.Scorer(theta, phi, device=local_rank)
    scores = torch.empty(per, dtype=torch.float64, device=device)
    ts = []
    for it in range(2 + steps):
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        idx, val = sc.score_flow_topk_device(ev[0].data_ptr(), ev[1].data_ptr(), ev[2].data_ptr(), ev[3].data_ptr(), per, 1.0,
                                             1000, scores.data_ptr())
        pad = 1000 - len(idx)
        mine = torch.tensor(np.concatenate([val, np.full(pad, np.inf)]), device=device)
        mine_i = torch.tensor(np.concatenate([idx + rank * per, np.full(pad, -1)]), device=device)
        allv = [torch.empty_like(mine) for _ in range(world)]
        alli = [torch.empty_like(mine_i) for _ in range(world)]
        dist.all_gather(allv, mine)
        dist.all_gather(alli, mine_i)
        if rank == 0:
            v, i = torch.cat(allv).cpu().numpy(), torch.cat(alli).cpu().numpy()
            order = np.lexsort((i, v))[:1000]
            top = (i[order], v[order])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if it >= 2:
            ts.append(dt)
    t = torch.tensor([float(np.mean(ts)), sc.stats()["kernel_ms"]], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sc.close()
    if rank != 0:
        return None
    sec, kms = float(t[0]), float(t[1])
    return {"events": per * world, "n_gpus": world, "events_per_sec": per * world / sec, "unit": "flow events/s",
            "ms": 1e3 * sec, "score_kernel_ms_max_over_ranks": kms, "most_suspicious_score": float(top[1][0]),
            "note": "events sharded contiguously, theta/phi replicated, per-rank top-1000 gathered and merged on rank 0; "
                    "wall time from barrier to merged rows, max over ranks"}


# ----------------------------------------------------------------- main

def algorithmic_bytes(nnz, num_docs, K):
    """B_EM = N(8 + 16K) + D(16 + 8K) per EM iteration."""
    return nnz * (8 + 16 * K) + num_docs * (16 + 8 * K)


_REAL_STDOUT = None


def emit(line):
    """The one JSON line goes to the process's real stdout."""
    data = (json.dumps(line, default=float) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    # Libraries (NCCL's version banner, for one) write to file descriptor 1.  stdout must carry exactly one
    # JSON line, so fd 1 is pointed at stderr for the duration and the line is written to a saved copy.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=18)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-double-buffer", action="store_true", help="skip the double-buffered variant of the end-to-end leg")
    ap.add_argument("--scale", type=float, default=1.0, help="development only: shrink the number of documents")
    ap.add_argument("--hub-frac", type=float, default=None, help="development only: fraction of hub documents")
    ap.add_argument("--no-e2e", action="store_true", help="development only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-scoring", action="store_true")
    ap.add_argument("--no-grouping", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the one-corpus-over-N-GPUs legs at N > 1")
    ap.add_argument("--events", type=int, default=1_000_000_000, help="flow events of the scoring leg (configs[4])")
    args = ap.parse_args()
    if args.warmup < 3:
        log("warm-up raised to 3 (timing rules)")
        args.warmup = 3

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return 0

    import torch
    cfg = dict(WORKLOADS[args.workload])
    cfg["num_docs"] = max(1000, int(cfg["num_docs"] * args.scale))
    K, V = cfg["num_topics"], cfg["num_terms"]
    if args.hub_frac is not None:
        cfg["hub_frac"] = args.hub_frac
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the synthetic corpus is generated there; libonilda has no CPU path)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)

    # ------------------------------------------------------------ reference arm: never loads libonilda
    if args.impl == "reference":
        t0 = time.time()
        offsets, words, counts = gen_corpus_torch(device=device, doc_seed=cfg["seed"], **cfg)
        log(f"corpus: D={len(offsets) - 1} N={int(offsets[-1])} ({time.time() - t0:.1f}s)")
        cores = host_cores()
        n_docs = reference_sample_docs(offsets, cores)
        n_iters = args.steps + args.warmup
        # as in the other arm: a run is the reference's own (settings.txt: em max iter 20 -> 21 iterations); steps beyond it
        # are the first iterations of another run from the same start model
        run_len = 21
        secs = []
        while len(secs) < n_iters:
            with tempfile.TemporaryDirectory() as td:
                ref = run_reference(td, offsets, words, counts, n_docs, K, V, min(run_len, n_iters - len(secs)), cores,
                                    start_model_numpy(K, V, 20161102))
            if ref is None:
                emit({"impl": "reference", "unavailable": "oracle/_ref/lda missing or failed on this box"})
                return 0
            secs += ref[0]
        timed = secs[args.warmup:] if len(secs) > args.warmup else secs
        nnz = int(offsets[n_docs])
        tok = int(np.asarray(counts[:nnz], np.int64).sum())
        val = tok / float(np.mean(timed))
        sample = (f"first {n_docs} docs ({nnz} nnz, {tok} tokens) of the workload, steps {len(secs) - len(timed) + 1}-{len(secs)} "
                  f"timed ({len(timed)}; a run has {run_len} EM iterations, the steps after it start a new run), "
                  f"{cores} forked ranks of the unmodified lda-c")
        emit({"metric": "lda_em_tokens_per_sec_per_iter", "value": val, "unit": "tokens/s", "n_gpus": 0,
              "steps": len(timed), "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(timed)),
              "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
              "impl": "reference",
              "config": {"workload": args.workload, "sample": sample, "nnz_per_sec": nnz / float(np.mean(timed)),
                         "inputs": "sample.dat and 000.* written with numpy; libonilda is not loaded by this process; the "
                                   "synthetic corpus is generated with PyTorch on the GPU, as in the other arm",
                         "sec_per_iter": [round(x, 3) for x in secs]},
              "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": sample},
              "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
              "gpu_launches": 0})
        return 0

    oni = importlib.import_module("open-network-insight_b200")
    if oni.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: libonilda has no CPU path")
    dist = None
    host_cpus = bind_near_gpu(local_rank) if world > 1 else 0
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")    # stdout carries the one JSON line only
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)

    def comm_id():
        """128 bytes of NCCL unique id from rank 0 to everyone (torch.distributed is the messenger only)."""
        buf = torch.zeros(oni.COMM_ID_BYTES, dtype=torch.uint8, device=device)
        if rank == 0:
            buf.copy_(torch.frombuffer(bytearray(oni.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        return buf.cpu().numpy().tobytes()

    t0 = time.time()
    offsets, words, counts = gen_corpus_torch(device=device, doc_seed=cfg["seed"] + 1000 * rank, **cfg)
    torch.cuda.empty_cache()
    corpus = oni.Corpus.from_csr(offsets, words, counts, num_terms=V)
    D, N, T = corpus.num_docs, corpus.nnz, corpus.tokens
    log(f"[rank {rank}] corpus: D={D} N={N} T={T} V={corpus.num_terms} max_len={corpus.max_len} ({time.time() - t0:.1f}s)")
    model0 = oni.Model.new(K, V)
    model0.init_random(20161102)
    model0.alpha = float(np.float32(ALPHA0))
    # the reference's settings.txt (em max iter 20: run_em stops after 21 iterations, lda-estimate.c:186).  A run of the
    # bench ends by that rule too; the step after it is iteration 1 of a new run from the same start model.  (Letting
    # one run go on to iteration 25, as rounds before did, times iterations the reference never executes - and they are
    # the heaviest: by then alpha has fallen below 0.3 and ordinary documents take 9-12 sweeps instead of 3.)
    settings = oni.Settings.default()

    # ------------------------------------------------------------ resident-in-HBM leg
    em = oni.EM(corpus, model0, settings, device=local_rank)
    if dist is not None:
        oni.comm_attach(em, comm_id(), world, rank)      # the exchange is the library's: reduce-scatter, lda_mle, all-gather

    run_lengths = []            # EM iterations of each finished run
    in_run = [0]

    def em_step():
        em.estep()
        out = em.mstep()
        in_run[0] += 1
        if not out[2]:          # run_em's loop condition no longer holds: the next step starts a new run
            run_lengths.append(in_run[0])
            in_run[0] = 0
            em.reload_model(model0)
        return out

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    liks = []
    for _ in range(args.warmup):
        liks.append(em_step()[0])
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    t_begin = time.perf_counter()
    estep_ms, mstep_ms, launches, word_steps = [], [], 0, []
    for _ in range(args.steps):
        liks.append(em_step()[0])
        s = em.stats()
        word_steps.append(em.word_steps())
        estep_ms.append(s["estep_ms"])
        mstep_ms.append(s["mstep_ms"])
        launches += s["launches"]
    barrier()
    elapsed = time.perf_counter() - t_begin
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([elapsed], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed = float(t.item())
        tt = torch.tensor([float(T), float(N), float(D)], device=device, dtype=torch.float64)
        dist.all_reduce(tt)
        T_all, N_all, D_all = (int(x) for x in tt.tolist())
    else:
        T_all, N_all, D_all = T, N, D
    ms_per_step = 1e3 * elapsed / args.steps
    value = T_all / (elapsed / args.steps)
    sweeps = em.stats(with_sweeps=True)["sweeps"]

    # ------------------------------------------------------------ e2e leg: host buffers through the C ABI
    corpus.pin()
    g_host = torch.empty((D, K), dtype=torch.float64, pin_memory=True).numpy()
    cur = oni.Model.new(K, V)
    cur.log_beta[...] = model0.log_beta
    cur.alpha = model0.alpha
    lib = oni.lib()

    # the end-to-end legs reset the handle's loop state at every step (reload), so they count the run's iterations here
    run_len = run_lengths[0] if run_lengths else settings.em_max_iter + 1
    e2e_in_run = [0]

    def next_model():
        e2e_in_run[0] += 1
        if e2e_in_run[0] >= run_len:          # the run is over: the next step starts from the start model again
            e2e_in_run[0] = 0
            cur.log_beta[...] = model0.log_beta
            cur.alpha = model0.alpha

    def e2e_step():
        em.reload(cur)                        # H2D: CSR words+counts+totals, model
        em.estep()
        out = em.mstep()                      # exchange (N > 1), M-step; D2H: likelihood + statistics tail
        oni._check(lib.onilda_em_get_model(em._h, cur._h))      # D2H: new model
        oni._check(lib.onilda_em_get_gamma(em._h, g_host))      # D2H: gamma
        next_model()
        return out

    e2e_liks = [float("nan")]
    e2e_elapsed = float("nan")
    if not args.no_e2e:
        for _ in range(args.warmup):
            e2e_liks.append(e2e_step()[0])
        barrier()
        t_begin = time.perf_counter()
        for _ in range(args.steps):
            e2e_liks.append(e2e_step()[0])
        barrier()
        e2e_elapsed = time.perf_counter() - t_begin
    if dist is not None:
        t = torch.tensor([e2e_elapsed], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_elapsed = float(t.item())
    h2d = N * 8 + D * 4 + K * V * 8
    d2h = K * V * 8 + D * K * 8 + 32 + 4
    e2e = {"value": T_all / (e2e_elapsed / args.steps), "unit": "tokens/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * e2e_elapsed / args.steps}
    if world > 1:
        e2e["host_cores_bound_per_rank"] = host_cpus      # bind_near_gpu: the cores NVML lists for the rank's GPU (0: not bound)

    # The same step with the NEXT step's corpus copy under way while this one computes (what a caller with a stream of
    # corpora in host memory does): two handles, each with its own stream and device buffers; the CSR arrays of step
    # i + 1 go up from a second host thread during step i (onilda_em_reload_corpus), the model, which step i produces,
    # follows when it is there (onilda_em_reload_model).  Every step moves the same bytes in and out, inside the timed
    # region.  This is the end-to-end value; the one-call-at-a-time figure stays beside it as e2e.serial.
    if not args.no_e2e and not args.no_double_buffer:
        import threading
        em2 = None
        try:
            em2 = oni.EM(corpus, model0, settings, device=local_rank)
            if dist is not None:
                oni.comm_attach(em2, comm_id(), world, rank)
        except oni.OnildaError as e:          # no room for a second set of device buffers: the serial figure stands
            log(f"[rank {rank}] double-buffered leg skipped: {e}")
            em2 = None
        ok = torch.tensor([1.0 if em2 is not None else 0.0], device=device, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok.item()) > 0:
            pair = [em, em2]
            cur.log_beta[...] = model0.log_beta
            cur.alpha = model0.alpha
            e2e_in_run[0] = 0
            pair[0].reload_corpus()

            def db_step(i):
                x, y = pair[i % 2], pair[(i + 1) % 2]
                x.reload_model(cur)               # H2D: model (before the big copy is queued: copies of one direction run in order)
                th = threading.Thread(target=y.reload_corpus)      # H2D: the next step's CSR words + counts + totals
                th.start()
                x.estep()
                out = x.mstep()
                oni._check(lib.onilda_em_get_model(x._h, cur._h))      # D2H: new model
                oni._check(lib.onilda_em_get_gamma(x._h, g_host))      # D2H: gamma
                next_model()
                th.join()
                return out

            db_liks = [db_step(i)[0] for i in range(args.warmup)]
            barrier()
            t_begin = time.perf_counter()
            for i in range(args.warmup, args.warmup + args.steps):
                db_liks.append(db_step(i)[0])
            barrier()
            db_elapsed = time.perf_counter() - t_begin
            if dist is not None:
                t = torch.tensor([db_elapsed], device=device, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                db_elapsed = float(t.item())
            e2e["serial"] = {"value": e2e["value"], "ms_per_step": e2e["ms_per_step"],
                             "note": "reload -> estep -> mstep -> get_model -> get_gamma, one call after the other on one handle, nothing overlapped"}
            e2e["value"] = T_all / (db_elapsed / args.steps)
            e2e["ms_per_step"] = 1e3 * db_elapsed / args.steps
            e2e["likelihoods_equal_to_serial"] = bool(np.allclose(db_liks, e2e_liks[1:], rtol=1e-9, atol=0.0))
            e2e["note"] = ("double-buffered: two EM handles; the CSR arrays of step i + 1 are copied host -> device "
                           "(onilda_em_reload_corpus, second host thread, the other handle's stream) while step i computes; "
                           "every step copies h2d_bytes_per_step in and d2h_bytes_per_step out inside the timed region")
        if em2 is not None:
            em2.close()

    gamma = em.gamma()
    log_beta = np.array(em.model().log_beta)
    exchange = oni.comm_stats(em) if dist is not None else None
    em.close()
    torch.cuda.empty_cache()

    # ------------------------------------------------------------ N > 1: north_star's split and the sharded scoring sweep
    strong = strong_dns = strong_proxy = sharded = None
    if dist is not None and not args.no_strong:
        try:
            strong = strong_leg(oni, torch, dist, device, local_rank, rank, world, DEFAULT_WORKLOAD, 8, 3, comm_id, args.scale)
            strong_dns = strong_leg(oni, torch, dist, device, local_rank, rank, world, "dns-2M-docs-200M-tokens-V50k-K20", 6, 3,
                                    comm_id, args.scale)
            # configs[3] (K = 50, V = 500k: a 224 MB statistics table, the exchange's heaviest case), documents scaled to a
            # modest share of what the N GPUs hold: 5 % of the 5 M documents per GPU
            strong_proxy = strong_leg(oni, torch, dist, device, local_rank, rank, world, "proxy-5M-docs-1B-tokens-V500k-K50", 4, 3,
                                      comm_id, args.scale * 0.05 * world)
        except Exception as e:          # the headline line must still be printed
            log(f"[rank {rank}] strong leg failed: {e!r}")
            strong = strong or {"error": repr(e)}
        if not args.no_scoring:
            try:
                g0 = torch.from_numpy(gamma).to(device)
                b0 = torch.from_numpy(log_beta).to(device)
                dist.broadcast(g0, 0)
                dist.broadcast(b0, 0)
                sharded = sharded_scoring_leg(oni, torch, dist, device, local_rank, rank, world, g0.cpu().numpy(), b0.cpu().numpy(),
                                              max(1_000_000, int(args.events * args.scale)), 5)
                del g0, b0
            except Exception as e:
                log(f"[rank {rank}] sharded scoring failed: {e!r}")
                sharded = {"error": repr(e)}

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ------------------------------------------------------------ roofline of the E-step
    peak, peak_src = peak_hbm()
    b_em = algorithmic_bytes(N, D, K)
    achieved = b_em / (np.mean(estep_ms) * 1e-3) / 1e9
    traffic = fp64 = None
    prof = ROOT / "profiles" / "estep_traffic.json"
    if prof.exists():
        try:
            pj = json.loads(prof.read_text())
            traffic, fp64 = pj.get("dram_bytes_per_launch"), pj.get("fp64_pipe_pct_of_active_cycles")
        except ValueError:
            pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": "E-step = estep_long_kernel<20> (documents >= 1024 entries, one per warp) || "
                                              "estep_kernel<4,5,20> (the rest, eight per warp), two concurrent launches timed as one",
                "algorithmic_bytes_per_launch": b_em, "kernel_ms": float(np.mean(estep_ms)), "peak_source": peak_src,
                "fp64_pipe_pct_of_active_cycles": fp64,
                "note": "nominally an HBM gather; measured bound is instruction issue on the serial per-word dependency chain "
                        "(Gauss-Seidel order of lda-inference.c:50-75) of the long documents, so frac against HBM is small by "
                        "construction: DESIGN.md 4.1"}

    # ------------------------------------------------------------ single-GPU side legs (rank 0 at N = 1)
    scoring = grouping = cfg1 = None
    cpu_baseline = parity = None
    if world == 1:
        if not args.no_scoring:
            n_events = max(1_000_000, int(args.events * args.scale))
            scoring = scoring_leg(oni, torch, device, gamma, log_beta, n_events, 1 << 28, 5, peak,
                                  cpu_cores=0 if args.no_cpu_baseline else host_cores())
        torch.cuda.empty_cache()
        if not args.no_grouping:
            _, w_all, c_all, _ = corpus.csr()
            grouping = grouping_leg(oni, torch, device, np.asarray(offsets), w_all, c_all, 5, peak, cpu=not args.no_cpu_baseline)
        if not args.no_cpu_baseline:
            pl = parity_leg(oni, torch, device, offsets, words, counts, K, V, host_cores(), local_rank)
            cpu_baseline, parity = pl["cpu_baseline"], pl["parity"]
            cfg1 = cfg1_leg(oni, local_rank)
    else:
        scoring = {"sharded": sharded} if sharded is not None else None
    del gamma, log_beta

    line = {"metric": "lda_em_tokens_per_sec_per_iter", "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "num_docs": D_all, "nnz": N_all, "tokens": T_all, "num_terms": V,
                       "num_topics": K, "docs_per_gpu": D, "settings": "var 20/1e-6, em 20/1e-6, alpha estimate",
                       "start": "random (MT19937 seed 20161102), alpha 2.5",
                       "l2": "inputs larger than L2 (CSR 0.7 GB + phi scratch 14 GB per GPU)",
                       "parallelism": f"dp{world} over documents; statistics: NCCL reduce-scatter -> lda_mle of the own rows -> all-gather "
                                      "inside libonilda" if world > 1 else "one GPU", "scale": args.scale},
            "nnz_per_sec": N_all / (elapsed / args.steps),
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu_baseline, "parity": parity, "cfg1": cfg1, "scoring": scoring, "grouping": grouping,
            "strong": strong, "strong_dns": strong_dns, "strong_proxy": strong_proxy, "exchange_in_step": exchange,
            "em": {"estep_ms": [round(x, 3) for x in estep_ms], "mstep_ms_mean": float(np.mean(mstep_ms)),
                   "likelihood_first": liks[0], "likelihood_last": liks[-1], "sweeps_last_iter": int(sweeps),
                   "e2e_likelihood_last": e2e_liks[-1],
                   "word_steps_per_iter": [int(x) for x in word_steps],
                   "runs_finished_in_resident_leg": run_lengths,
                   "estep_ms_note": "steps warmup+1 .. warmup+steps.  A run ends by the reference's own rule (settings.txt: em max "
                                    "iter 20, i.e. 21 iterations, or convergence) and the next step is iteration 1 of a new run "
                                    "from the same start model, so with steps + warmup > 21 the list wraps around.  "
                                    "word_steps_per_iter = sum over documents of entries x variational sweeps (all ranks), the "
                                    "E-step's unit of work: it grows as the topics sharpen and more documents run towards the "
                                    "20-sweep cap, and estep_ms follows it"}}
    emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
