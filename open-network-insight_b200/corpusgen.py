"""Seeded synthetic per-IP "document" corpora and flow-event tables with the
shapes BASELINE.json names (there is no network for real netflow/DNS/proxy data).

The generative recipe is LDA itself, so EM has structure to find:
  tokens per doc  n_d = clip(lognormal(mu, 0.8), 5, max_tokens), mu set for the mean
  theta_d ~ Dir(0.1 * 1_K*),   beta*_k ~ Dir(0.05 * 1_V)
  word ids ~ theta_d^T beta*; entries = unique ids with counts, ascending id
  dup_frac of the docs get one entry split in two (duplicate word ids in a doc are
      legal input: netflow/FlowPreLDA.scala:139 unions src and dst counts)
  hub_frac of the docs are "hub" IPs with hub_tokens tokens (long-document path)

The corpus is what `OniLDACWrapper.createModel` writes to model.dat
(OniLDACWrapper.scala:146-168): line d = "<#entries> <word>:<count> ...".
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class Corpus:
    offsets: np.ndarray   # int64 [D+1]
    words: np.ndarray     # int32 [N]
    counts: np.ndarray    # int32 [N]
    num_terms: int        # max word id + 1 (lda-data.c:49,55)

    @property
    def num_docs(self) -> int:
        return len(self.offsets) - 1

    @property
    def nnz(self) -> int:
        return int(self.offsets[-1])

    @property
    def tokens(self) -> int:
        return int(self.counts.sum(dtype=np.int64))

    @property
    def max_len(self) -> int:
        return int(np.diff(self.offsets).max()) if self.num_docs else 0


# name -> (D, mean tokens, V, K, seed); SURVEY.md section 8(d)
CONFIGS = {
    "cfg1": dict(num_docs=10_000, mean_tokens=100, num_terms=5_000, num_topics=20, seed=20161101),
    "cfg2": dict(num_docs=1_000_000, mean_tokens=100, num_terms=100_000, num_topics=20, seed=20161102),
    "cfg3": dict(num_docs=2_000_000, mean_tokens=100, num_terms=50_000, num_topics=20, seed=20161103),
    "cfg4": dict(num_docs=5_000_000, mean_tokens=200, num_terms=500_000, num_topics=50, seed=20161104),
}


def _doc_token_counts(rng, num_docs, mean_tokens, max_tokens, hub_frac, hub_tokens):
    sigma = 0.8
    mu = np.log(mean_tokens) - 0.5 * sigma * sigma
    n = np.clip(rng.lognormal(mu, sigma, num_docs), 5, max_tokens).astype(np.int64)
    n_hub = int(round(hub_frac * num_docs))
    if n_hub > 0:
        hubs = rng.choice(num_docs, n_hub, replace=False)
        n[hubs] = rng.integers(hub_tokens // 4, hub_tokens + 1, n_hub)
    return n


def generate(num_docs, mean_tokens, num_terms, num_topics, seed, max_tokens=20_000, hub_frac=0.001,
             hub_tokens=20_000, dup_frac=0.01, chunk_docs=200_000) -> Corpus:
    rng = np.random.default_rng(seed)
    K, V = num_topics, num_terms
    beta = rng.dirichlet(np.full(V, 0.05), K)                 # [K][V]
    cdf = np.cumsum(beta, axis=1)
    cdf /= cdf[:, -1:]
    n_tok = _doc_token_counts(rng, num_docs, mean_tokens, max_tokens, hub_frac, hub_tokens)

    lens, words_out, counts_out = [], [], []
    for lo in range(0, num_docs, chunk_docs):
        hi = min(num_docs, lo + chunk_docs)
        nd = hi - lo
        theta = rng.dirichlet(np.full(K, 0.1), nd)
        per_topic = rng.multinomial(n_tok[lo:hi], theta)       # [nd][K]
        doc_of, word_of = [], []
        for k in range(K):
            m = int(per_topic[:, k].sum())
            if m == 0:
                continue
            w = np.searchsorted(cdf[k], rng.random(m), side="right")
            np.minimum(w, V - 1, out=w)
            doc_of.append(np.repeat(np.arange(nd, dtype=np.int64), per_topic[:, k]))
            word_of.append(w.astype(np.int64))
        key = np.concatenate(doc_of) * V + np.concatenate(word_of)
        uniq, cnt = np.unique(key, return_counts=True)
        d_idx = uniq // V
        lens.append(np.bincount(d_idx, minlength=nd).astype(np.int64))
        words_out.append((uniq % V).astype(np.int32))
        counts_out.append(cnt.astype(np.int32))
    lens = np.concatenate(lens)
    words = np.concatenate(words_out)
    counts = np.concatenate(counts_out)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)

    offsets, words, counts = add_duplicates(offsets, words, counts, rng, dup_frac)
    return Corpus(offsets, words.astype(np.int32), counts.astype(np.int32), int(words.max()) + 1 if len(words) else 0)


def add_duplicates(offsets, words, counts, rng, dup_frac):
    """Duplicate word ids inside a document: in dup_frac of the documents one entry with
    count >= 2 is split in two, the second half appended at the end of the document."""
    num_docs = len(offsets) - 1
    n_dup = int(round(dup_frac * num_docs))
    if n_dup <= 0:
        return offsets, words, counts
    docs = np.sort(rng.choice(num_docs, n_dup, replace=False))
    ins_pos, ins_word, ins_cnt, ins_doc = [], [], [], []
    for d in docs:
        a, b = offsets[d], offsets[d + 1]
        cand = np.nonzero(counts[a:b] >= 2)[0]
        if len(cand) == 0:
            continue
        e = a + cand[rng.integers(len(cand))]
        take = counts[e] // 2
        counts[e] -= take
        ins_pos.append(b)
        ins_word.append(words[e])
        ins_cnt.append(take)
        ins_doc.append(d)
    if ins_pos:
        ins_pos = np.asarray(ins_pos)
        words = np.insert(words, ins_pos, np.asarray(ins_word, np.int32))
        counts = np.insert(counts, ins_pos, np.asarray(ins_cnt, np.int32))
        add = np.zeros(num_docs + 1, np.int64)
        np.add.at(add, np.asarray(ins_doc) + 1, 1)
        offsets = offsets + np.cumsum(add)
    return offsets, words, counts


def generate_named(name: str, scale: float = 1.0, **overrides) -> Corpus:
    cfg = dict(CONFIGS[name])
    cfg["num_docs"] = max(1, int(cfg["num_docs"] * scale))
    cfg.update(overrides)
    return generate(**cfg)


def write_dat(corpus: Corpus, path) -> None:
    """model.dat as OniLDACWrapper.createModel emits it (OniLDACWrapper.scala:146-168)."""
    off, w, c = corpus.offsets, corpus.words, corpus.counts
    with open(path, "w") as f:
        for d in range(corpus.num_docs):
            a, b = off[d], off[d + 1]
            f.write(str(b - a))
            if b > a:
                f.write(" ")
                f.write(" ".join(f"{x}:{y}" for x, y in zip(w[a:b].tolist(), c[a:b].tolist())))
            f.write("\n")


def generate_flow_events(num_events, theta, phi_wk, seed, unknown_frac=0.001):
    """Flow events (src_doc, src_word, dst_doc, dst_word) int32, docs Zipf(1.1) over D,
    words drawn from the doc's theta^T phi (cfg 5).  unknown_frac of the ids are -1
    (unseen ip / word => uniform 1/K, SuspiciousConnectsScoreFunction.scala:13-18)."""
    rng = np.random.default_rng(seed)
    D, K = theta.shape
    V = phi_wk.shape[0]
    rank = np.arange(1, D + 1, dtype=np.float64) ** -1.1
    doc_cdf = np.cumsum(rank)
    doc_cdf /= doc_cdf[-1]
    topic_cdf = np.cumsum(phi_wk, axis=0)
    topic_cdf /= topic_cdf[-1:, :]

    def side():
        doc = np.minimum(np.searchsorted(doc_cdf, rng.random(num_events)), D - 1)
        th = np.maximum(theta[doc], 0)
        tc = np.cumsum(th, axis=1)
        tc /= np.maximum(tc[:, -1:], 1e-300)
        z = (rng.random(num_events)[:, None] > tc).sum(axis=1).clip(0, K - 1)
        u = rng.random(num_events)
        word = np.empty(num_events, np.int64)
        for k in range(K):
            sel = np.nonzero(z == k)[0]
            word[sel] = np.minimum(np.searchsorted(topic_cdf[:, k], u[sel]), V - 1)
        doc = doc.astype(np.int32)
        word = word.astype(np.int32)
        doc[rng.random(num_events) < unknown_frac] = -1
        word[rng.random(num_events) < unknown_frac] = -1
        return doc, word

    sd, sw = side()
    dd, dw = side()
    return sd, sw, dd, dw
