"""open-network-insight_b200: B200-native LDA + suspicious-connects scoring.

This module is a thin ctypes view of the C ABI in include/onilda.h
(lib/libonilda.so: hand-written sm_100a kernels behind `extern "C"` entry
points).  It exists for the tests and the benchmark; production callers use the
`lda` program or bind the C ABI directly.

There is no CPU fallback: importing works anywhere, but every compute call
raises OnildaError on a machine without a usable CUDA device, and loading fails
loudly if the shared library has not been built (`make -C open-network-insight_b200`).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "lib" / "libonilda.so"
LDA_BIN = HERE / "bin" / "lda"


class OnildaError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"onilda error {code}: {message}")
        self.code = code


class Settings(C.Structure):
    """The five values of lda-c's settings file (lda-estimate.c:331-350)."""
    _fields_ = [("var_max_iter", C.c_int), ("var_converged", C.c_float),
                ("em_max_iter", C.c_int), ("em_converged", C.c_float),
                ("estimate_alpha", C.c_int)]

    @classmethod
    def default(cls):   # oni-ml/oni-lda-c/settings.txt
        return cls(20, 1e-6, 20, 1e-6, 1)


def build(verbose: bool = False) -> None:
    """Compile libonilda.so and bin/lda for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", str(HERE), "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libonilda.so failed")


_lib = None
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_vp = C.c_void_p


def lib():
    """The loaded C ABI.  Raises if lib/libonilda.so is missing: no silent fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise OnildaError(-3, f"{LIB_PATH} is not built; run `make -C {HERE}` (or __graft_entry__.build())")
    L = C.CDLL(str(LIB_PATH))
    L.onilda_last_error.restype = C.c_char_p
    L.onilda_version.restype = C.c_char_p
    L.onilda_device_count.restype = C.c_int
    L.onilda_settings_read.argtypes = [C.c_char_p, C.POINTER(Settings)]
    L.onilda_corpus_from_dat.argtypes = [C.c_char_p, C.POINTER(_vp)]
    L.onilda_corpus_from_csr.argtypes = [_i64p, _vp, _vp, C.c_int64, C.c_int32, C.POINTER(_vp)]
    L.onilda_corpus_from_triples.argtypes = [_vp, _vp, _vp, C.c_int64, C.c_int64, C.c_int32, C.c_int, C.POINTER(_vp)]
    L.onilda_flow_prelda.argtypes = [_vp] * 9 + [C.c_int64, C.c_int64, C.c_int, _vp, C.POINTER(_vp), _vp, C.c_int32, C.POINTER(C.c_int32), _vp, _vp]
    L.onilda_dns_words.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, C.c_int64, C.c_int, _vp, _vp]
    L.onilda_dns_words_features.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, C.c_int64, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
    L.onilda_dns_word_string.argtypes = [C.c_int64, C.c_char_p, C.c_size_t]
    L.onilda_flow_results.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int64, C.c_double, C.c_int64, _vp, _vp, _vp, C.POINTER(C.c_int64)]
    L.onilda_quantiles_i64.argtypes = [_vp, C.c_int64, _vp, C.c_int, C.c_int, _vp]
    L.onilda_quantiles_i64_device.argtypes = [_vp, C.c_int64, _vp, C.c_int, _vp, _vp]
    L.onilda_flow_words.argtypes = [_vp] * 7 + [C.c_int64, C.c_int, _vp, _vp, _vp]
    L.onilda_flow_words_device.argtypes = [_vp] * 7 + [C.c_int64, _vp, _vp, _vp, _vp]
    L.onilda_flow_word_string.argtypes = [C.c_int32, C.c_char_p, C.c_size_t]
    L.onilda_pairs_count.argtypes = [_vp, _vp, C.c_int64, C.c_int64, C.c_int32, C.c_int, _vp, _vp, _vp, C.POINTER(C.c_int64)]
    L.onilda_pairs_count_device.argtypes = [_vp, _vp, C.c_int64, C.c_int64, C.c_int32, _vp, _vp, _vp, _vp, _vp]
    L.onilda_triples_group_device.argtypes = [_vp, _vp, _vp, C.c_int64, C.c_int64, _vp, _vp, _vp, _vp, _vp, _vp]
    L.onilda_corpus_dims.argtypes = [_vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int32),
                                     C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
    L.onilda_corpus_csr.argtypes = [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp)]
    L.onilda_corpus_pin.argtypes = [_vp]
    L.onilda_corpus_save_dat.argtypes = [_vp, C.c_int64, C.c_int64, C.c_char_p]
    L.onilda_corpus_free.argtypes = [_vp]
    L.onilda_corpus_free.restype = None
    L.onilda_model_new.argtypes = [C.c_int, C.c_int, C.POINTER(_vp)]
    L.onilda_model_dims.argtypes = [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.onilda_model_alpha.argtypes = [_vp]
    L.onilda_model_alpha.restype = C.c_double
    L.onilda_model_set_alpha.argtypes = [_vp, C.c_double]
    L.onilda_model_log_beta.argtypes = [_vp]
    L.onilda_model_log_beta.restype = C.POINTER(C.c_double)
    L.onilda_model_init_random.argtypes = [_vp, C.c_uint32]
    L.onilda_model_init_seeded.argtypes = [_vp, _vp, C.c_uint32]
    L.onilda_model_load.argtypes = [C.c_char_p, C.POINTER(_vp)]
    L.onilda_model_save.argtypes = [_vp, C.c_char_p]
    L.onilda_model_free.argtypes = [_vp]
    L.onilda_model_free.restype = None
    L.onilda_gamma_save.argtypes = [C.c_char_p, _f64p, C.c_int64, C.c_int]
    L.onilda_em_create.argtypes = [_vp, C.c_int64, C.c_int64, _vp, C.POINTER(Settings), C.c_int, C.POINTER(_vp)]
    L.onilda_em_destroy.argtypes = [_vp]
    L.onilda_em_destroy.restype = None
    L.onilda_em_estep.argtypes = [_vp]
    L.onilda_em_reload.argtypes = [_vp, _vp]
    L.onilda_em_reload_corpus.argtypes = [_vp]
    L.onilda_em_reload_model.argtypes = [_vp, _vp]
    L.onilda_em_suffstats_device.argtypes = [_vp, C.POINTER(_vp), C.POINTER(C.c_int64)]
    L.onilda_em_mstep.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int)]
    L.onilda_em_iterate.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int)]
    L.onilda_em_allreduce_peer.argtypes = [C.POINTER(_vp), C.c_int]
    L.onilda_comm_unique_id.argtypes = [C.c_char_p]
    L.onilda_em_comm_attach.argtypes = [_vp, C.c_char_p, C.c_int, C.c_int]
    L.onilda_em_comm_attach_local.argtypes = [C.POINTER(_vp), C.c_int]
    L.onilda_em_mstep_group.argtypes = [C.POINTER(_vp), C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int)]
    L.onilda_em_comm_stats.argtypes = [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.onilda_em_collective_time.argtypes = [C.POINTER(_vp), C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.onilda_em_get_model.argtypes = [_vp, _vp]
    L.onilda_em_get_gamma.argtypes = [_vp, _f64p]
    L.onilda_em_get_doc_likelihood.argtypes = [_vp, _f64p]
    L.onilda_em_get_doc_sweeps.argtypes = [_vp, _i32p]
    L.onilda_em_stats.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int64),
                                  C.POINTER(C.c_int64)]
    L.onilda_em_log_space_words.argtypes = [_vp, C.POINTER(C.c_int64)]
    L.onilda_em_word_steps.argtypes = [_vp, C.POINTER(C.c_double)]
    L.onilda_em_stream.argtypes = [_vp]
    L.onilda_em_stream.restype = _vp
    L.onilda_run_em.argtypes = [C.c_char_p, C.c_char_p, _vp, C.c_int, C.c_double, C.POINTER(Settings), C.c_int,
                                C.c_uint32]
    L.onilda_infer.argtypes = [_vp, _vp, C.POINTER(Settings), C.c_int, _f64p, _f64p]
    L.onilda_infer_files.argtypes = [C.c_char_p, C.c_char_p, _vp, C.POINTER(Settings), C.c_int]
    L.onilda_theta_from_gamma.argtypes = [_f64p, C.c_int64, C.c_int, C.c_int, _f64p]
    L.onilda_phi_from_log_beta.argtypes = [_f64p, C.c_int, C.c_int, C.c_int, _f64p]
    L.onilda_scorer_create.argtypes = [_vp, C.c_int64, _vp, C.c_int32, C.c_int, C.c_int, C.POINTER(_vp)]
    L.onilda_scorer_create_from_em.argtypes = [_vp, C.POINTER(_vp)]
    L.onilda_scorer_destroy.argtypes = [_vp]
    L.onilda_results_rows.argtypes = [_vp, _vp, _vp, C.c_int64, C.c_double, C.c_int64, _vp, _vp, C.POINTER(C.c_int64)]
    L.onilda_results_write.argtypes = [C.c_char_p, C.POINTER(C.c_char_p), C.c_int64, _vp, _vp, C.c_char_p]
    L.onilda_format_java_double.argtypes = [C.c_double, C.c_char_p]
    L.onilda_scorer_destroy.restype = None
    L.onilda_score.argtypes = [_vp, _vp, _vp, C.c_int64, _vp]
    L.onilda_score_flow_min.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int64, _vp]
    L.onilda_topk.argtypes = [_vp, _vp, C.c_int64, C.c_double, C.c_int64, _i64p, _f64p, C.POINTER(C.c_int64)]
    L.onilda_score_flow_topk_device.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int64, C.c_double, C.c_int64, _i64p,
                                                _f64p, C.POINTER(C.c_int64), _vp]
    L.onilda_scorer_stats.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.onilda_topk_device_noalloc.argtypes = [_vp, C.c_int64, C.c_double, C.c_int64, _i64p, _f64p, C.POINTER(C.c_int64)]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise OnildaError(rc, lib().onilda_last_error().decode())


def device_count() -> int:
    return lib().onilda_device_count()


def read_settings(path) -> Settings:
    s = Settings()
    _check(lib().onilda_settings_read(str(path).encode(), C.byref(s)))
    return s


def _ptr(a):
    return a.ctypes.data_as(_vp) if a is not None else None


class Corpus:
    """CSR corpus (replaces lda-c's `corpus`/`document`, lda.h:23-37)."""

    def __init__(self, handle):
        self._h = handle
        D, N, V, T, M = C.c_int64(), C.c_int64(), C.c_int32(), C.c_int64(), C.c_int32()
        _check(lib().onilda_corpus_dims(self._h, C.byref(D), C.byref(N), C.byref(V), C.byref(T), C.byref(M)))
        self.num_docs, self.nnz, self.num_terms, self.tokens, self.max_len = D.value, N.value, V.value, T.value, M.value

    @classmethod
    def from_dat(cls, path):
        h = _vp()
        _check(lib().onilda_corpus_from_dat(str(path).encode(), C.byref(h)))
        return cls(h)

    @classmethod
    def from_csr(cls, offsets, words, counts, num_terms=0):
        offsets = np.ascontiguousarray(offsets, np.int64)
        words = np.ascontiguousarray(words, np.int32)
        counts = np.ascontiguousarray(counts, np.int32)
        h = _vp()
        _check(lib().onilda_corpus_from_csr(offsets, _ptr(words), _ptr(counts), len(offsets) - 1, num_terms, C.byref(h)))
        return cls(h)

    @classmethod
    def from_triples(cls, doc, word, count, num_docs=0, num_terms=0, device=-1):
        """(document id, word id, count) triples -> corpus, grouped on the GPU
        (OniLDACWrapper.createModel, OniLDACWrapper.scala:146-168)."""
        doc = np.ascontiguousarray(doc, np.int32)
        word = np.ascontiguousarray(word, np.int32)
        count = np.ascontiguousarray(count, np.int32)
        if not (len(doc) == len(word) == len(count)):
            raise ValueError("doc, word and count must have the same length")
        h = _vp()
        _check(lib().onilda_corpus_from_triples(_ptr(doc), _ptr(word), _ptr(count), len(doc), num_docs, num_terms,
                                                device, C.byref(h)))
        return cls(h)

    def save_dat(self, path, doc_begin=0, doc_end=None):
        _check(lib().onilda_corpus_save_dat(self._h, doc_begin, self.num_docs if doc_end is None else doc_end,
                                            str(path).encode()))

    def pin(self):
        _check(lib().onilda_corpus_pin(self._h))

    def csr(self):
        po, pw, pc, pt = _vp(), _vp(), _vp(), _vp()
        _check(lib().onilda_corpus_csr(self._h, C.byref(po), C.byref(pw), C.byref(pc), C.byref(pt)))

        def view(p, n, ct, dt):
            if n == 0:
                return np.zeros(0, dt)
            return np.ctypeslib.as_array(C.cast(p, C.POINTER(ct)), (n,)).copy()
        return (view(po, self.num_docs + 1, C.c_int64, np.int64), view(pw, self.nnz, C.c_int32, np.int32),
                view(pc, self.nnz, C.c_int32, np.int32), view(pt, self.num_docs, C.c_int32, np.int32))

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.onilda_corpus_free(self._h)
            self._h = None


class Model:
    """alpha + log_beta [K][V] (replaces lda_model, lda.h:40-47)."""

    def __init__(self, handle):
        self._h = handle
        K, V = C.c_int(), C.c_int()
        _check(lib().onilda_model_dims(self._h, C.byref(K), C.byref(V)))
        self.num_topics, self.num_terms = K.value, V.value

    @classmethod
    def new(cls, num_topics, num_terms):
        h = _vp()
        _check(lib().onilda_model_new(num_topics, num_terms, C.byref(h)))
        return cls(h)

    @classmethod
    def load(cls, root):
        h = _vp()
        _check(lib().onilda_model_load(str(root).encode(), C.byref(h)))
        return cls(h)

    @classmethod
    def from_arrays(cls, alpha, log_beta):
        log_beta = np.asarray(log_beta, np.float64)
        m = cls.new(log_beta.shape[0], log_beta.shape[1])
        m.log_beta[...] = log_beta
        m.alpha = alpha
        return m

    def save(self, root):
        _check(lib().onilda_model_save(self._h, str(root).encode()))

    @property
    def alpha(self):
        return lib().onilda_model_alpha(self._h)

    @alpha.setter
    def alpha(self, v):
        _check(lib().onilda_model_set_alpha(self._h, float(v)))

    @property
    def log_beta(self):
        """Writable numpy view [K][V] of the model's own storage."""
        n = self.num_topics * self.num_terms
        if n == 0:
            return np.zeros((self.num_topics, self.num_terms))
        p = lib().onilda_model_log_beta(self._h)
        buf = (C.c_double * n).from_address(C.addressof(p.contents))
        buf._owner = self          # the view keeps the model (and its storage) alive
        return np.frombuffer(buf, dtype=np.float64).reshape(self.num_topics, self.num_terms)

    def init_random(self, seed):
        _check(lib().onilda_model_init_random(self._h, seed))

    def init_seeded(self, corpus: Corpus, seed):
        _check(lib().onilda_model_init_seeded(self._h, corpus._h, seed))

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.onilda_model_free(self._h)
            self._h = None


class FlowCuts(C.Structure):
    _fields_ = [("time", C.c_double * 10), ("ibyt", C.c_double * 10), ("ipkt", C.c_double * 5)]


def quantiles(values, thresholds, device=-1):
    """Quantiles.computeQuantiles (Quantiles.scala:53-75) of an integer column."""
    values = np.ascontiguousarray(values, np.int64)
    thresholds = np.ascontiguousarray(thresholds, np.float64)
    out = np.empty(len(thresholds), np.float64)
    _check(lib().onilda_quantiles_i64(_ptr(values), len(values), _ptr(thresholds), len(thresholds), device, _ptr(out)))
    return out


def flow_words(hour, minute, second, ibyt, ipkt, sport, dport, device=-1):
    """FlowWordCreation.flowWordCreation (FlowWordCreation.scala:14-146) on integer columns.
    Returns (cuts, src_code, dst_code); flow_word_string(code) gives the reference's word."""
    cols32 = [np.ascontiguousarray(x, np.int32) for x in (hour, minute, second)]
    cols64 = [np.ascontiguousarray(x, np.int64) for x in (ibyt, ipkt)]
    ports = [np.ascontiguousarray(x, np.int32) for x in (sport, dport)]
    n = len(cols32[0])
    if any(len(x) != n for x in cols32 + cols64 + ports):
        raise ValueError("columns must have the same length")
    cuts = FlowCuts()
    sc, dc = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
    _check(lib().onilda_flow_words(*[_ptr(x) for x in cols32 + cols64 + ports], n, device, C.byref(cuts), _ptr(sc), _ptr(dc)))
    return ({"time": np.array(cuts.time), "ibyt": np.array(cuts.ibyt), "ipkt": np.array(cuts.ipkt)}, sc[:n], dc[:n])


def flow_prelda(hour, minute, second, ibyt, ipkt, sport, dport, sip, dip, num_docs=0, device=-1, vocab_cap=1 << 20):
    """FlowPreLDA + word dictionary + createModel in one call (FlowPreLDA.scala:128-139, OniLDACWrapper.scala:46-68,146-168).
    Returns (cuts, Corpus, vocab codes, source word ids, destination word ids)."""
    c32 = [np.ascontiguousarray(x, np.int32) for x in (hour, minute, second)]
    c64 = [np.ascontiguousarray(x, np.int64) for x in (ibyt, ipkt)]
    rest = [np.ascontiguousarray(x, np.int32) for x in (sport, dport, sip, dip)]
    n = len(c32[0])
    if any(len(x) != n for x in c32 + c64 + rest):
        raise ValueError("columns must have the same length")
    cuts = FlowCuts()
    h = _vp()
    vocab = np.empty(max(vocab_cap, 1), np.int32)
    V = C.c_int32()
    sw, dw = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
    _check(lib().onilda_flow_prelda(*[_ptr(x) for x in c32 + c64 + rest], n, num_docs, device, C.byref(cuts), C.byref(h),
                                    _ptr(vocab), vocab_cap, C.byref(V), _ptr(sw), _ptr(dw)))
    return ({"time": np.array(cuts.time), "ibyt": np.array(cuts.ibyt), "ipkt": np.array(cuts.ipkt)}, Corpus(h),
            vocab[:V.value].copy(), sw[:n], dw[:n])


def flow_word_string(code):
    buf = C.create_string_buffer(64)
    n = lib().onilda_flow_word_string(int(code), buf, 64)
    if n < 0:
        _check(n)
    return buf.value.decode()


class DnsCuts(C.Structure):
    _fields_ = [("time", C.c_double * 10), ("frame_length", C.c_double * 10), ("subdomain_length", C.c_double * 5),
                ("entropy", C.c_double * 5), ("num_periods", C.c_double * 5)]


def _pack_strings(strings):
    raw = [x if isinstance(x, bytes) else str(x).encode() for x in strings]
    off = np.zeros(len(raw) + 1, np.int64)
    np.cumsum([len(x) for x in raw], out=off[1:])
    return np.frombuffer(b"".join(raw) or b"\0", np.uint8).copy(), off


def dns_words(names, frame_len, unix_tstamp, qry_type, qry_rcode, top_domains=(), device=-1, features=False):
    """DNSWordCreation.dnsWordCreation (dns/DNSWordCreation.scala:32-236) on columns.
    Returns (cuts, word_code int64[, features]); dns_word_string(code) gives the reference's word."""
    nb, noff = _pack_strings(names)
    tb, toff = _pack_strings(top_domains)
    n = len(noff) - 1
    fl, qt, rc = (np.ascontiguousarray(x, np.int32) for x in (frame_len, qry_type, qry_rcode))
    ts = np.ascontiguousarray(unix_tstamp, np.int64)
    if not (len(fl) == len(qt) == len(rc) == len(ts) == n):
        raise ValueError("columns must have the same length")
    cuts = DnsCuts()
    code = np.empty(max(n, 1), np.int64)
    m = max(n, 1)
    f = {"subdomain_length": np.zeros(m), "num_parts": np.zeros(m), "entropy": np.zeros(m), "top": np.zeros(m, np.int32),
         "spans": np.zeros((m, 3), np.int32)} if features else None
    _check(lib().onilda_dns_words_features(_ptr(nb), _ptr(noff), _ptr(fl), _ptr(ts), _ptr(qt), _ptr(rc), n, _ptr(tb), _ptr(toff),
                                           len(toff) - 1, device, C.byref(cuts), _ptr(code),
                                           *([_ptr(f[k]) for k in ("subdomain_length", "num_parts", "entropy", "top", "spans")]
                                             if features else [None] * 5)))
    out = ({k: np.array(getattr(cuts, k)) for k, _ in DnsCuts._fields_}, code[:n])
    return out + ({k: v[:n] for k, v in f.items()},) if features else out


def dns_word_string(code):
    buf = C.create_string_buffer(64)
    n = lib().onilda_dns_word_string(int(code), buf, 64)
    if n < 0:
        _check(n)
    return buf.value.decode()


def pairs_count(doc, word, num_docs=0, num_terms=0, device=-1):
    """(document id, word id) occurrences -> (doc, word, count) of the distinct pairs, sorted by (doc, word):
    the reduceByKey of FlowPreLDA.scala:131-137 / DNSPreLDA.scala:114-120, on the GPU."""
    doc = np.ascontiguousarray(doc, np.int32)
    word = np.ascontiguousarray(word, np.int32)
    if len(doc) != len(word):
        raise ValueError("doc and word must have the same length")
    n = len(doc)
    od, ow, oc = (np.empty(max(n, 1), np.int32) for _ in range(3))
    m = C.c_int64()
    _check(lib().onilda_pairs_count(_ptr(doc), _ptr(word), n, num_docs, num_terms, device, _ptr(od), _ptr(ow), _ptr(oc), C.byref(m)))
    return od[:m.value].copy(), ow[:m.value].copy(), oc[:m.value].copy()


def save_gamma(path, gamma):
    gamma = np.ascontiguousarray(gamma, np.float64)
    _check(lib().onilda_gamma_save(str(path).encode(), gamma, gamma.shape[0], gamma.shape[1]))


class EM:
    """run_em's state for one shard on one device (lda-estimate.c:112-321)."""

    def __init__(self, corpus: Corpus, model: Model, settings: Settings, device=-1, doc_begin=0, doc_end=None):
        self.corpus, self.settings = corpus, settings
        self.doc_begin = doc_begin
        self.doc_end = corpus.num_docs if doc_end is None else doc_end
        self.num_topics, self.num_terms = model.num_topics, model.num_terms
        h = _vp()
        _check(lib().onilda_em_create(corpus._h, self.doc_begin, self.doc_end, model._h, C.byref(settings), device,
                                      C.byref(h)))
        self._h = h

    def estep(self):
        _check(lib().onilda_em_estep(self._h))

    def reload(self, model: "Model"):
        _check(lib().onilda_em_reload(self._h, model._h))

    def reload_corpus(self):
        """host -> device copy of the shard's CSR arrays alone (the first half of reload)"""
        _check(lib().onilda_em_reload_corpus(self._h))

    def reload_model(self, model: "Model"):
        """alpha + log_beta host -> device and the EM loop state reset (the second half of reload)"""
        _check(lib().onilda_em_reload_model(self._h, model._h))

    def mstep(self):
        lik, conv, keep = C.c_double(), C.c_double(), C.c_int()
        _check(lib().onilda_em_mstep(self._h, C.byref(lik), C.byref(conv), C.byref(keep)))
        return lik.value, conv.value, bool(keep.value)

    def iterate(self):
        lik, conv, keep = C.c_double(), C.c_double(), C.c_int()
        _check(lib().onilda_em_iterate(self._h, C.byref(lik), C.byref(conv), C.byref(keep)))
        return lik.value, conv.value, bool(keep.value)

    def suffstats_device(self):
        """(device pointer, count of doubles) of the packed statistics buffer."""
        p, n = _vp(), C.c_int64()
        _check(lib().onilda_em_suffstats_device(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def stream(self):
        return lib().onilda_em_stream(self._h)

    def model(self) -> Model:
        m = Model.new(self.num_topics, self.num_terms)
        _check(lib().onilda_em_get_model(self._h, m._h))
        return m

    def gamma(self):
        g = np.zeros((self.doc_end - self.doc_begin, self.num_topics))
        _check(lib().onilda_em_get_gamma(self._h, g))
        return g

    def doc_likelihood(self):
        out = np.zeros(self.doc_end - self.doc_begin)
        _check(lib().onilda_em_get_doc_likelihood(self._h, out))
        return out

    def doc_sweeps(self):
        out = np.zeros(self.doc_end - self.doc_begin, np.int32)
        _check(lib().onilda_em_get_doc_sweeps(self._h, out))
        return out

    def word_steps(self):
        v = C.c_double()
        _check(lib().onilda_em_word_steps(self._h, C.byref(v)))
        return v.value

    def log_space_words(self):
        n = C.c_int64()
        _check(lib().onilda_em_log_space_words(self._h, C.byref(n)))
        return n.value

    def stats(self, with_sweeps=False):
        e, m, l, s = C.c_double(), C.c_double(), C.c_int64(), C.c_int64()
        _check(lib().onilda_em_stats(self._h, C.byref(e), C.byref(m), C.byref(l), C.byref(s) if with_sweeps else None))
        return dict(estep_ms=e.value, mstep_ms=m.value, launches=l.value, sweeps=s.value)

    def close(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.onilda_em_destroy(self._h)
            self._h = None

    __del__ = close


def allreduce_peer(ems):
    arr = (_vp * len(ems))(*[e._h for e in ems])
    _check(lib().onilda_em_allreduce_peer(arr, len(ems)))


COMM_ID_BYTES = 128


def _nccl_of_this_process():
    """libonilda binds NCCL at the first communicator call (csrc/nccl_dyn.h) and takes the libnccl.so.2 the process
    already holds.  A Python process that will also import PyTorch must hold PyTorch's own NCCL by then - the
    system library is older, and loading it first makes a later `import torch` fail on missing symbols - so
    PyTorch, where installed, is imported before the first such call."""
    try:
        import torch  # noqa: F401
    except ImportError:
        pass


def comm_unique_id() -> bytes:
    """The 128 bytes rank 0 ships to every other rank before comm_attach (ncclGetUniqueId)."""
    _nccl_of_this_process()
    buf = C.create_string_buffer(COMM_ID_BYTES)
    _check(lib().onilda_comm_unique_id(buf))
    return buf.raw


def comm_attach(em: "EM", comm_id: bytes, nranks: int, rank: int):
    """One process per GPU: collective over all ranks."""
    assert len(comm_id) == COMM_ID_BYTES
    _nccl_of_this_process()
    _check(lib().onilda_em_comm_attach(em._h, comm_id, nranks, rank))


def comm_attach_local(ems):
    """One process, several devices."""
    _nccl_of_this_process()
    arr = (_vp * len(ems))(*[e._h for e in ems])
    _check(lib().onilda_em_comm_attach_local(arr, len(ems)))


def mstep_group(ems):
    arr = (_vp * len(ems))(*[e._h for e in ems])
    lik, conv, keep = C.c_double(), C.c_double(), C.c_int()
    _check(lib().onilda_em_mstep_group(arr, len(ems), C.byref(lik), C.byref(conv), C.byref(keep)))
    return lik.value, conv.value, bool(keep.value)


def comm_stats(em: "EM"):
    n, r, rs, ag = C.c_int(), C.c_int(), C.c_double(), C.c_double()
    _check(lib().onilda_em_comm_stats(em._h, C.byref(n), C.byref(r), C.byref(rs), C.byref(ag)))
    return dict(nranks=n.value, rank=r.value, reduce_scatter_ms=rs.value, all_gather_ms=ag.value)


def collective_time(ems, repeats=10):
    """Per-call ms of the M-step's reduce-scatter and of its three all-gathers, and the table bytes of one call."""
    arr = (_vp * len(ems))(*[e._h for e in ems])
    rs, ag, nb = C.c_double(), C.c_double(), C.c_int64()
    _check(lib().onilda_em_collective_time(arr, len(ems), repeats, C.byref(rs), C.byref(ag), C.byref(nb)))
    return dict(reduce_scatter_ms=rs.value, all_gather_ms=ag.value, table_bytes=nb.value)


def run_em(start, directory, corpus: Corpus, num_topics, initial_alpha, settings: Settings, ndev=1, seed=0):
    _check(lib().onilda_run_em(str(start).encode(), str(directory).encode(), corpus._h, num_topics,
                               float(initial_alpha), C.byref(settings), ndev, seed))


def infer(model: Model, corpus: Corpus, settings: Settings, device=-1):
    gamma = np.zeros((corpus.num_docs, model.num_topics))
    lik = np.zeros(corpus.num_docs)
    _check(lib().onilda_infer(model._h, corpus._h, C.byref(settings), device, gamma, lik))
    return gamma, lik


def infer_files(model_root, name, corpus: Corpus, settings: Settings, device=-1):
    _check(lib().onilda_infer_files(str(model_root).encode(), str(name).encode(), corpus._h, C.byref(settings), device))


def theta_from_gamma(gamma, device=-1):
    gamma = np.ascontiguousarray(gamma, np.float64)
    out = np.zeros_like(gamma)
    _check(lib().onilda_theta_from_gamma(gamma, gamma.shape[0], gamma.shape[1], device, out))
    return out


def phi_from_log_beta(log_beta, device=-1):
    log_beta = np.ascontiguousarray(log_beta, np.float64)
    K, V = log_beta.shape
    out = np.zeros((V, K))
    _check(lib().onilda_phi_from_log_beta(log_beta, K, V, device, out))
    return out


class Scorer:
    """theta/phi resident on one device (replaces the broadcast maps of
    SuspiciousConnectsScoreFunction.scala:7-9)."""

    def __init__(self, theta, phi_wk, device=-1):
        theta = np.ascontiguousarray(theta, np.float64)
        phi_wk = np.ascontiguousarray(phi_wk, np.float64)
        assert theta.shape[1] == phi_wk.shape[1]
        self.K = theta.shape[1]
        h = _vp()
        _check(lib().onilda_scorer_create(_ptr(theta), theta.shape[0], _ptr(phi_wk), phi_wk.shape[0], self.K, device,
                                          C.byref(h)))
        self._h = h

    @classmethod
    def from_em(cls, em: "EM"):
        """theta / phi straight from an EM handle's device memory (no text files, no host copies)."""
        self = cls.__new__(cls)
        self.K = em.num_topics
        h = _vp()
        _check(lib().onilda_scorer_create_from_em(em._h, C.byref(h)))
        self._h = h
        return self

    def score(self, doc, word, out=None):
        doc = np.ascontiguousarray(doc, np.int32)
        word = np.ascontiguousarray(word, np.int32)
        out = np.zeros(len(doc)) if out is None else out
        _check(lib().onilda_score(self._h, _ptr(doc), _ptr(word), len(doc), _ptr(out)))
        return out

    def score_flow_min(self, sdoc, sword, ddoc, dword, out=None):
        a = [np.ascontiguousarray(x, np.int32) for x in (sdoc, sword, ddoc, dword)]
        out = np.zeros(len(a[0])) if out is None else out
        _check(lib().onilda_score_flow_min(self._h, _ptr(a[0]), _ptr(a[1]), _ptr(a[2]), _ptr(a[3]), len(a[0]), _ptr(out)))
        return out

    def flow_results(self, sdoc, sword, ddoc, dword, threshold, k):
        """FlowPostLDA.scala:75-80: (event index, source score, destination score) of the kept rows, most suspicious first."""
        a = [np.ascontiguousarray(x, np.int32) for x in (sdoc, sword, ddoc, dword)]
        n = len(a[0])
        cap = max(1, n if k == -1 else min(k, n))
        idx, ss, ds = np.zeros(cap, np.int64), np.zeros(cap), np.zeros(cap)
        m = C.c_int64()
        _check(lib().onilda_flow_results(self._h, _ptr(a[0]), _ptr(a[1]), _ptr(a[2]), _ptr(a[3]), n, threshold, k,
                                         _ptr(idx), _ptr(ss), _ptr(ds), C.byref(m)))
        return idx[:m.value].copy(), ss[:m.value].copy(), ds[:m.value].copy()

    def results_rows(self, doc, word, threshold, k):
        """dns / proxy: (event index, Score) of the kept rows, most suspicious first (DNSPostLDA.scala:44-70)."""
        a = [np.ascontiguousarray(x, np.int32) for x in (doc, word)]
        n = len(a[0])
        cap = max(1, n if k == -1 else min(k, n))
        idx, sc = np.zeros(cap, np.int64), np.zeros(cap)
        m = C.c_int64()
        _check(lib().onilda_results_rows(self._h, _ptr(a[0]), _ptr(a[1]), n, threshold, k, _ptr(idx), _ptr(sc), C.byref(m)))
        return idx[:m.value].copy(), sc[:m.value].copy()

    def score_raw(self, sdoc_ptr, sword_ptr, ddoc_ptr, dword_ptr, n, out_ptr):
        """Host pointers given as integers (e.g. pinned torch tensors' data_ptr())."""
        if ddoc_ptr:
            _check(lib().onilda_score_flow_min(self._h, sdoc_ptr, sword_ptr, ddoc_ptr, dword_ptr, n, out_ptr))
        else:
            _check(lib().onilda_score(self._h, sdoc_ptr, sword_ptr, n, out_ptr))

    def topk(self, score, threshold, k):
        score = np.ascontiguousarray(score, np.float64)
        n = len(score)
        cap = n if k == -1 else min(k, n)
        idx, sc, m = np.zeros(max(cap, 1), np.int64), np.zeros(max(cap, 1)), C.c_int64()
        _check(lib().onilda_topk(self._h, _ptr(score), n, threshold, k, idx, sc, C.byref(m)))
        return idx[:m.value].copy(), sc[:m.value].copy()

    def score_flow_topk_device(self, sdoc_dev, sword_dev, ddoc_dev, dword_dev, n, threshold, k, score_dev_out=None):
        """Device pointers (ints).  Returns (idx, score) of the k most suspicious events."""
        cap = n if k == -1 else min(k, n)
        idx, sc, m = np.zeros(max(cap, 1), np.int64), np.zeros(max(cap, 1)), C.c_int64()
        _check(lib().onilda_score_flow_topk_device(self._h, sdoc_dev, sword_dev, ddoc_dev, dword_dev, n, threshold, k,
                                                   idx, sc, C.byref(m), score_dev_out))
        return idx[:m.value].copy(), sc[:m.value].copy()

    def stats(self):
        ms, l = C.c_double(), C.c_int64()
        _check(lib().onilda_scorer_stats(self._h, C.byref(ms), C.byref(l)))
        return dict(kernel_ms=ms.value, launches=l.value)

    def close(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.onilda_scorer_destroy(self._h)
            self._h = None

    __del__ = close


def java_double(v: float) -> str:
    buf = C.create_string_buffer(40)
    n = lib().onilda_format_java_double(float(v), buf)
    return buf.raw[:n].decode()


def results_write(path, rows, score_a, score_b=None, delimiter=","):
    """<type>_results.csv: rows[i] + delimiter + score_a[i] [+ delimiter + score_b[i]]."""
    rows = [r.encode() if isinstance(r, str) else r for r in rows]
    arr = (C.c_char_p * max(len(rows), 1))(*rows)
    a = np.ascontiguousarray(score_a, np.float64)
    b = None if score_b is None else np.ascontiguousarray(score_b, np.float64)
    _check(lib().onilda_results_write(str(path).encode(), arr, len(rows), _ptr(a), _ptr(b) if b is not None else None,
                                      delimiter.encode()))


NOALLOC_MAX_K = 4096


def topk_device_noalloc(score_dev_ptr, n, threshold, k):
    """onilda_topk_device_noalloc: scores already on the current device (pointer as int); the
    call allocates nothing in the library (the two small result arrays here are the caller's)."""
    cap = max(min(k, n), 1) if k >= 0 else 1
    idx, sc, m = np.zeros(cap, np.int64), np.zeros(cap), C.c_int64()
    _check(lib().onilda_topk_device_noalloc(score_dev_ptr, n, threshold, k, idx, sc, C.byref(m)))
    return idx[:m.value].copy(), sc[:m.value].copy()
