"""ctypes view of oracle/liboracle.so (the CPU restatement) and a runner for
oracle/_ref/lda (the unmodified reference binary).  TEST INFRASTRUCTURE: import
this only from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "liboracle.so"
REF_BIN = HERE / "_ref" / "lda"

_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


class Settings(C.Structure):
    _fields_ = [("var_max_iter", C.c_int), ("var_converged", C.c_float),
                ("em_max_iter", C.c_int), ("em_converged", C.c_float),
                ("estimate_alpha", C.c_int)]


def build(force: bool = False) -> None:
    """make liboracle.so, and _ref/lda when the reference sources are present."""
    if force or not LIB_PATH.exists():
        subprocess.run(["make", "-C", str(HERE), "liboracle.so"], check=True, capture_output=True)
    if os.path.isdir("/root/reference/oni-ml/oni-lda-c") and (force or not REF_BIN.exists()):
        subprocess.run(["make", "-C", str(HERE), "ref"], check=True, capture_output=True)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(str(LIB_PATH))
    L.oracle_log_sum.restype = C.c_double
    L.oracle_log_sum.argtypes = [C.c_double, C.c_double]
    for f in (L.oracle_digamma, L.oracle_trigamma):
        f.restype = C.c_double
        f.argtypes = [C.c_double]
    L.oracle_mt_seed.argtypes = [C.c_ulong]
    L.oracle_mt_next.restype = C.c_ulong
    L.oracle_myrand.restype = C.c_double
    L.oracle_read_settings.argtypes = [C.c_char_p, C.POINTER(Settings)]
    L.oracle_read_dat.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_int64)), C.POINTER(C.POINTER(C.c_int32)),
                                  C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.POINTER(C.c_int32)),
                                  C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
    L.oracle_free.argtypes = [C.c_void_p]
    L.oracle_opt_alpha.restype = C.c_double
    L.oracle_opt_alpha.argtypes = [C.c_double, C.c_int, C.c_int]
    L.oracle_lda_mle.argtypes = [C.c_int, C.c_int, _f64p, _f64p, _f64p]
    L.oracle_random_init.argtypes = [C.c_int, C.c_int, _f64p]
    L.oracle_seeded_init.argtypes = [C.c_int, C.c_int, C.c_int64, _i64p, _i32p, _i32p, _f64p]
    L.oracle_save_model.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_double, _f64p]
    L.oracle_load_model_dims.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.oracle_load_model.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_double), _f64p]
    L.oracle_save_gamma.argtypes = [C.c_char_p, C.c_int64, C.c_int, _f64p]
    L.oracle_compute_likelihood.restype = C.c_double
    L.oracle_compute_likelihood.argtypes = [C.c_int, _i32p, _i32p, C.c_int, C.c_int, C.c_double, _f64p, _f64p, _f64p]
    L.oracle_lda_inference.restype = C.c_double
    L.oracle_lda_inference.argtypes = [C.c_int, _i32p, _i32p, C.c_int, C.c_int, C.c_int, C.c_double, _f64p,
                                       C.c_int, C.c_float, _f64p, _f64p, C.POINTER(C.c_int)]
    L.oracle_e_step.restype = C.c_double
    L.oracle_e_step.argtypes = [C.c_int64, _i64p, _i32p, _i32p, _i32p, C.c_int, C.c_int, C.c_double, _f64p,
                                C.c_int, C.c_float, C.c_int, C.c_int, _f64p, _f64p, _f64p,
                                C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p]
    L.oracle_run_em.argtypes = [C.c_int64, _i64p, _i32p, _i32p, _i32p, C.c_int, C.c_int, C.POINTER(C.c_double),
                                _f64p, C.POINTER(Settings), C.c_int, _f64p, _f64p, _f64p, C.POINTER(C.c_int)]
    L.oracle_infer.argtypes = [C.c_int64, _i64p, _i32p, _i32p, _i32p, C.c_int, C.c_int, C.c_double, _f64p,
                               C.POINTER(Settings), _f64p, _f64p]
    L.oracle_doc_sweeps.argtypes = [C.c_int64, _i64p, _i32p, _i32p, _i32p, C.c_int, C.c_int, C.c_double, _f64p,
                                    C.c_int, C.c_float, _i32p]
    L.oracle_set_threads.argtypes = [C.c_int]
    L.oracle_theta_from_gamma.argtypes = [C.c_int64, C.c_int, _f64p, _f64p]
    L.oracle_phi_from_logbeta.argtypes = [C.c_int, C.c_int, _f64p, _f64p]
    L.oracle_score.argtypes = [C.c_int64, _i32p, _i32p, C.c_int64, C.c_int, C.c_int, _f64p, _f64p, _f64p]
    L.oracle_score_flow_min.argtypes = [C.c_int64, _i32p, _i32p, _i32p, _i32p, C.c_int64, C.c_int, C.c_int,
                                        _f64p, _f64p, _f64p]
    L.oracle_topk.restype = C.c_int64
    L.oracle_topk.argtypes = [C.c_int64, _f64p, C.c_double, C.c_int64, _i64p, _f64p]
    _lib = L
    return L


def settings(var_max_iter=20, var_converged=1e-6, em_max_iter=20, em_converged=1e-6, estimate_alpha=1) -> Settings:
    return Settings(var_max_iter, var_converged, em_max_iter, em_converged, estimate_alpha)


def read_settings(path) -> Settings:
    s = Settings()
    if lib().oracle_read_settings(str(path).encode(), C.byref(s)) != 0:
        raise FileNotFoundError(path)
    return s


def read_dat(path):
    """-> (offsets int64[D+1], words int32[N], counts int32[N], totals int32[D], V)"""
    L = lib()
    po, pw, pc, pt = (C.POINTER(C.c_int64)(), C.POINTER(C.c_int32)(), C.POINTER(C.c_int32)(), C.POINTER(C.c_int32)())
    D, N, V = C.c_int64(), C.c_int64(), C.c_int32()
    if L.oracle_read_dat(str(path).encode(), C.byref(po), C.byref(pw), C.byref(pc), C.byref(pt),
                         C.byref(D), C.byref(N), C.byref(V)) != 0:
        raise FileNotFoundError(path)
    off = np.ctypeslib.as_array(po, (D.value + 1,)).copy()
    words = np.ctypeslib.as_array(pw, (max(N.value, 1),))[:N.value].copy()
    counts = np.ctypeslib.as_array(pc, (max(N.value, 1),))[:N.value].copy()
    totals = np.ctypeslib.as_array(pt, (max(D.value, 1),))[:D.value].copy()
    for p in (po, pw, pc, pt):
        L.oracle_free(C.cast(p, C.c_void_p))
    return off, words, counts, totals, V.value


def totals_of(offsets, counts):
    # int32 wrap-around matches the reference's int `total` (lda.h:27)
    cs = np.concatenate([[0], np.cumsum(counts.astype(np.int64))])
    return (cs[offsets[1:]] - cs[offsets[:-1]]).astype(np.int32)


def load_model(root):
    L = lib()
    K, V = C.c_int(), C.c_int()
    if L.oracle_load_model_dims(str(root).encode(), C.byref(K), C.byref(V)) != 0:
        raise FileNotFoundError(root)
    lb = np.zeros((K.value, V.value))
    a = C.c_double()
    rc = L.oracle_load_model(str(root).encode(), K.value, V.value, C.byref(a), lb)
    if rc != 0:
        raise RuntimeError(f"oracle_load_model rc={rc}")
    return a.value, lb


def save_model(root, alpha, log_beta):
    K, V = log_beta.shape
    if lib().oracle_save_model(str(root).encode(), K, V, float(alpha), np.ascontiguousarray(log_beta)) != 0:
        raise OSError(root)


def run_em(offsets, words, counts, V, alpha, log_beta, st: Settings, nproc=1):
    """-> dict(alpha, log_beta, gamma, lik, conv)"""
    L = lib()
    D = len(offsets) - 1
    K = log_beta.shape[0]
    lb = np.ascontiguousarray(log_beta, dtype=np.float64).copy()
    gamma = np.zeros((D, K))
    cap = st.em_max_iter + 2
    lik, conv = np.zeros(cap), np.zeros(cap)
    a = C.c_double(alpha)
    n = C.c_int()
    totals = totals_of(offsets, counts)
    L.oracle_run_em(D, offsets, words, counts, totals, K, V, C.byref(a), lb, C.byref(st), nproc,
                    gamma, lik, conv, C.byref(n))
    return dict(alpha=a.value, log_beta=lb, gamma=gamma, lik=lik[:n.value].copy(), conv=conv[:n.value].copy())


def e_step(offsets, words, counts, V, alpha, log_beta, var_max_iter, var_converged):
    """One full E-step pass -> dict(gamma, class_word[K][V], class_total, alpha_ss, lik, doc_lik)"""
    L = lib()
    D = len(offsets) - 1
    K = log_beta.shape[0]
    gamma = np.zeros((D, K))
    cw, ct = np.zeros((K, V)), np.zeros(K)
    a_ss, nd = C.c_double(0.0), C.c_int(0)
    doc_lik = np.zeros(D)
    totals = totals_of(offsets, counts)
    lik = L.oracle_e_step(D, offsets, words, counts, totals, K, V, alpha, np.ascontiguousarray(log_beta),
                          var_max_iter, var_converged, 0, 1, gamma, cw, ct, C.byref(a_ss), C.byref(nd),
                          doc_lik.ctypes.data_as(C.c_void_p))
    return dict(gamma=gamma, class_word=cw, class_total=ct, alpha_ss=a_ss.value, num_docs=nd.value,
                lik=lik, doc_lik=doc_lik)


def doc_sweeps(offsets, words, counts, V, alpha, log_beta, var_max_iter, var_converged):
    D = len(offsets) - 1
    out = np.zeros(D, np.int32)
    lib().oracle_doc_sweeps(D, offsets, words, counts, totals_of(offsets, counts), log_beta.shape[0], V, alpha,
                            np.ascontiguousarray(log_beta), var_max_iter, var_converged, out)
    return out


def infer(offsets, words, counts, V, alpha, log_beta, st: Settings):
    L = lib()
    D = len(offsets) - 1
    K = log_beta.shape[0]
    gamma, doc_lik = np.zeros((D, K)), np.zeros(D)
    L.oracle_infer(D, offsets, words, counts, totals_of(offsets, counts), K, V, alpha,
                   np.ascontiguousarray(log_beta), C.byref(st), gamma, doc_lik)
    return gamma, doc_lik


def theta_from_gamma(gamma):
    D, K = gamma.shape
    out = np.zeros_like(gamma)
    lib().oracle_theta_from_gamma(D, K, np.ascontiguousarray(gamma), out)
    return out


def phi_from_logbeta(log_beta):
    K, V = log_beta.shape
    out = np.zeros((V, K))
    lib().oracle_phi_from_logbeta(K, V, np.ascontiguousarray(log_beta), out)
    return out


def score(doc, word, theta, phi_wk, threads=1):
    L = lib()
    L.oracle_set_threads(threads)
    out = np.zeros(len(doc))
    L.oracle_score(len(doc), doc, word, theta.shape[0], phi_wk.shape[0], theta.shape[1], theta, phi_wk, out)
    return out


def score_flow_min(sdoc, sword, ddoc, dword, theta, phi_wk, threads=1):
    L = lib()
    L.oracle_set_threads(threads)
    out = np.zeros(len(sdoc))
    L.oracle_score_flow_min(len(sdoc), sdoc, sword, ddoc, dword, theta.shape[0], phi_wk.shape[0],
                            theta.shape[1], theta, phi_wk, out)
    return out


def topk(scores, threshold, k):
    n = len(scores)
    idx, sc = np.zeros(max(n, 1), np.int64), np.zeros(max(n, 1))
    m = lib().oracle_topk(n, np.ascontiguousarray(scores), threshold, k, idx, sc)
    return idx[:m].copy(), sc[:m].copy()


def compute_ecdf(data):
    """Quantiles.computeEcdf (Quantiles.scala:21-41): (value, fraction of the data <= value) for each distinct value,
    ascending; the running count is a double sum and the fraction a double division, as there."""
    data = np.asarray(data, np.float64)
    if len(data) == 0:
        return np.zeros(0), np.zeros(0)
    values, counts = np.unique(data, return_counts=True)
    return values, np.cumsum(counts.astype(np.float64)) / float(len(data))


def compute_quantiles(data, thresholds):
    """Quantiles.computeQuantiles (Quantiles.scala:53-75): per threshold the least value whose ecdf is >= it,
    +infinity when there is none."""
    values, cdf = compute_ecdf(data)
    out = []
    for t in thresholds:
        hit = np.nonzero(cdf >= t)[0]
        out.append(float(values[hit[0]]) if len(hit) else float("inf"))
    return np.array(out)


DECILES = [0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0]      # Quantiles.scala:86
QUINTILES = [0.2, 0.4, 0.6, 0.8, 1.0]                              # Quantiles.scala:94


def quantile_bin(value, cuts):
    """Quantiles.bin (Quantiles.scala:97-99): index of the first cut with value <= cut, -1 if none."""
    for i, c in enumerate(cuts):
        if value <= c:
            return i
    return -1


def scala_double(x):
    """Scala's Double.toString for the values that occur in words: whole numbers print as '23.0'."""
    return "%d.0" % int(x) if float(x) == int(x) else repr(float(x))


def adjust_port(source_ip, destination_ip, destination_port, source_port, ipkt_bin, ibyt_bin, time_bin):
    """FlowWordCreation.adjustPort (FlowWordCreation.scala:92-146), statement by statement.
    Returns [wordPort, ipPair, sourceWord, destinationWord]."""
    word_port = 111111.0
    ip_pair = destination_ip + " " + source_ip
    if source_ip < destination_ip and source_ip != 0:
        ip_pair = source_ip + " " + destination_ip
    dp, sp = destination_port, source_port
    if (dp <= 1024 or sp <= 1024) and (dp > 1024 or sp > 1024) and min(dp, sp) != 0:
        port_case, word_port = 2, float(min(dp, sp))
    elif dp > 1024 and sp > 1024:
        port_case, word_port = 3, 333333.0
    elif dp == 0 and sp != 0:
        port_case, word_port = 4, float(sp)
    elif sp == 0 and dp != 0:
        port_case, word_port = 4, float(dp)
    else:
        port_case = 1
        word_port = float(max(dp, sp)) if min(dp, sp) == 0 else 111111.0
    word = "_".join([scala_double(word_port), scala_double(time_bin), scala_double(ibyt_bin), scala_double(ipkt_bin)])
    source_word = destination_word = word
    if port_case == 2 and dp < sp:
        destination_word = "-1_" + destination_word
    elif port_case == 2 and sp < dp:
        source_word = "-1_" + source_word
    elif port_case == 4 and dp == 0:
        source_word = "-1_" + source_word
    elif port_case == 4 and sp == 0:
        destination_word = "-1_" + destination_word
    return [scala_double(word_port), ip_pair, source_word, destination_word]


def flow_words(hour, minute, second, ibyt, ipkt, sport, dport):
    """FlowWordCreation.flowWordCreation (FlowWordCreation.scala:14-84) on columns: the three cut arrays and the
    (sourceWord, destinationWord) strings per event."""
    num_time = [float(h) + float(m) / 60 + float(s) / 3600 for h, m, s in zip(hour, minute, second)]   # addTime, :88-90
    time_cuts = compute_quantiles(num_time, DECILES)
    ibyt_cuts = compute_quantiles(np.asarray(ibyt, np.float64), DECILES)
    ipkt_cuts = compute_quantiles(np.asarray(ipkt, np.float64), QUINTILES)
    src, dst = [], []
    for i in range(len(num_time)):
        r = adjust_port("a", "b", int(dport[i]), int(sport[i]), float(quantile_bin(float(ipkt[i]), ipkt_cuts)),
                        float(quantile_bin(float(ibyt[i]), ibyt_cuts)), float(quantile_bin(num_time[i], time_cuts)))
        src.append(r[2])
        dst.append(r[3])
    return {"time": time_cuts, "ibyt": ibyt_cuts, "ipkt": ipkt_cuts}, src, dst


# ISO country-code labels of DNSWordCreation.scala:17-30 (the set there also holds "" - never a last label)
COUNTRY_CODES = set("""ac ad ae af ag ai al am an ao aq ar as at au aw ax az ba bb bd be bf bg bh bi bj bm bn bo bq br bs bt
bv bw by bz ca cc cd cf cg ch ci ck cl cm cn co cr cu cv cw cx cy cz de dj dk dm do dz ec ee eg eh er es et eu fi fj fk fm
fo fr ga gb gd ge gf gg gh gi gl gm gn gp gq gr gs gt gu gw gy hk hm hn hr ht hu id ie il im in io iq ir is it je jm jo jp
ke kg kh ki km kn kp kr krd kw ky kz la lb lc li lk lr ls lt lu lv ly ma mc md me mg mh mk ml mm mn mo mp mq mr ms mt mu mv
mw mx my mz na nc ne nf ng ni nl no np nr nu nz om pa pe pf pg ph pk pl pm pn pr ps pt pw py qa re ro rs ru rw sa sb sc sd
se sg sh si sj sk sl sm sn so sr ss st su sv sx sy sz tc td tf tg th tj tk tl tm tn to tp tr tt tv tw tz ua ug uk us uy uz
va vc ve vg vi vn vu wf ws ye yt za zm zw""".split()) | {""}


def java_split_dots(url):
    """String.split("[.]"): trailing empty strings are removed; no dot -> the string itself."""
    parts = url.split(".")
    if len(parts) > 1:
        while parts and parts[-1] == "":
            parts.pop()
    return parts


def extract_subdomain(url):
    """DNSWordCreation.extractSubdomain (dns/DNSWordCreation.scala:155-196): [domain, subdomain, subdomain length, parts]."""
    parts = java_split_dots(url)
    n = len(parts)
    domain = subdomain = "None"
    is_ip = ("IP" if parts[n - 1] == "arpa" and parts[n - 2] == "in-addr" else "Name") if n > 2 else "Unknown"
    if n > 2 and is_ip != "IP":
        if parts[n - 1] in COUNTRY_CODES:
            domain = parts[n - 3]
            if 1 <= n - 3:
                subdomain = ".".join(parts[0:n - 3])
        else:
            domain = parts[n - 2]
            subdomain = ".".join(parts[0:n - 2])
    return [domain, subdomain, float(len(subdomain)) if subdomain != "None" else 0.0, float(n)]


def string_entropy(v):
    """Entropy.stringEntropy (utilities/Entropy.scala:16-23): sum over distinct characters of -p log10(p) / log10(2).
    The reference adds in the iteration order of a Scala Map (unspecified); here the largest count goes first, so that
    the value depends on the counts only."""
    import math
    counts = {}
    for ch in v:
        counts[ch] = counts.get(ch, 0) + 1
    total = 0.0
    for k in sorted(counts.values(), reverse=True):
        p = float(k) / len(v)
        total += -p * math.log10(p) / math.log10(2)
    return total


def get_top_domain(top_domains, domain):
    """DNSWordCreation.getTopDomain (:147-153)."""
    return "2" if domain == "intel" else ("1" if domain in top_domains else "0")


def dns_words(names, frame_len, unix_tstamp, qry_type, qry_rcode, top_domains=()):
    """DNSWordCreation.dnsWordCreation (:32-145) + dnsWord (:213-235) on columns: the five cut arrays and the words."""
    top = set(top_domains)
    feats = [extract_subdomain(u) for u in names]
    entropy = [string_entropy(f[1]) for f in feats]
    cuts = {
        "time": compute_quantiles(np.asarray(unix_tstamp, np.float64), DECILES),
        "frame_length": compute_quantiles(np.asarray(frame_len, np.float64), DECILES),
        "subdomain_length": compute_quantiles([f[2] for f in feats if f[2] > 0], QUINTILES),
        "entropy": compute_quantiles([e for e in entropy if e > 0], QUINTILES),
        "num_periods": compute_quantiles([f[3] for f in feats if f[3] > 0], QUINTILES),
    }
    words = []
    for i, f in enumerate(feats):
        words.append("_".join(str(x) for x in [
            get_top_domain(top, f[0]), quantile_bin(float(frame_len[i]), cuts["frame_length"]),
            quantile_bin(float(unix_tstamp[i]), cuts["time"]), quantile_bin(f[2], cuts["subdomain_length"]),
            quantile_bin(entropy[i], cuts["entropy"]), quantile_bin(f[3], cuts["num_periods"]),
            int(qry_type[i]), int(qry_rcode[i])]))
    return cuts, words


def pairs_count(doc, word):
    """`.map((ip, word) -> 1).reduceByKey(_ + _)` (FlowPreLDA.scala:131-137, DNSPreLDA.scala:114-120) on integer ids;
    Spark's output order is unspecified, this one is sorted by (doc, word).  Returns (doc, word, count)."""
    doc = np.asarray(doc, np.int64)
    word = np.asarray(word, np.int64)
    if len(doc) == 0:
        return np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.int32)
    span = int(word.max()) + 1
    key, cnt = np.unique(doc * span + word, return_counts=True)
    return (key // span).astype(np.int32), (key % span).astype(np.int32), cnt.astype(np.int32)


def triples_to_csr(doc, word, count, num_docs=0):
    """OniLDACWrapper.createModel (OniLDACWrapper.scala:146-168) followed by read_data
    (lda-data.c:22-59), without the text in between: documents in id order (the position in
    `distinctDocument`), entries of a document in input order, repeated words kept.
    Returns (offsets, words, counts)."""
    doc = np.asarray(doc, np.int64)
    if num_docs == 0:
        num_docs = int(doc.max()) + 1 if len(doc) else 0
    order = np.argsort(doc, kind="stable")
    lens = np.bincount(doc, minlength=num_docs) if len(doc) else np.zeros(num_docs, np.int64)
    offsets = np.zeros(num_docs + 1, np.int64)
    np.cumsum(lens, out=offsets[1:])
    return offsets, np.asarray(word, np.int32)[order], np.asarray(count, np.int32)[order]


def dat_lines(offsets, words, counts):
    """The model.dat lines createModel returns for a CSR corpus."""
    return [" ".join([str(int(offsets[d + 1] - offsets[d]))] +
                     ["%d:%d" % (w, c) for w, c in zip(words[offsets[d]:offsets[d + 1]], counts[offsets[d]:offsets[d + 1]])])
            for d in range(len(offsets) - 1)]


def run_ref(args, cwd, seed=None, nproc=1, timeout=3600):
    """Run the unmodified reference binary (oracle/_ref/lda).  Returns CompletedProcess."""
    if not REF_BIN.exists():
        raise FileNotFoundError(f"{REF_BIN} not built (needs /root/reference; run make -C oracle ref)")
    env = dict(os.environ)
    if seed is not None:
        env["ONI_ORACLE_SEED"] = str(seed)
    if nproc > 1:
        env["ONI_MPI_STUB_NP"] = str(nproc)
    else:
        env.pop("ONI_MPI_STUB_NP", None)
    return subprocess.run([str(REF_BIN)] + [str(a) for a in args], cwd=str(cwd), env=env,
                          capture_output=True, text=True, timeout=timeout, check=True)


def java_double_str(v):
    """Java's Double.toString(v) as JDK 19+ prints it (shortest digits that read back as v; older JDKs print one
    more digit for a small share of values): decimal for 1e-3 <= |v| < 1e7, else d.dddE[-]n.  Digits from Python's
    repr (shortest round trip).  This is the form of the score columns of <type>_results.csv: the reference writes
    rows with `_.mkString(outputDelimiter)` (netflow/FlowPostLDA.scala:80, dns/DNSPostLDA.scala:69)."""
    import math
    from decimal import Decimal
    if math.isnan(v):
        return "NaN"
    sign = "-" if math.copysign(1.0, v) < 0 else ""
    v = abs(v)
    if math.isinf(v):
        return sign + "Infinity"
    if v == 0:
        return sign + "0.0"
    t = Decimal(repr(v)).as_tuple()
    if len(t.digits) == 1:                          # Java prints two digits at least, the closest ones
        t = Decimal("%.1e" % v).as_tuple()
    digits = "".join(map(str, t.digits)).rstrip("0") or "0"
    e10 = len(t.digits) + t.exponent - 1            # exponent of the first digit
    if 1e-3 <= v < 1e7:
        if e10 >= 0:
            ip = (digits + "0" * (e10 + 1))[:e10 + 1]
            fp = digits[e10 + 1:] or "0"
            return f"{sign}{ip}.{fp}"
        return f"{sign}0.{'0' * (-e10 - 1)}{digits}"
    return f"{sign}{digits[0]}.{digits[1:] or '0'}E{e10}"


def results_lines(rows, score_a, score_b=None, delimiter=","):
    """Lines of <type>_results.csv: the event's own columns (already joined), then the score column(s)."""
    out = []
    for i, r in enumerate(rows):
        cols = [r, java_double_str(float(score_a[i]))]
        if score_b is not None:
            cols.append(java_double_str(float(score_b[i])))
        out.append(delimiter.join(cols))
    return out
