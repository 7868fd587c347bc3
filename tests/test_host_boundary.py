"""CPU-side tests of the drop-in boundary: the C ABI loads and exports what
include/onilda.h declares, and the host logic (settings, `.dat` parser, model /
gamma text I/O, MT19937 start models, alpha Newton) is bit/byte exact against
the oracle and the reference-written golden files.  No compute kernels run here."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"
SEEDS = {"k5": (1001, 2.5), "k20": (2002, 2.5), "k20_fixed": (3003, 0.1), "k50": (4004, 2.5)}


def test_library_exports_every_declared_symbol(oni):
    header = (ROOT / "include" / "onilda.h").read_text()
    names = sorted(set(re.findall(r"\b(onilda_[a-z0-9_]+)\s*\(", header)))
    assert len(names) > 40
    L = oni.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert b"sm_100a" in L.onilda_version()


def test_no_oracle_in_product():
    """The product must not link or load anything under oracle/."""
    out = subprocess.run(["ldd", str(ROOT / "open-network-insight_b200" / "lib" / "libonilda.so")],
                         capture_output=True, text=True).stdout
    assert "oracle" not in out
    for p in (ROOT / "open-network-insight_b200").rglob("*"):
        if p.suffix in (".c", ".cu", ".cuh", ".h", ".py") and p.is_file():
            txt = p.read_text()
            assert "liboracle" not in txt and "oracle_py" not in txt and "oracle.h" not in txt, p


def test_settings(oni, oracle, tmp_path):
    for case in SEEDS:
        for name in ("settings.txt", "inf-settings.txt"):
            a, b = oni.read_settings(GOLD / case / name), oracle.read_settings(GOLD / case / name)
            for f, _ in a._fields_:
                assert getattr(a, f) == getattr(b, f)
    s = oni.read_settings(GOLD / "k20_fixed" / "settings.txt")
    assert s.estimate_alpha == 0 and s.var_max_iter == 20
    assert np.float32(s.var_converged) == np.float32(1e-6)
    with pytest.raises(oni.OnildaError):
        oni.read_settings(tmp_path / "missing.txt")
    (tmp_path / "bad.txt").write_text("var max iter 20\nnonsense\n")
    with pytest.raises(oni.OnildaError):
        oni.read_settings(tmp_path / "bad.txt")


@pytest.mark.parametrize("case", list(SEEDS))
def test_parser_bit_exact(oni, oracle, case):
    c = oni.Corpus.from_dat(GOLD / case / "c.dat")
    off, w, cn, tot, V = oracle.read_dat(GOLD / case / "c.dat")
    po, pw, pc, pt = c.csr()
    assert c.num_terms == V and c.num_docs == len(off) - 1 and c.nnz == len(w)
    assert np.array_equal(po, off) and np.array_equal(pw, w) and np.array_equal(pc, cn) and np.array_equal(pt, tot)
    assert c.tokens == int(cn.sum()) and c.max_len == int(np.diff(off).max())


def test_parser_edge_cases(oni, oracle, tmp_path):
    texts = {
        "empty.dat": "",
        "blank.dat": "\n\n  \n",
        "one.dat": "1 0:1",
        "ragged.dat": "0\n0\n3 1:2\n 2:1\n5:4\n2 0:0 4:1\n0",
        "tabs.dat": "2\t7:3\t9:1\r\n1 2:5\r\n",
        "wide.dat": "1 12345678901:7\n",              # "%10d" stops after ten characters
        "plus.dat": "2 +3:+4 5:6\n",
        "truncated.dat": "3 1:2 2:3",                  # EOF inside a document: stale values are stored
        "big.dat": "1 5:2147483647\n1 5:3\n",
    }
    for name, text in texts.items():
        p = tmp_path / name
        p.write_text(text)
        if name == "wide.dat":
            # the reference (and the fscanf-based oracle) spins forever on the unconsumed '1';
            # the product reports the malformed entry instead
            with pytest.raises(oni.OnildaError):
                oni.Corpus.from_dat(p)
            continue
        off, w, cn, tot, V = oracle.read_dat(p)
        c = oni.Corpus.from_dat(p)
        po, pw, pc, pt = c.csr()
        assert np.array_equal(po, off), name
        assert np.array_equal(pw, w) and np.array_equal(pc, cn) and np.array_equal(pt, tot), name
        assert c.num_terms == V, name
    with pytest.raises(oni.OnildaError):
        oni.Corpus.from_dat(tmp_path / "nope.dat")
    (tmp_path / "neg.dat").write_text("1 -4:2\n")
    with pytest.raises(oni.OnildaError):
        oni.Corpus.from_dat(tmp_path / "neg.dat")


def test_save_dat_matches_createmodel_format(oni, corpusgen, tmp_path):
    g = corpusgen.generate(300, 30, 200, 5, seed=4, hub_tokens=900)
    corpusgen.write_dat(g, tmp_path / "py.dat")
    c = oni.Corpus.from_csr(g.offsets, g.words, g.counts)
    c.save_dat(tmp_path / "c.dat")
    assert (tmp_path / "c.dat").read_bytes() == (tmp_path / "py.dat").read_bytes()
    c.save_dat(tmp_path / "part.dat", 10, 20)
    p = oni.Corpus.from_dat(tmp_path / "part.dat")
    assert p.num_docs == 10 and np.array_equal(p.csr()[1], g.words[g.offsets[10]:g.offsets[20]])
    (tmp_path / "m.dat").write_text("2 0:8 3:5\n1 1:4\n1 2:2\n")          # OniLDACWrapperTest.scala:56-74
    oni.Corpus.from_dat(tmp_path / "m.dat").save_dat(tmp_path / "m2.dat")
    assert (tmp_path / "m2.dat").read_text() == "2 0:8 3:5\n1 1:4\n1 2:2\n"


def test_corpus_from_csr_roundtrip(oni, corpusgen):
    g = corpusgen.generate(500, 30, 200, 5, seed=3, hub_tokens=900)
    c = oni.Corpus.from_csr(g.offsets, g.words, g.counts)
    po, pw, pc, pt = c.csr()
    assert np.array_equal(po, g.offsets) and np.array_equal(pw, g.words) and np.array_equal(pc, g.counts)
    assert c.num_terms == g.num_terms and c.tokens == g.tokens and c.max_len == g.max_len
    assert oni.Corpus.from_csr(g.offsets, g.words, g.counts, num_terms=1000).num_terms == 1000
    with pytest.raises(oni.OnildaError):
        oni.Corpus.from_csr(np.array([0, 2, 1], np.int64), np.zeros(2, np.int32), np.ones(2, np.int32))
    e = oni.Corpus.from_csr(np.array([0], np.int64), np.zeros(0, np.int32), np.zeros(0, np.int32))
    assert e.num_docs == 0 and e.nnz == 0


@pytest.mark.parametrize("case", list(SEEDS))
def test_model_load_is_float_rounded_and_save_is_byte_exact(oni, oracle, case, tmp_path):
    m = oni.Model.load(GOLD / case / "final")
    alpha, lb = oracle.load_model(GOLD / case / "final")
    assert m.alpha == alpha and np.array_equal(m.log_beta, lb)
    assert np.array_equal(m.log_beta, m.log_beta.astype(np.float32).astype(np.float64))   # lda-model.c:225
    m.save(tmp_path / "x")
    oracle.save_model(tmp_path / "y", alpha, lb)
    assert (tmp_path / "x.beta").read_bytes() == (tmp_path / "y.beta").read_bytes()
    assert (tmp_path / "x.other").read_bytes() == (tmp_path / "y.other").read_bytes()


def test_fixed10_formatter_matches_printf(oni, tmp_path):
    rng = np.random.default_rng(0)
    vals = np.concatenate([
        rng.normal(0, 1, 20000), -np.exp(rng.uniform(-30, 7, 20000)), np.exp(rng.uniform(-30, 13, 20000)),
        np.array([0.0, -0.0, -100.0, 1e-11, -1e-11, 5e-11, 4.9999999999e-11, 0.99999999995, 0.99999999994999,
                  123456.78901234565, 899999.99999999995, 9e5, 1e6, 123456789.125, 1e-300, 2.5, 0.00000000005,
                  0.12345678905, 0.12345678915, 1.00000000005, np.nextafter(0.5e-10, 1), np.nextafter(0.5e-10, 0)]),
        np.round(rng.uniform(0, 50, 5000), 10), np.round(rng.uniform(0, 50, 5000), 11),
    ])
    g = vals.reshape(-1, 1).copy()
    oni.save_gamma(tmp_path / "g.txt", g)
    got = (tmp_path / "g.txt").read_text().splitlines()
    want = ["%5.10f" % v for v in vals]
    assert got == want


@pytest.mark.parametrize("case", list(SEEDS))
def test_random_init_reproduces_reference_000(oni, case, tmp_path):
    """MT19937 stream + random_initialize_ss + lda_mle: the reference's 000.beta for the same seed."""
    seed, alpha0 = SEEDS[case]
    ref = oni.Model.load(GOLD / case / "000")
    m = oni.Model.new(ref.num_topics, ref.num_terms)
    m.init_random(seed)
    m.alpha = float(np.float32(alpha0))       # INITIAL_ALPHA comes from a float (lda-estimate.c:425)
    m.save(tmp_path / "000")
    assert (tmp_path / "000.beta").read_bytes() == (GOLD / case / "000.beta").read_bytes()
    assert (tmp_path / "000.other").read_bytes() == (GOLD / case / "000.other").read_bytes()


def test_seeded_init_matches_oracle(oni, oracle):
    c = oni.Corpus.from_dat(GOLD / "k20" / "c.dat")
    off, w, cn, tot, V = oracle.read_dat(GOLD / "k20" / "c.dat")
    m = oni.Model.new(20, V)
    m.init_seeded(c, 777)
    L = oracle.lib()
    L.oracle_mt_seed(777)
    lb = np.zeros((20, V))
    L.oracle_seeded_init(20, V, len(off) - 1, off, w, cn, lb)
    assert np.array_equal(m.log_beta, lb)


def test_opt_alpha_bit_exact(oni, oracle):
    f = oni.lib().onilda_opt_alpha
    f.restype = C.c_double
    f.argtypes = [C.c_double, C.c_int, C.c_int]
    g = oracle.lib().oracle_opt_alpha
    finite = 0
    for D, K in [(100, 20), (10000, 20), (10, 5), (50000, 50), (1, 2)]:
        for per in (1.0, 3.0, 10.0, 50.0, 400.0):
            ss = -D * K * per
            a, b = f(ss, D, K), g(ss, D, K)
            assert a == b or (np.isnan(a) and np.isnan(b)), (ss, D, K, a, b)
            finite += int(np.isfinite(a))
    assert finite >= 15


def test_cli_usage_and_errors(oni, tmp_path):
    r = subprocess.run([str(oni.LDA_BIN)], capture_output=True, text=True)
    assert r.returncode == 0
    assert r.stdout.splitlines()[0].startswith("usage : lda est [alpha] [k] [settings] [#processes] [data]")
    r = subprocess.run([str(oni.LDA_BIN), "est", "2.5", "20", str(tmp_path / "none.txt"), "1", "x", "random", "o"],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "settings" in r.stderr
    # a second copy started by an MPI launcher does nothing
    r = subprocess.run([str(oni.LDA_BIN), "est"], capture_output=True, text=True, env={"PMI_RANK": "3"})
    assert r.returncode == 0 and r.stdout == ""


def test_compute_fails_loudly_without_gpu(oni):
    if oni.device_count() > 0:
        pytest.skip("a GPU is present")
    c = oni.Corpus.from_dat(GOLD / "k5" / "c.dat")
    m = oni.Model.load(GOLD / "k5" / "000")
    with pytest.raises(oni.OnildaError) as e:
        oni.EM(c, m, oni.Settings.default())
    assert "no CPU path" in str(e.value)
    with pytest.raises(oni.OnildaError):
        oni.Scorer(np.ones((2, 3)) / 3, np.ones((4, 3)) / 4)
    with pytest.raises(oni.OnildaError):
        oni.infer(m, c, oni.Settings.default())


def _csr_equal(a, b):
    return (a.num_docs == b.num_docs and a.num_terms == b.num_terms and a.tokens == b.tokens and a.max_len == b.max_len
            and all(np.array_equal(x, y) for x, y in zip(a.csr(), b.csr())))


def test_parallel_reader_equals_sequential(oni, tmp_path, monkeypatch):
    """Files in createModel's layout are parsed by several host threads; the result must be what the
    sequential scanf emulation (the definition of the format) gives, and anything else must fall back."""
    rng = np.random.default_rng(11)
    D = 120_000
    lens = rng.integers(0, 12, D)
    lens[::1000] = 400
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    w = rng.integers(0, 999_999_999, off[-1]).astype(np.int32)      # up to nine digits
    cnt = rng.integers(0, 5000, off[-1]).astype(np.int32)
    big = tmp_path / "big.dat"
    oni.Corpus.from_csr(off, w, cnt).save_dat(big)
    assert big.stat().st_size > 6 << 20                              # several chunks
    variants = {"plain": big.read_bytes()}
    variants["no_final_newline"] = variants["plain"][:-1]
    variants["crlf_and_blank_lines"] = variants["plain"].replace(b"\n", b"\r\n", 50).replace(b"\n4", b"\n\n\n4", 30)
    # not canonical somewhere in the middle: a tab, a sign, a pair moved to the next line, a ten-digit count
    mid = len(variants["plain"]) // 2
    nl = variants["plain"].index(b"\n", mid) + 1
    while True:                                                      # a document with at least two entries
        line_end = variants["plain"].index(b"\n", nl)
        line = variants["plain"][nl:line_end]
        if line.count(b" ") >= 2:
            break
        nl = line_end + 1
    def patched(new_line):
        return variants["plain"][:nl] + new_line + variants["plain"][line_end:]
    variants["tab"] = patched(line.replace(b" ", b"\t", 1))
    variants["plus"] = patched(line.replace(b" ", b" +", 1))
    variants["split_line"] = patched(line[::-1].replace(b" ", b"\n", 1)[::-1])
    variants["ten_digits"] = patched(line + b"\n1 7:1234567890")
    for name, data in variants.items():
        p = tmp_path / (name + ".dat")
        p.write_bytes(data)
        monkeypatch.setenv("ONILDA_IO_THREADS", "1")
        seq = oni.Corpus.from_dat(p)
        for threads in ("2", "7"):
            monkeypatch.setenv("ONILDA_IO_THREADS", threads)
            assert _csr_equal(oni.Corpus.from_dat(p), seq), (name, threads)
    assert 1234567890 in seq.csr()[2]                                # the fallback really parsed the odd line


def test_parallel_writer_is_byte_identical(oni, tmp_path, monkeypatch):
    rng = np.random.default_rng(3)
    g = np.exp(rng.uniform(-25, 12, (30_011, 20)))
    g[5, 3] = 1e17                                                   # long "%5.10f" output inside a chunk
    m = oni.Model.new(7, 40_003)
    m.log_beta[...] = -np.exp(rng.uniform(-3, 5, (7, 40_003)))
    out = {}
    for threads in ("1", "5"):
        monkeypatch.setenv("ONILDA_IO_THREADS", threads)
        oni.save_gamma(tmp_path / f"g{threads}", g)
        m.save(tmp_path / f"m{threads}")
        out[threads] = ((tmp_path / f"g{threads}").read_bytes(), (tmp_path / f"m{threads}.beta").read_bytes())
    assert out["1"] == out["5"]
    first = out["1"][0].split(b"\n", 1)[0].decode()
    assert first == " ".join("%5.10f" % v for v in g[0])
    assert out["1"][1].split(b"\n", 1)[0].decode() == "".join(" %5.10f" % v for v in m.log_beta[0])


def test_java_double_layout(oni, oracle):
    """The score columns of <type>_results.csv are Scala Doubles printed by mkString, i.e. Double.toString: the
    documented examples of its layout, the doubles of OniLDACWrapperTest.scala, and random values against the
    restatement (which takes its digits from Python's shortest repr)."""
    known = {1.0: "1.0", 0.05: "0.05", 100.0: "100.0", 1234567.0: "1234567.0", 1.0e7: "1.0E7", 12345678.0: "1.2345678E7",
             0.001: "0.001", 0.00099: "9.9E-4", 1e-4: "1.0E-4", 6.379973716629045E-4: "6.379973716629045E-4",
             0.010333787125924218: "0.010333787125924218", 0.9993620026283371: "0.9993620026283371", 0.0: "0.0",
             4.9e-324: "4.9E-324", 1.7976931348623157e308: "1.7976931348623157E308", -2.5: "-2.5", 3.0e-10: "3.0E-10",
             float("inf"): "Infinity", 123456.789: "123456.789", 0.5: "0.5", 2e-3: "0.002"}
    for v, want in known.items():
        assert oni.java_double(v) == want == oracle.java_double_str(v), v
    rng = np.random.default_rng(9)
    vals = np.concatenate([rng.random(3000), np.exp(rng.uniform(-40, 20, 3000)), rng.random(500) * 1e-3, rng.random(500) * 2e7])
    for v in vals:
        got = oni.java_double(float(v))
        assert got == oracle.java_double_str(float(v)) and float(got.replace("E", "e")) == v


def test_results_file_writer(oni, oracle, tmp_path):
    """flow_results.csv (two score columns) and dns_results.csv (one): the caller's rows, then the scores."""
    rng = np.random.default_rng(10)
    rows = [f"2016-07-08 0{i % 10}:11:2{i % 7},10.0.{i % 5}.{i},172.16.0.{i % 9},{i * 37 % 65536},80,TCP,{i}_word,-1_{i}_word" for i in range(200)]
    a, b = rng.random(200) ** 8, rng.random(200) ** 8 * 1e-3
    oni.results_write(tmp_path / "flow_results.csv", rows, a, b, ",")
    assert (tmp_path / "flow_results.csv").read_text().splitlines() == oracle.results_lines(rows, a, b, ",")
    oni.results_write(tmp_path / "dns_results.csv", rows[:50], a[:50], None, "\t")
    assert (tmp_path / "dns_results.csv").read_text().splitlines() == oracle.results_lines(rows[:50], a[:50], None, "\t")
    oni.results_write(tmp_path / "empty.csv", [], [], None, ",")
    assert (tmp_path / "empty.csv").read_text() == ""
    assert not list(tmp_path.glob("*.tmp"))


def test_jni_shim_compiles(tmp_path):
    """The JNI shim has no JDK to build against in this image; it is at least compiled (-Wall -Wextra -Werror, and
    linked against libonilda so that every entry point it calls exists with that signature) against a
    declaration-only subset of jni.h kept in tests/jni_stub/."""
    import subprocess
    pkg = ROOT / "open-network-insight_b200"
    out = tmp_path / "libonilda_jni_check.so"
    r = subprocess.run(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-fPIC", "-shared", "-I", str(ROOT / "tests" / "jni_stub"),
                        "-I", str(ROOT / "include"), str(pkg / "jni" / "onilda_jni.c"), "-L", str(pkg / "lib"), "-lonilda",
                        "-Wl,--no-undefined", "-o", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    syms = subprocess.run(["nm", "-D", "--defined-only", str(out)], capture_output=True, text=True).stdout
    for name in ("create", "score", "scoreFlowMin", "topK", "destroy", "flowWords", "flowResults", "resultsRows", "corpusFromTriples"):
        assert f"Java_org_opennetworkinsight_OniGpuScorer_00024_{name}" in syms
