"""The oracle against (a) fixtures written by the unmodified reference binary
(tests/golden/*, made by tests/golden/make_golden.py) and (b) the exact doubles
in the reference's own OniLDACWrapperTest.scala.  Runs anywhere (no reference,
no GPU needed)."""
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden"
CASES = ["k5", "k20", "k20_fixed", "k50"]


def fmt_model(alpha, lb):
    beta = "".join("".join(" %5.10f" % x for x in row) + "\n" for row in lb)
    other = "num_topics %d\nnum_terms %d\nalpha %5.10f\n" % (lb.shape[0], lb.shape[1], alpha)
    return beta, other


def fmt_gamma(g):
    return "".join(" ".join("%5.10f" % x for x in row) + "\n" for row in g)


@pytest.mark.parametrize("case", CASES)
def test_em_from_000_reproduces_reference_files(oracle, case):
    d = GOLD / case
    st = oracle.read_settings(d / "settings.txt")
    off, w, cn, tot, V = oracle.read_dat(d / "c.dat")
    alpha0, lb0 = oracle.load_model(d / "000")
    res = oracle.run_em(off, w, cn, V, alpha0, lb0, st)
    beta, other = fmt_model(res["alpha"], res["log_beta"])
    assert beta == (d / "final.beta").read_text()
    assert other == (d / "final.other").read_text()
    assert fmt_gamma(res["gamma"]) == (d / "final.gamma").read_text()
    lik = "".join("%10.10f\t%5.5e\n" % (l, c) for l, c in zip(res["lik"], res["conv"]))
    assert lik == (d / "likelihood.dat").read_text()


@pytest.mark.parametrize("case", CASES)
def test_infer_reproduces_reference_files(oracle, case):
    d = GOLD / case
    st = oracle.read_settings(d / "inf-settings.txt")
    off, w, cn, tot, V = oracle.read_dat(d / "c.dat")
    alpha, lb = oracle.load_model(d / "final")
    gamma, lik = oracle.infer(off, w, cn, V, alpha, lb, st)
    assert fmt_gamma(gamma) == (d / "inf-gamma.dat").read_text()
    assert "".join("%5.5f\n" % x for x in lik) == (d / "inf-lda-lhood.dat").read_text()


SEEDS = {"k5": (1001, 2.5), "k20": (2002, 2.5), "k20_fixed": (3003, 0.1), "k50": (4004, 2.5)}


@pytest.mark.parametrize("case", CASES)
def test_random_start_run_reproduces_reference_files(oracle, case):
    """`lda est ... random` with the fixture's RNG seed: MT19937 init + the whole EM, no float round trip."""
    d = GOLD / case
    seed, alpha0 = SEEDS[case]
    st = oracle.read_settings(d / "settings.txt")
    off, w, cn, tot, V = oracle.read_dat(d / "c.dat")
    K = int((d / "000.other").read_text().split()[1])
    L = oracle.lib()
    L.oracle_mt_seed(seed)
    lb = np.zeros((K, V))
    L.oracle_random_init(K, V, lb)
    res = oracle.run_em(off, w, cn, V, float(np.float32(alpha0)), lb, st)
    lik = "".join("%10.10f\t%5.5e\n" % (l, c) for l, c in zip(res["lik"], res["conv"]))
    assert lik == (d / "random.likelihood.dat").read_text()
    assert fmt_gamma(res["gamma"]) == (d / "random.final.gamma").read_text()
    assert fmt_model(res["alpha"], res["log_beta"])[1] == (d / "random.final.other").read_text()


def test_parser_edge_cases_in_k5(oracle):
    """empty doc, single entry, entries split over lines, zero count (appended by make_golden.py)."""
    off, w, cn, tot, V = oracle.read_dat(GOLD / "k5" / "c.dat")
    lens = np.diff(off)
    assert list(lens[-4:]) == [0, 1, 3, 2]
    assert list(w[off[-4]:]) == [3, 1, 2, 5, 0, 4]
    assert list(cn[off[-4]:]) == [7, 2, 1, 4, 0, 1]
    assert list(tot[-4:]) == [0, 7, 7, 1]


def test_readme_format_example(oracle, tmp_path):
    """createModel's lines (OniLDACWrapperTest.scala:56-74)."""
    (tmp_path / "m.dat").write_text("2 0:8 3:5\n1 1:4\n1 2:2\n")
    off, w, cn, tot, V = oracle.read_dat(tmp_path / "m.dat")
    assert list(off) == [0, 2, 3, 4] and list(w) == [0, 3, 1, 2] and list(cn) == [8, 5, 4, 2] and V == 4


# ---- post-LDA normalisation: OniLDACWrapperTest.scala ----

def test_phi_exp_normalise_line(oracle):
    """OniLDACWrapperTest.scala:14-23"""
    phi = oracle.phi_from_logbeta(np.array([[1.0, 2.0, 3.0, 4.0, 5.0]]))
    assert list(phi[:, 0]) == [0.011656230956039607, 0.03168492079612427, 0.0861285444362687,
                               0.23412165725273662, 0.6364086465588308]


GAMMA_LINE = ("0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 "
              "0.0124531442 0.0124531442 0.0124531442 23983.5532262138 0.0124531442 0.0124531442 0.0124531442 0.0124531442 "
              "0.0124531442 0.0124531442 22999.4716800747 0.0124531442")


def test_theta_from_gamma_line(oracle):
    """OniLDACWrapperTest.scala:28-51"""
    g = np.array([[float(x) for x in GAMMA_LINE.split()],
                  [0.0, 0.0, 1.0, -1.0] + [0.0] * 16])
    th = oracle.theta_from_gamma(g)
    small = 2.6505498126219955E-7
    expect = [small] * 20
    expect[11] = 0.5104702996191969
    expect[18] = 0.48952492939114034
    assert list(th[0]) == expect
    assert list(th[1]) == [0.0] * 20


def test_theta_four_docs(oracle):
    """OniLDACWrapperTest.scala:81-115 (rows 1..3 of the gamma file)"""
    rows = [
        "0.0124531442 0.0124531442 45583.9162803520 0.0124531442 24896.2502172987 0.0124531442 0.0124531442 0.0124531442 "
        "38659.6152616495 0.0124531442 0.0124531442 0.0124531442 0.0124531442 64804.4886288180 25162.0192468565 "
        "0.0124531442 0.0124531442 36361.7850838895 0.0124531442 0.0124531442",
        "0.0124531442 0.0124531442 53816.0629371489 0.0124531442 24404.5440320289 0.0124531442 0.0124531442 0.0124531442 "
        "28303.9934577136 0.0124531442 0.0124531442 0.0124531442 0.0124531442 71151.8773513492 26093.1297387104 "
        "0.0124531442 0.0124531442 29142.4672019139 0.0124531442 0.0124531442",
        "0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 "
        "0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 0.0124531442 12.0124531442 "
        "0.0124531442 0.0124531442 0.0124531442 0.0124531442"]
    th = oracle.theta_from_gamma(np.array([[float(x) for x in r.split()] for r in rows]))
    assert th[0][2] == 0.19358837746391266 and th[0][0] == 5.28867235797652E-8 and th[0][13] == 0.27521540117076093
    assert th[1][2] == 0.2310572464680428 and th[1][0] == 5.346710724792232E-8 and th[1][17] == 0.12512208919525633
    assert th[2][15] == 0.9806834414974662 and th[2][0] == 0.0010166609738175624


BETA_20x4 = """-15.3937807051 -536.5105632673 -532.4503365466 -10.8318318078
-532.7002781469 -447.4890465231 -525.1690620213 -13.9180467610
-19.5667065907 -5.4442353782 -8.0252973255 -12.4288410371
-537.3544437825 -15.4983056878 -22.4722439197 -14.7817718408
-13.3791705206 -5.3657121205 -9.6286562391 -10.8083455362
-539.4466499733 -535.9756606126 -122.4199569787 -535.7615756409
-14.9824699328 -532.0369117371 -527.9766850251 -11.1874512926
-15.1889345209 -534.5750685563 -530.5148418392 -13.5374760178
-20.4194054373 -5.3920297886 -9.0613037095 -11.5228207125
-16.7232163605 -533.1029777067 -529.0427509894 -11.7742458954
-15.4319201624 -533.2550983208 -529.1948716024 -10.7487557143
-536.2164790868 -532.7454897639 -528.6852629847 -532.5314047894
-14.9148351018 -535.3169147724 -531.2566880561 -10.9572759378
-19.2062800183 -5.8636196300 -8.3638265558 -10.6807087581
-20.4589237180 -5.2246884025 -8.1727002832 -10.7430483178
-341.6123784319 -339.5204175204 -337.1106321483 -339.3909978330
-534.8483200081 -531.3773307914 -527.3171040363 -17.0420121428
-19.9682608636 -5.4274731739 -8.0755820355 -12.0653946611
-535.7333037656 -532.2623144682 -528.2020876890 -532.0482294927
-18.4350359818 -534.4612736041 -530.4010468817 -11.0784977885"""


def test_phi_word_major_20x4(oracle):
    """OniLDACWrapperTest.scala:121-150 ; dictionary: word 0 = "23.0_7.0_7.0_4.0", word 3 = "-1_23.0_7.0_7.0_4.0" """
    lb = np.array([[float(x) for x in line.split()] for line in BETA_20x4.splitlines()])
    phi = oracle.phi_from_logbeta(lb)
    assert phi.shape == (4, 20)
    assert phi[0][0] == 0.010333787125924218
    assert phi[0][19] == 6.379973716629045E-4
    assert phi[3][0] == 0.9896662128740757
    assert phi[3][19] == 0.9993620026283371


def test_score_restatement_small(oracle):
    """SuspiciousConnectsScoreFunction.scala:11-23 on a hand-checkable case, incl. unknown ids."""
    theta = np.array([[0.5, 0.5], [1.0, 0.0]])
    phi = np.array([[0.2, 0.4], [0.8, 0.6]])
    doc = np.array([0, 1, -1, 0], np.int32)
    word = np.array([0, 1, 1, -1], np.int32)
    s = oracle.score(doc, word, theta, phi)
    assert list(s) == [0.5 * 0.2 + 0.5 * 0.4, 1.0 * 0.8 + 0.0 * 0.6, 0.5 * 0.8 + 0.5 * 0.6, 0.5 * 0.5 + 0.5 * 0.5]
    m = oracle.score_flow_min(doc, word, doc[::-1].copy(), word[::-1].copy(), theta, phi)
    # flow joins the events with the document table (FlowPostLDA.scala:94): an event with an unknown ip on either
    # side is gone from the reference's results; per event it is +inf here.  An unknown WORD stays (uniform row, :100).
    gone = (doc < 0) | (doc[::-1] < 0)
    assert list(m) == [np.inf if g else min(a, b) for a, b, g in zip(s, s[::-1], gone)]
    idx, sc = oracle.topk(np.array([0.3, 0.1, 0.9, 0.1, 0.2]), 0.3, 3)
    assert list(idx) == [1, 3, 4] and list(sc) == [0.1, 0.1, 0.2]
    idx, sc = oracle.topk(np.array([0.3, 0.1, 0.9, 0.1, 0.2]), 0.3, -1)
    assert list(idx) == [1, 3, 4, 0]


# OniLDACWrapperTest.scala:53-75 ("createModel"): document ids are positions in distinctDocument
# (192.168.1.1 -> 0, 10.10.98.123 -> 1, 66.23.45.11 -> 2), word ids come from wordDictionary.
CREATE_MODEL_TRIPLES = ([0, 1, 2, 0], [0, 1, 2, 3], [8, 4, 2, 5])
CREATE_MODEL_LINES = ["2 0:8 3:5", "1 1:4", "1 2:2"]


def test_create_model_golden(oracle):
    off, w, c = oracle.triples_to_csr(*CREATE_MODEL_TRIPLES)
    assert oracle.dat_lines(off, w, c) == CREATE_MODEL_LINES
    # ids that never occur are empty documents; order inside a document is input order
    off, w, c = oracle.triples_to_csr([3, 0, 3, 3], [5, 6, 1, 5], [1, 2, 3, 4], num_docs=5)
    assert list(off) == [0, 1, 1, 1, 4, 4]
    assert list(w) == [6, 5, 1, 5] and list(c) == [2, 1, 3, 4]


def test_pairs_count_restatement(oracle):
    """FlowPreLDA.scala:131-137 on a hand-checkable case: three distinct (ip, word) pairs."""
    d, w, c = oracle.pairs_count([2, 0, 2, 2, 0], [5, 1, 5, 0, 1])
    assert list(d) == [0, 2, 2] and list(w) == [1, 0, 5] and list(c) == [2, 1, 2]


# ---- QuantilesTest.scala: every case
def test_quantiles_golden(oracle):
    all_ones, ones_twos, to_ten = [1.0] * 5, [1.0, 2.0, 1.0, 2.0], [float(i) for i in range(1, 11)]
    assert len(oracle.compute_ecdf([])[0]) == 0
    assert [list(x) for x in oracle.compute_ecdf(all_ones)] == [[1.0], [1.0]]
    assert [list(x) for x in oracle.compute_ecdf(ones_twos)] == [[1.0, 2.0], [0.5, 1.0]]
    v, c = oracle.compute_ecdf(to_ten)
    assert list(v) == to_ten and list(c) == [0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0]
    assert list(oracle.compute_quantiles([], [0.0, 0.5])) == [float("inf")] * 2
    assert list(oracle.compute_quantiles(all_ones, [0.0, 0.5])) == [1.0, 1.0]
    assert list(oracle.compute_quantiles(ones_twos, [0.0, 0.5])) == [1.0, 1.0]
    assert list(oracle.compute_quantiles(ones_twos, [0.0, 0.6, 0.99])) == [1.0, 2.0, 2.0]
    assert list(oracle.compute_quantiles(to_ten, [0.0, 0.5])) == [1.0, 5.0]
    assert list(oracle.compute_quantiles(to_ten, oracle.DECILES)) == to_ten
    assert list(oracle.compute_quantiles(to_ten, oracle.QUINTILES)) == [2.0, 4.0, 6.0, 8.0, 10.0]
    assert oracle.quantile_bin(3.5, [1.0, 2.0, 3.0, 4.0, 5.0]) == 3
    assert oracle.quantile_bin(0.0, [1.0, 2.0, 3.0, 4.0, 5.0]) == 0


# FlowWordCreationTest.scala: (sport, dport) -> (wordPort, sourceWord, destinationWord) with ipkt/ibyt/time bins 4/7/7;
# the first seven with sip < dip, the last seven with dip < sip (the ip pair is the same sorted string in all)
ADJUST_PORT_GOLDEN = [
    (2132, 23, "23.0", "23.0_7.0_7.0_4.0", "-1_23.0_7.0_7.0_4.0"),
    (23, 2132, "23.0", "-1_23.0_7.0_7.0_4.0", "23.0_7.0_7.0_4.0"),
    (8392, 9874, "333333.0", "333333.0_7.0_7.0_4.0", "333333.0_7.0_7.0_4.0"),
    (80, 0, "80.0", "-1_80.0_7.0_7.0_4.0", "80.0_7.0_7.0_4.0"),
    (0, 43, "43.0", "43.0_7.0_7.0_4.0", "-1_43.0_7.0_7.0_4.0"),
    (1024, 80, "111111.0", "111111.0_7.0_7.0_4.0", "111111.0_7.0_7.0_4.0"),
    (0, 0, "0.0", "0.0_7.0_7.0_4.0", "0.0_7.0_7.0_4.0"),
    (3245, 43, "43.0", "43.0_7.0_7.0_4.0", "-1_43.0_7.0_7.0_4.0"),
    (80, 2435, "80.0", "-1_80.0_7.0_7.0_4.0", "80.0_7.0_7.0_4.0"),
    (2354, 2435, "333333.0", "333333.0_7.0_7.0_4.0", "333333.0_7.0_7.0_4.0"),
    (80, 0, "80.0", "-1_80.0_7.0_7.0_4.0", "80.0_7.0_7.0_4.0"),
    (0, 2435, "2435.0", "2435.0_7.0_7.0_4.0", "-1_2435.0_7.0_7.0_4.0"),
    (80, 1024, "111111.0", "111111.0_7.0_7.0_4.0", "111111.0_7.0_7.0_4.0"),
    (0, 0, "0.0", "0.0_7.0_7.0_4.0", "0.0_7.0_7.0_4.0"),
]


def test_adjust_port_golden(oracle):
    for i, (sport, dport, port, sword, dword) in enumerate(ADJUST_PORT_GOLDEN):
        sip, dip = ("10.0.2.115", "172.16.0.107") if i < 7 else ("172.16.0.107", "10.0.2.115")
        r = oracle.adjust_port(sip, dip, dport, sport, 4.0, 7.0, 7.0)
        assert r == [port, "10.0.2.115 172.16.0.107", sword, dword], (i, r)


# ---- DNSWordCreationTest.scala: every case
def test_dns_word_creation_golden(oracle):
    assert oracle.extract_subdomain("123.103.104.10.in-addr.arpa") == ["None", "None", 0.0, 6.0]
    assert oracle.extract_subdomain("services.amazon.com.mx") == ["amazon", "services", 8.0, 4.0]
    assert oracle.extract_subdomain("amazon.com.mx") == ["amazon", "None", 0.0, 3.0]
    assert oracle.extract_subdomain("services.amazon.com") == ["amazon", "services", 8.0, 3.0]
    assert oracle.extract_subdomain("amazon.com") == ["None", "None", 0.0, 2.0]
    assert oracle.string_entropy("abcdefg") == 2.807354922057604
    assert oracle.string_entropy("None") == 2.0 and oracle.string_entropy("") == 0.0


def test_country_codes_match_reference(oracle):
    """The oracle's and the kernel's label lists against the reference's own set, where the reference is present."""
    import re
    root = Path(__file__).resolve().parent.parent
    src = (root / "open-network-insight_b200" / "csrc" / "words.cu").read_text()
    body = "".join(re.findall(r'"([a-z]+)"', src[src.index("kCountry2[] ="):src.index("constexpr int kCountry2Count")]))
    kernel = {body[i:i + 2] for i in range(0, len(body), 2)} | {"krd", ""}
    assert len(body) % 2 == 0 and kernel == oracle.COUNTRY_CODES
    ref = Path("/root/reference/oni-ml/src/main/scala/org/opennetworkinsight/dns/DNSWordCreation.scala")
    if ref.exists():
        text = ref.read_text()
        listed = set(re.findall(r'"([a-z]*)"', text[text.index("val countryCodes = Set("):text.index("def dnsWordCreation")]))
        assert listed == oracle.COUNTRY_CODES


def test_oracle_first_iteration_of_cfg1_fixture(oracle, tmp_path):
    """The C restatement's first E-step from tests/golden/cfg1/000 (10,000 documents, V = 5,000, K = 20) gives the
    first line of the likelihood.dat the unmodified reference wrote (tests/golden/make_golden_cfg1.py)."""
    import gzip
    import shutil
    d = GOLD / "cfg1"
    for name in ("c.dat", "000.beta"):
        with gzip.open(d / (name + ".gz"), "rb") as f, open(tmp_path / name, "wb") as g:
            shutil.copyfileobj(f, g)
    shutil.copy(d / "000.other", tmp_path / "000.other")
    off, w, cn, tot, V = oracle.read_dat(tmp_path / "c.dat")
    alpha, lb = oracle.load_model(tmp_path / "000")
    ref = oracle.e_step(off, w, cn, V, alpha, lb, 20, float(np.float32(1e-6)))
    want = float((d / "likelihood.dat").read_text().split()[0])
    assert abs(ref["lik"] - want) <= 1e-9 * abs(want)
