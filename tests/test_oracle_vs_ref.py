"""Pin the C restatement (oracle/liboracle.so) to the UNMODIFIED reference binary
(oracle/_ref/lda, compiled from /root/reference by oracle/Makefile).

These tests need the reference sources, so they run in the build container and
skip on the GPU box; the fixtures they protect are re-checked everywhere by
tests/test_oracle_golden.py.
"""
import filecmp
import shutil

import numpy as np
import pytest

from conftest import needs_reference

pytestmark = needs_reference

SETTINGS = "var max iter 20\nvar convergence 1e-6\nem max iter {em}\nem convergence 1e-6\nalpha {alpha}\n"


def _small_corpus(corpusgen, tmp_path, seed=11, D=300, V=120, K=6):
    c = corpusgen.generate(D, 30, V, K, seed=seed, hub_tokens=1500, hub_frac=0.01, dup_frac=0.05)
    corpusgen.write_dat(c, tmp_path / "c.dat")
    return c


@pytest.mark.parametrize("alpha_mode,em", [("estimate", 6), ("fixed", 4)])
def test_run_em_bytes_identical_to_reference(oracle, corpusgen, tmp_path, alpha_mode, em):
    """ref `random` -> 000.*; then ref and restatement both start from 000 (lda-estimate.c:163-167)."""
    K = 6
    c = _small_corpus(corpusgen, tmp_path)
    (tmp_path / "settings.txt").write_text(SETTINGS.format(em=em, alpha=alpha_mode))
    oracle.run_ref(["est", 2.5, K, "settings.txt", 1, "c.dat", "random", "boot"], tmp_path, seed=4242)
    oracle.run_ref(["est", 2.5, K, "settings.txt", 1, "c.dat", "boot/000", "ref"], tmp_path)

    st = oracle.read_settings(tmp_path / "settings.txt")
    off, w, cn, tot, V = oracle.read_dat(tmp_path / "c.dat")
    assert V == c.num_terms and np.array_equal(off, c.offsets)
    alpha0, lb0 = oracle.load_model(tmp_path / "boot" / "000")
    res = oracle.run_em(off, w, cn, V, alpha0, lb0, st)

    out = tmp_path / "mine"
    out.mkdir()
    oracle.save_model(out / "final", res["alpha"], res["log_beta"])
    oracle.lib().oracle_save_gamma(str(out / "final.gamma").encode(), len(off) - 1, K, res["gamma"])
    with open(out / "likelihood.dat", "w") as f:
        for l, cv in zip(res["lik"], res["conv"]):
            f.write("%10.10f\t%5.5e\n" % (l, cv))
    for name in ("final.beta", "final.other", "final.gamma", "likelihood.dat"):
        assert filecmp.cmp(tmp_path / "ref" / name, out / name, shallow=False), name
    assert len(res["lik"]) == len((tmp_path / "ref" / "likelihood.dat").read_text().splitlines())


def test_multirank_emulation_matches_forked_reference(oracle, corpusgen, tmp_path):
    """Round-robin sharding d = rank; d += nproc with a rank-ordered Allreduce (lda-estimate.c:201-252)."""
    K = 5
    _small_corpus(corpusgen, tmp_path, seed=5, D=200, V=80, K=K)
    (tmp_path / "settings.txt").write_text(SETTINGS.format(em=3, alpha="estimate"))
    oracle.run_ref(["est", 2.5, K, "settings.txt", 1, "c.dat", "random", "boot"], tmp_path, seed=7)
    oracle.run_ref(["est", 2.5, K, "settings.txt", 3, "c.dat", "boot/000", "ref3"], tmp_path, nproc=3)
    st = oracle.read_settings(tmp_path / "settings.txt")
    off, w, cn, tot, V = oracle.read_dat(tmp_path / "c.dat")
    alpha0, lb0 = oracle.load_model(tmp_path / "boot" / "000")
    res = oracle.run_em(off, w, cn, V, alpha0, lb0, st, nproc=3)
    ref_lik = np.loadtxt(tmp_path / "ref3" / "likelihood.dat")[:, 0]
    assert np.array_equal(np.array(["%10.10f" % x for x in res["lik"]]), np.array(["%10.10f" % x for x in ref_lik]))
    out = tmp_path / "mine"
    out.mkdir()
    oracle.save_model(out / "final", res["alpha"], res["log_beta"])
    oracle.lib().oracle_save_gamma(str(out / "final.gamma").encode(), len(off) - 1, K, res["gamma"])
    for name in ("final.beta", "final.other", "final.gamma"):
        assert filecmp.cmp(tmp_path / "ref3" / name, out / name, shallow=False), name


@pytest.mark.parametrize("init", ["random", "seeded"])
def test_init_matches_reference_with_fixed_seed(oracle, corpusgen, tmp_path, init):
    """MT19937 stream + random_initialize_ss / corpus_initialize_ss (lda-model.c:104-144, cokus.c)."""
    K = 4
    _small_corpus(corpusgen, tmp_path, seed=3, D=60, V=50, K=K)
    (tmp_path / "settings.txt").write_text(SETTINGS.format(em=0, alpha="fixed"))
    oracle.run_ref(["est", 0.5, K, "settings.txt", 1, "c.dat", init, "boot"], tmp_path, seed=20161101)
    off, w, cn, tot, V = oracle.read_dat(tmp_path / "c.dat")
    L = oracle.lib()
    L.oracle_mt_seed(20161101)
    lb = np.zeros((K, V))
    if init == "random":
        L.oracle_random_init(K, V, lb)
    else:
        L.oracle_seeded_init(K, V, len(off) - 1, off, w, cn, lb)
    oracle.save_model(tmp_path / "mine000", float(np.float32(0.5)), lb)
    assert filecmp.cmp(tmp_path / "boot" / "000.beta", tmp_path / "mine000.beta", shallow=False)
    assert filecmp.cmp(tmp_path / "boot" / "000.other", tmp_path / "mine000.other", shallow=False)


def test_infer_matches_reference(oracle, corpusgen, tmp_path):
    K = 6
    _small_corpus(corpusgen, tmp_path, seed=9, D=120, V=90, K=K)
    (tmp_path / "settings.txt").write_text(SETTINGS.format(em=3, alpha="estimate"))
    shutil.copy("/root/reference/oni-ml/oni-lda-c/inf-settings.txt", tmp_path / "inf-settings.txt")
    oracle.run_ref(["est", 2.5, K, "settings.txt", 1, "c.dat", "random", "m"], tmp_path, seed=1)
    oracle.run_ref(["inf", "inf-settings.txt", "m/final", "c.dat", "inf"], tmp_path)
    st = oracle.read_settings(tmp_path / "inf-settings.txt")
    assert st.var_max_iter == -1
    off, w, cn, tot, V = oracle.read_dat(tmp_path / "c.dat")
    alpha, lb = oracle.load_model(tmp_path / "m" / "final")
    gamma, lik = oracle.infer(off, w, cn, V, alpha, lb, st)
    oracle.lib().oracle_save_gamma(str(tmp_path / "mine-gamma.dat").encode(), len(off) - 1, K, gamma)
    assert filecmp.cmp(tmp_path / "inf-gamma.dat", tmp_path / "mine-gamma.dat", shallow=False)
    mine = "".join("%5.5f\n" % x for x in lik)
    assert mine == (tmp_path / "inf-lda-lhood.dat").read_text()
