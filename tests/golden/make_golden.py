"""Regenerate tests/golden/* by running the UNMODIFIED reference binary
(oracle/_ref/lda, built from /root/reference by `make -C oracle ref`).

    python tests/golden/make_golden.py

Each case directory holds: c.dat, settings.txt, 000.{beta,other} (the start model the
reference wrote with a fixed RNG seed), and the reference's outputs from that start:
final.{beta,other,gamma}, likelihood.dat, random.{likelihood.dat,final.gamma,final.other} (what the
`random` run with that seed wrote itself), plus `lda inf` outputs inf-gamma.dat and
inf-lda-lhood.dat (model = final, settings = inf-settings.txt).
The parity protocol is SURVEY.md section 8(c): both sides start from 000.
"""
import importlib
import shutil
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))

import oracle_py  # noqa: E402

cg = importlib.import_module("open-network-insight_b200.corpusgen")

SETTINGS = "var max iter 20\nvar convergence 1e-6\nem max iter {em}\nem convergence 1e-6\nalpha {alpha}\n"
INF_SETTINGS = "var max iter -1\nvar convergence 1e-6\nem max iter 100\nem convergence 1e-4\nalpha estimate\n"

CASES = {
    # name: (generator kwargs, K, em iters, alpha mode, alpha0, rng seed)
    "k5": (dict(num_docs=120, mean_tokens=25, num_terms=80, num_topics=5, seed=101, hub_tokens=600,
                hub_frac=0.02, dup_frac=0.05), 5, 8, "estimate", 2.5, 1001),
    "k20": (dict(num_docs=160, mean_tokens=40, num_terms=150, num_topics=20, seed=202, hub_tokens=1200,
                 hub_frac=0.02, dup_frac=0.05), 20, 6, "estimate", 2.5, 2002),
    "k20_fixed": (dict(num_docs=100, mean_tokens=30, num_terms=120, num_topics=20, seed=303, hub_tokens=400,
                       hub_frac=0.02, dup_frac=0.05), 20, 4, "fixed", 0.1, 3003),
    "k50": (dict(num_docs=80, mean_tokens=60, num_terms=200, num_topics=50, seed=404, hub_tokens=900,
                 hub_frac=0.02, dup_frac=0.05), 50, 3, "estimate", 2.5, 4004),
}


def main():
    oracle_py.build()
    for name, (gen, K, em, alpha_mode, alpha0, seed) in CASES.items():
        out = HERE / name
        out.mkdir(exist_ok=True)
        with tempfile.TemporaryDirectory() as td:
            td = Path(td)
            c = cg.generate(**gen)
            cg.write_dat(c, td / "c.dat")
            if name == "k5":
                # edge cases the token-based parser must keep: an empty document, a one-entry
                # document, entries spread over several lines, a zero count
                with open(td / "c.dat", "a") as f:
                    f.write("0\n1 3:7\n3 1:2\n 2:1\n5:4\n2 0:0 4:1\n")
            (td / "settings.txt").write_text(SETTINGS.format(em=em, alpha=alpha_mode))
            (td / "inf-settings.txt").write_text(INF_SETTINGS)
            oracle_py.run_ref(["est", alpha0, K, "settings.txt", 1, "c.dat", "random", "boot"], td, seed=seed)
            oracle_py.run_ref(["est", alpha0, K, "settings.txt", 1, "c.dat", "boot/000", "ref"], td)
            oracle_py.run_ref(["inf", "inf-settings.txt", "ref/final", "c.dat", "inf"], td)
            for f in ("c.dat", "settings.txt", "inf-settings.txt"):
                shutil.copy(td / f, out / f)
            for f in ("000.beta", "000.other"):
                shutil.copy(td / "boot" / f, out / f)
            # the `random` run itself (double-precision start, no float round trip through 000.*)
            for f in ("likelihood.dat", "final.gamma", "final.other"):
                shutil.copy(td / "boot" / f, out / ("random." + f))
            for f in ("final.beta", "final.other", "final.gamma", "likelihood.dat"):
                shutil.copy(td / "ref" / f, out / f)
            for f in ("inf-gamma.dat", "inf-lda-lhood.dat"):
                shutil.copy(td / f, out / f)
        print(name, "D", c.num_docs, "N", c.nnz, "V", c.num_terms, "iters",
              len((out / "likelihood.dat").read_text().splitlines()))


if __name__ == "__main__":
    main()
