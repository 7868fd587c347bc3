"""Regenerate tests/golden/cfg1/: BASELINE.json configs[0] as SURVEY.md 8(d) states it (10,000 documents,
~1 M tokens, V = 5,000, K = 20, seed 20161101, settings.txt defaults, all EM iterations) run by the
UNMODIFIED reference binary (oracle/_ref/lda, single process).

    python tests/golden/make_golden_cfg1.py        # about four minutes of one core

Protocol of SURVEY.md 8(c): a `random` run with a fixed RNG seed writes boot/000.*; the reference is then
started from that 000 (model-start path, float-parsed like every resume) and its likelihood.dat and
final.* are the fixture.  Large text files are stored gzip-compressed, byte for byte.
"""
import gzip
import importlib
import shutil
import sys
import tempfile
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))

import oracle_py  # noqa: E402

cg = importlib.import_module("open-network-insight_b200.corpusgen")

SETTINGS = "var max iter 20\nvar convergence 1e-6\nem max iter 20\nem convergence 1e-6\nalpha estimate\n"


def gz(src: Path, dst: Path):
    with open(src, "rb") as f, gzip.GzipFile(dst, "wb", compresslevel=9, mtime=0) as g:
        shutil.copyfileobj(f, g)


def main():
    oracle_py.build()
    out = HERE / "cfg1"
    out.mkdir(exist_ok=True)
    c = cg.generate(**cg.CONFIGS["cfg1"])
    with tempfile.TemporaryDirectory() as td:
        td = Path(td)
        cg.write_dat(c, td / "c.dat")
        (td / "settings.txt").write_text(SETTINGS)
        # `em max iter 0` runs exactly one iteration: enough to have the reference write 000.* from its own RNG
        (td / "boot-settings.txt").write_text(SETTINGS.replace("em max iter 20", "em max iter 0"))
        t0 = time.time()
        oracle_py.run_ref(["est", 2.5, 20, "boot-settings.txt", 1, "c.dat", "random", "boot"], td, seed=20161101)
        oracle_py.run_ref(["est", 2.5, 20, "settings.txt", 1, "c.dat", "boot/000", "ref"], td)
        secs = time.time() - t0
        shutil.copy(td / "settings.txt", out / "settings.txt")
        gz(td / "c.dat", out / "c.dat.gz")
        gz(td / "boot" / "000.beta", out / "000.beta.gz")
        shutil.copy(td / "boot" / "000.other", out / "000.other")
        for f in ("likelihood.dat", "final.other"):
            shutil.copy(td / "ref" / f, out / f)
        for f in ("final.beta", "final.gamma"):
            gz(td / "ref" / f, out / (f + ".gz"))
    n_iter = len((out / "likelihood.dat").read_text().splitlines())
    (out / "README").write_text(
        f"written by tests/golden/make_golden_cfg1.py: D={c.num_docs} N={c.nnz} T={c.tokens} V={c.num_terms} K=20, "
        f"{n_iter} EM iterations, reference wall time {secs:.0f} s (both runs, one core)\n")
    print((out / "README").read_text())


if __name__ == "__main__":
    main()
