"""The reference's OWN acceptance suite, unmodified, against the drop-in package (SURVEY §2
row 16, §7 step 2): ``tests/shared/test_processing.py`` + ``tests/conftest.py`` of
matthewhoung/inference-arena (71 tests), run in a subprocess with ``PYTHONPATH`` pointing at
``inference-arena_b200`` so that ``shared.processing`` is this repository's GPU-backed package.

The two files are staged byte for byte under the git-ignored ``baseline/_ref`` by
``scripts/stage_reference.py`` (``__graft_entry__.build()`` runs it wherever the reference tree
exists); the test is skipped when they are absent.  The pass count is written to
``gpurun_out/reference_suite.txt``.
"""

import os
import re
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu

REPO = Path(__file__).resolve().parent.parent
REF = REPO / "baseline" / "_ref"
SUITE = REF / "tests" / "shared" / "test_processing.py"


@pytest.mark.skipif(not SUITE.exists(), reason="baseline/_ref not staged (scripts/stage_reference.py)")
def test_reference_processing_suite_passes_against_the_drop_in():
    env = dict(os.environ)
    env["PYTHONPATH"] = str(REPO / "inference-arena_b200")  # `shared` = the drop-in, not the reference's src/
    env.pop("IA_EXPERIMENT_YAML", None)
    res = subprocess.run([sys.executable, "-m", "pytest", str(SUITE), "-q", "-p", "no:cacheprovider",
                          "--rootdir", str(REF), "-o", "addopts="],
                         cwd=REF, env=env, capture_output=True, text=True, timeout=900)
    tail = "\n".join(res.stdout.strip().splitlines()[-15:])
    out = REPO / "gpurun_out"
    out.mkdir(exist_ok=True)
    (out / "reference_suite.txt").write_text(res.stdout[-4000:] + "\n" + res.stderr[-2000:])
    assert res.returncode == 0, tail + "\n" + res.stderr[-2000:]
    m = re.search(r"(\d+) passed", res.stdout)
    assert m and int(m.group(1)) >= 71, tail
    # the suite imported THIS package, not the reference's
    probe = subprocess.run([sys.executable, "-c", "import shared.processing as s; print(s.__file__)"],
                           cwd=REF, env=env, capture_output=True, text=True, timeout=300)
    assert "inference-arena_b200" in probe.stdout, probe.stdout + probe.stderr
