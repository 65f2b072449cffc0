"""The reference's OWN Catch2 tests (tests/convert, tests/simple, tests/PDFs of the reference checkout), compiled unmodified where they
lie (`make examples`: one executable per file, as upstream's goofit_catch_test() builds them) and linked against this repo's
libraries. Each test fits or evaluates a PDF on toy data and CHECKs the reference's own expectations; the process exit status is the
number of failed assertions. tests/ref_catch/ stands in for Catch2, which upstream carries as an empty submodule."""
import glob
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "examples", "bin", "reftests")
TESTS = sorted(os.path.relpath(p, BIN) for p in glob.glob(os.path.join(BIN, "*", "*")) if os.access(p, os.X_OK))


def test_suite_was_built():
    if not os.path.isdir(BIN):
        pytest.skip("examples/bin/reftests not built (needs the reference checkout at build time)")
    assert len(TESTS) >= 20, TESTS


@pytest.mark.parametrize("name", TESTS)
def test_reference_catch_test(name):
    r = subprocess.run([os.path.join(BIN, name)], capture_output=True, text=True, timeout=900)
    tail = (r.stdout[-1500:] + "\n" + r.stderr[-3000:])
    assert r.returncode == 0, tail
    m = re.search(r"(\d+) test cases, (\d+) assertions, (\d+) failed", r.stderr)
    assert m, tail
    assert int(m.group(1)) >= 1 and int(m.group(2)) >= 1 and int(m.group(3)) == 0, tail
