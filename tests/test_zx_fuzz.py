"""A short randomised sweep on every GPU test run (tools/fuzz_parity.py: random picture
sizes, layouts, maps, flags through the C-ABI against the checker, bit-exact)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("what,extra", [("tdlt", []), ("tdlt", ["--max-frames", "64"]), ("pix", []),
                                        ("dering", []), ("coding", [])])
def test_random_cases_are_bit_exact(what, extra):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), "--what", what,
                        "--seconds", "6", "--seed", "101"] + extra, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "all bit-exact" in r.stdout, r.stdout[-500:]
