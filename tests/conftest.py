import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REF = "/root/reference"          # read-only reference tree; absent on the GPU box
FIXTURES = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built_libraries():
    """Build the host library, the oracle and the CUDA library when they are missing."""
    need = [os.path.join(ROOT, "rs-vgaligner_b200", "libvga_host.so"),
            os.path.join(ROOT, "oracle", "_build", "liboracle.so"),
            os.path.join(ROOT, "rs-vgaligner_b200", "libvgaligner_gpu.so")]
    if not all(os.path.exists(p) for p in need):
        subprocess.run(["make", "-j8", "host", "oracle", "gpu"], cwd=ROOT, check=True, stdout=subprocess.DEVNULL)


def gfa_text(nodes, edges):
    """nodes: list of sequences (ids 1..n); edges: list of (from, to) or (from, fo, to, to_o)."""
    out = ["H\tVN:Z:1.0"]
    for i, s in enumerate(nodes):
        out.append("S\t%d\t%s" % (i + 1, s))
    for e in edges:
        if len(e) == 2:
            out.append("L\t%d\t+\t%d\t+\t0M" % e)
        else:
            out.append("L\t%d\t%s\t%d\t%s\t0M" % e)
    return "\n".join(out) + "\n"


def fixture_path(name):
    return os.path.join(FIXTURES, name)
