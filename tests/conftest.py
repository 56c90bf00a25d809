import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_terminal_summary(terminalreporter):
    """rel-L2 and max-norm relative error of every whole-model parity comparison of the run"""
    try:
        import gpu_util
    except Exception:
        return
    if not gpu_util.REPORT:
        return
    tr = terminalreporter
    tr.write_line("parity report (relative error vs oracle / golden: L2 norm, max norm)")
    worst = {}
    for what, precision, e2, em in gpu_util.REPORT:
        tr.write_line(f"  [{precision}] {what}: rel_l2={e2:.3e} rel_max={em:.3e}")
        w = worst.setdefault(precision, [0.0, 0.0])
        w[0], w[1] = max(w[0], e2), max(w[1], em)
    for p, (a, b) in worst.items():
        tr.write_line(f"  worst [{p}]: rel_l2={a:.3e} rel_max={b:.3e}")


@pytest.fixture(scope="session")
def sd():
    """The shared synthetic LDM weights (same seed as the golden fixtures)."""
    import cases
    from oracle import ldm_oracle as O
    return O.make_state_dict(cases.WEIGHT_SEED)
