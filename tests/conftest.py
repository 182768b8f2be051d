import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


def pytest_terminal_summary(terminalreporter):
    """Parity REPORT (not only assertions): observed deviations of the full-size BASELINE configs from the oracle."""
    from tests.util import PARITY_LOG
    if not PARITY_LOG:
        return
    tr = terminalreporter
    tr.write_sep("=", "parity vs oracle on identical noise (bars: costs 1e-4 relative, update 1e-3 absolute)")
    tr.write_line(f"{'case':58s} {'cost rel':>9s} {'|dbest|':>9s} {'tol used':>9s} {'plain 1e-3':>10s} {'delta':>9s} {'live':>5s}")
    for r in PARITY_LOG:
        delta = "-" if r["delta"] is None else f"{r['delta']:.2e}"
        live = "-" if r["live"] is None else str(r["live"])
        tr.write_line(f"{r['tag'][:58]:58s} {r['cost_rel']:9.2e} {r['best_abs']:9.2e} {r['tol']:9.2e} "
                      f"{'held' if r['best_abs'] < 1e-3 else 'WIDENED':>10s} {delta:>9s} {live:>5s}")
