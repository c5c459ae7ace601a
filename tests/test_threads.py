"""Two host threads, one solver handle each, running at the same time (what bench.py's two-instance
e2e figure does): error state, side streams and cached device properties are per thread, so the
results must be the oracle's bits regardless of interleaving."""
import threading

import numpy as np
import pytest

from oracle import Oracle
from tests import helpers as H


@pytest.mark.gpu
def test_two_threads_two_handles():
    from fluid_simulation_b200 import ParallelNavierStokes
    cases = [(700, 900, 3, 17, 4), (900, 700, 2, 24, 8)]
    want, errors = [], []
    for X, Y, S, K, kmax in cases:
        o = Oracle(X, Y).set_cavity()
        o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-6, S, K
        o.solve()
        want.append(o)

    def run(i):
        try:
            X, Y, S, K, kmax = cases[i]
            for rep in range(3):
                s = ParallelNavierStokes(X, Y).set_cavity()
                s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-6, S, K
                s.set_sweeps_per_launch(kmax)
                s.migrateSolve()
                H.compare_with(s, want[i], tag=f"thread{i}/rep{rep}")
                s.close()
        except BaseException as ex:  # noqa: BLE001 - reported by the main thread
            errors.append((i, ex))

    th = [threading.Thread(target=run, args=(i,)) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors
