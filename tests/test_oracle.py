"""CPU-only: pin oracle/fluid_oracle.c against the reference.

(a) golden digests / raw fields recorded from the reference build (tests/golden/),
(b) the hex spot values SURVEY.md section 8(c) recorded from the reference,
(c) when oracle/_ref is present, the reference itself, live (serial == safeSolve == threaded,
    fluid_cpp/cpp/SerialNavierStokes.cpp:20-67, ThreadedNavierStokes.cpp:35-72).
"""
import numpy as np
import pytest

import oracle
from oracle import Oracle, fnv1a64
from tests.helpers import assert_bit_equal, golden_case_setup, load_fields

ALL = ("u", "v", "p", "u_next", "v_next", "p_next", "rhs")

# SURVEY.md section 8(c) known answers (C hex floats at [x, y]) recorded from the reference.
SURVEY_SPOTS = {
    "cavity41_S1000_K50": {"u": {(20, 20): "-0x1.459c1cp-3"}, "v": {(20, 20): "0x1.774d3p-8"},
                           "p": {(20, 20): "-0x1.15b582p-5", (0, 0): "-0x1.7e75cp-5",
                                 (40, 0): "0x1.830efap-6", (39, 39): "0x1.e7b7acp+1"}},
    "cavity100_S50_K50": {"u": {(50, 50): "-0x1.913c94p-8"}, "p": {(99, 0): "0x1.4d672ap-7"}},
    "cavity300_S10_K100": {"u": {(1, 1): "-0x1.6p-143"}, "v": {(1, 1): "0x1.14p-135"}},
}
CASE_NAMES = [
    "cavity41_S1000_K50", "cavity41_S10_K50", "cavity41_S1_K1", "cavity41_S3_K7",
    "cavity129_S100_K50", "cavity257_S20_K100", "cavity100_S50_K50", "cavity300_S10_K100",
    "cavity67x45_S25_K30", "cavity45x131_S12_K33", "cavity41_f64_S100_K50",
    "cavity100_f64_S20_K50", "general53x37_S20_K30", "general53x37_f64_S20_K30",
]


def test_fnv_known_answers():
    # textbook FNV-1a-64 vectors
    assert fnv1a64(np.frombuffer(b"", np.uint8)) == "cbf29ce484222325"
    assert fnv1a64(np.frombuffer(b"a", np.uint8)) == "af63dc4c8601ec8c"
    assert fnv1a64(np.frombuffer(b"foobar", np.uint8)) == "85944171f73967e8"


@pytest.mark.parametrize("name", CASE_NAMES)
def test_oracle_matches_golden(golden, name):
    rec = golden[name]
    o = golden_case_setup(rec, Oracle).solve()
    for n in ALL:
        assert fnv1a64(getattr(o, n)) == rec["digest"][n], f"{name}:{n}"
    for n, pts in rec["spots"].items():
        for key, hx in pts.items():
            x, y = map(int, key.split(","))
            assert float(getattr(o, n)[x, y]).hex() == hx
    for n, pts in SURVEY_SPOTS.get(name, {}).items():
        for (x, y), hx in pts.items():
            assert float(getattr(o, n)[x, y]) == float.fromhex(hx), f"{name}:{n}[{x},{y}]"
    if "fields_file" in rec:
        want = load_fields(rec)
        for n in ALL:
            assert_bit_equal(getattr(o, n), want[n], f"{name}:{n}")


def test_survey_sums(golden):
    o = golden_case_setup(golden["cavity41_S1000_K50"], Oracle).solve()
    assert abs(o.u.astype(np.float64).sum() - 72.35788106) < 1e-7
    assert abs(o.p.astype(np.float64).sum() - -28.84153472) < 1e-7


def test_stale_corner_rule(golden):
    # p(40,0) != p(40,1) after the 41^2 default run (corner lags one sweep)
    o = golden_case_setup(golden["cavity41_S1000_K50"], Oracle).solve()
    assert float(o.p[40, 0]).hex() == "0x1.830efa0000000p-6"
    assert float(o.p[40, 1]).hex() == "0x1.830ef40000000p-6"


def test_phase_split_equals_solve(golden):
    rec = golden["cavity41_S3_K7"]
    a = golden_case_setup(rec, Oracle).solve()
    b = golden_case_setup(rec, Oracle)
    for _ in range(rec["S"]):
        b.predictor().poisson_sweeps(rec["K"]).corrector()
    for n in ALL:
        assert_bit_equal(getattr(b, n), getattr(a, n), n)


needs_ref = pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("kind", ["serial", "safe", "threaded"])
@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_reference_live_cavity(kind, dtype):
    X, Y, S, K = 41, 41, 60, 50
    o = Oracle(X, Y, dtype=dtype).set_cavity()
    o.num_iterations, o.num_poisson_iterations = S, K
    o.solve()
    got, _ = oracle.ref_solve(kind, X, Y, S=S, K=K, dt=float(np.dtype(dtype).type(0.001)),
                              dtype=dtype, want=ALL)
    for n in ("u", "v", "p"):
        assert_bit_equal(getattr(o, n), got[n], f"{kind}:{n}")


@needs_ref
@pytest.mark.parametrize("name", ["general53x37_S20_K30", "cavity67x45_S25_K30"])
def test_reference_live_general(golden, name):
    rec = golden[name]
    o = golden_case_setup(rec, Oracle)
    got, _ = oracle.ref_solve("serial", rec["X"], rec["Y"], dt=float.fromhex(rec["dt_hex"]),
                              K=rec["K"], S=rec["S"], bcs=o, want=ALL)
    o.solve()
    for n in ALL:
        assert_bit_equal(getattr(o, n), got[n], f"{name}:{n}")
