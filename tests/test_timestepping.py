"""Host logic: the BDF time-stepping restatement against sequences recorded from the reference
class (tests/golden/bdf_time_stepping.json <- tools/make_golden.py, which contain the pinned
tables of the reference's tests/test_bdf_time_stepping.py:67-114).  Bit-exact."""
import json
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "bdf_time_stepping.json")) as f:
    GOLDEN = json.load(f)


def replay(ts, rows, step_sizes):
    for row in rows:
        assert not ts.is_at_end()
        if step_sizes is not None:
            ts.set_desired_next_step_size(step_sizes[ts.step_number])
        ts.update_coefficients()
        assert ts.step_number == row["step"]
        assert ts.current_time == row["t"] and ts.next_time == row["t_next"]
        assert ts.get_next_step_size() == row["k"] and ts.get_previous_step_size() == row["k_prev"]
        assert list(ts.coefficients(1)) == row["a1"]
        assert list(ts.coefficients(2)) == row["a2"]
        assert ts.coefficients_changed(1) == row["changed1"] and ts.coefficients_changed(2) == row["changed2"]
        assert ts.n_levels() == row["n_levels"]
        assert str(ts) == row["text"]
        ts.advance_time()


@pytest.mark.parametrize("case", GOLDEN["cases"], ids=[c["name"] for c in GOLDEN["cases"]])
def test_matches_reference_sequences(case):
    import nsb200
    ts = nsb200.BDFTimeStepping(*case["args"], order=case["order"], desired_start_time_step=case["dt0"])
    replay(ts, case["rows"], case["step_sizes"])
    if case["rows_after_restart"] is not None:
        assert ts.is_at_end()
        ts.restart()
        replay(ts, case["rows_after_restart"], case["step_sizes"])


def test_reference_pinned_tables():
    """the literal tables of the reference test (second order, first derivative)"""
    import nsb200
    ts = nsb200.BDFTimeStepping(0.0, 9.0, order=2)
    expect = [(1.0, -1.0, 0.0), (3 / 2, -2.0, 1 / 2), (5 / 3, -3.0, 4 / 3), (3 / 2, -2.0, 1 / 2),
              (4 / 3, -3 / 2, 1 / 6), (3 / 2, -2.0, 1 / 2), (3 / 2, -2.0, 1 / 2)]
    for k, a in zip([1.0, 1.0, 2.0, 2.0, 1.0, 1.0, 1.0], expect):
        ts.set_desired_next_step_size(k)
        ts.update_coefficients()
        assert ts.coefficients(1) == a
        ts.advance_time()
    assert ts.is_at_end()
