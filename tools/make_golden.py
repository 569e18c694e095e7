#!/usr/bin/env python3
"""Generate tests/golden/bdf_time_stepping.json by RUNNING the reference's pure-Python
``BDFTimeStepping`` (/root/reference/source/bdf_time_stepping.py, imported unchanged) through
the step-size patterns of its own test (tests/test_bdf_time_stepping.py:16-17,67-114) and two
solver-style constant-step runs.  Run in the build container only (the reference tree does not
travel to the GPU box); the JSON is committed."""
import json
import os
import sys

sys.path.insert(0, "/root/reference/source")
from bdf_time_stepping import BDFTimeStepping   # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def record(ts, step_sizes=None, max_steps=50):
    rows = []
    while not ts.is_at_end() and len(rows) < max_steps:
        if step_sizes is not None:
            ts.set_desired_next_step_size(step_sizes[ts.step_number])
        ts.update_coefficients()
        rows.append(dict(step=ts.step_number, t=ts.current_time, t_next=ts.next_time, k=ts.get_next_step_size(),
                         k_prev=ts.get_previous_step_size(), a1=list(ts.coefficients(1)),
                         a2=list(ts.coefficients(2)), changed1=ts.coefficients_changed(1),
                         changed2=ts.coefficients_changed(2), n_levels=ts.n_levels(), text=str(ts)))
        ts.advance_time()
    return rows


def main():
    pattern = [1.0, 1.0, 2.0, 2.0, 1.0, 1.0, 1.0]          # tests/test_bdf_time_stepping.py:16
    cases = []
    for order in (1, 2):
        ts = BDFTimeStepping(0.0, 9.0, order=order)
        first = record(ts, pattern)
        ts.restart()
        cases.append(dict(name="variable_order%d" % order, args=[0.0, 9.0], order=order, dt0=0.0, step_sizes=pattern,
                          rows=first, rows_after_restart=record(ts, pattern)))
    for order, dt, end in ((2, 0.01, 0.105), (2, 0.002, 0.02), (1, 0.25, 1.0)):
        ts = BDFTimeStepping(0.0, end, order=order, desired_start_time_step=dt)
        cases.append(dict(name="constant_dt%g_order%d" % (dt, order), args=[0.0, end], order=order, dt0=dt,
                          step_sizes=None, rows=record(ts), rows_after_restart=None))
    out = os.path.join(ROOT, "tests", "golden", "bdf_time_stepping.json")
    with open(out, "w") as f:
        json.dump(dict(source="/root/reference/source/bdf_time_stepping.py via tools/make_golden.py", cases=cases), f,
                  indent=1)
    print("wrote", out)


if __name__ == "__main__":
    main()
