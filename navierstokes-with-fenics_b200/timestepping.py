"""Variable-step BDF-1/2 bookkeeping with the interface of the reference's
``BDFTimeStepping`` (source/bdf_time_stepping.py:7-186, source/discrete_time.py:29-184) as far
as the solver classes use it (``coefficients(1)``, ``get_next_step_size``, ``next_time``,
``n_levels``, ``update_coefficients``, ``advance_time``, ``is_at_end``).

The reference module is pure Python and stays usable unchanged: the solvers here only
duck-type the object, so ``from bdf_time_stepping import BDFTimeStepping`` (with
``/root/reference/source`` on ``sys.path``) works as a drop-in.  This restatement exists
because the reference tree does not travel to the GPU box; ``tests/test_timestepping.py``
pins it against the reference's golden tables (tests/test_bdf_time_stepping.py:67-114).

First-derivative coefficients for step ratio w = k_{n+1}/k_n:
  BDF-2: alpha = ((1+2w)/(1+w), -(1+w), w^2/(1+w));  first step (and order 1): (1, -1, 0).
"""
import math


class BDFTimeStepping:
    def __init__(self, start_time, end_time, order=2, desired_start_time_step=0.0):
        assert end_time > start_time and order in (1, 2)
        self._start, self._end, self._order = float(start_time), float(end_time), int(order)
        self._t = float(start_time)
        self._step = 0
        self._k_prev = 0.0
        self._k_desired = float(desired_start_time_step) if desired_start_time_step > 0.0 else 0.0
        self._k_next = self._clip(self._k_desired)
        self._alpha = [1.0, -1.0, 0.0]
        self._omega = None

    # ---- DiscreteTime part --------------------------------------------------------------------
    def _clip(self, k):
        if k <= 0.0:
            return 0.0
        rest = self._end - self._t
        # land exactly on the end time when the next step would (nearly) pass it
        if self._t + k > self._end or abs(self._t + k - self._end) <= 1e-12 * max(1.0, abs(self._end)):
            return rest
        return k

    current_time = property(lambda self: self._t)
    next_time = property(lambda self: self._t + self._k_next)
    previous_time = property(lambda self: self._t - self._k_prev)
    start_time = property(lambda self: self._start)
    end_time = property(lambda self: self._end)
    step_number = property(lambda self: self._step)

    def is_at_start(self):
        return self._step == 0

    def is_at_end(self):
        return self._t >= self._end or abs(self._t - self._end) <= 1e-12 * max(1.0, abs(self._end))

    def get_next_step_size(self):
        return self._k_next

    def get_previous_step_size(self):
        return self._k_prev

    def set_desired_next_step_size(self, k):
        assert k > 0.0 and math.isfinite(k)
        self._k_desired = float(k)
        self._k_next = self._clip(self._k_desired)

    def advance_time(self):
        assert self._k_next > 0.0
        self._t += self._k_next
        self._k_prev = self._k_next
        self._step += 1
        self._k_next = self._clip(self._k_desired)

    def restart(self):
        self._t, self._step, self._k_prev = self._start, 0, 0.0
        self._k_next = self._clip(self._k_desired)
        self._alpha, self._omega = [1.0, -1.0, 0.0], None

    # ---- BDF part ----------------------------------------------------------------------------
    def n_levels(self, derivative=1):
        return self._order if derivative == 1 else self._order + 1

    def update_coefficients(self):
        """No change in the very first step (BDF-1 start-up)."""
        if self._step == 0:
            return
        w = self._k_next / self._k_prev
        assert math.isfinite(w) and w > 0.0
        self._omega = w
        if self._order == 1:
            self._alpha = [1.0, -1.0, 0.0]
        else:
            self._alpha = [(1.0 + 2.0 * w) / (1.0 + w), -(1.0 + w), w * w / (1.0 + w)]

    def coefficients(self, derivative=1):
        assert derivative == 1, "the solvers only use first-derivative coefficients"
        return tuple(self._alpha[:self._order + 1]) if self._order == 2 else (self._alpha[0], self._alpha[1])

    def coefficients_changed(self, derivative=1):
        return True

    def __str__(self):
        return "step number = {0:8d}, current time = {1:10.2e}, next step size = {2:10.2e}".format(
            self._step, self._t, self._k_next)
