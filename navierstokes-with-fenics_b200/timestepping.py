"""Variable-step BDF-1/2 bookkeeping with the interface of the reference's
``BDFTimeStepping`` (source/bdf_time_stepping.py:7-186) and its base ``DiscreteTime``
(source/discrete_time.py:29-184).

The reference modules are pure Python and stay usable unchanged: the solvers here only
duck-type the object (``coefficients(1)``, ``get_next_step_size``, ``next_time``, ``n_levels``,
``update_coefficients``, ``advance_time``, ``is_at_end`` ...), so
``from bdf_time_stepping import BDFTimeStepping`` with ``/root/reference/source`` on
``sys.path`` is a drop-in.  This restatement exists because the reference tree does not travel
to the GPU box; ``tests/test_timestepping.py`` pins it step by step against sequences recorded
from the reference class (``tests/golden/bdf_time_stepping.json``, made by
``tools/make_golden.py``), which contain the tables of tests/test_bdf_time_stepping.py:67-114.

With w = k_{n+1}/k_n and W = k_n/k_{n-1} (previous ratio):
  1st derivative, BDF-2: ((1+2w)/(1+w), -(1+w), w^2/(1+w));   BDF-1 / first step: (1, -1[, 0])
  2nd derivative, order 1: (2w/(1+w), -2w, 2w^2/(1+w));       first step: (1, -2, 1)
  2nd derivative, order 2: four-level formula of source/bdf_time_stepping.py:106-118
"""
import math

_SNAP = 0.05      # a step ending within 5 % of its size before the end time is stretched to it


def _next_time(t, k, t_end):
    """source/discrete_time.py:5-26"""
    assert k >= 0.0 and t_end >= t
    return t_end if t + k > t_end - _SNAP * k else t + k


class BDFTimeStepping:
    def __init__(self, start_time, end_time, order=2, desired_start_time_step=0.0):
        assert isinstance(start_time, float) and isinstance(end_time, float) and start_time < end_time
        assert isinstance(desired_start_time_step, float) and desired_start_time_step >= 0.0
        assert isinstance(order, int) and order > 0
        if order > 2:
            raise NotImplementedError()
        self._order = order
        self._t0, self._t_end = start_time, end_time
        self._k0 = _next_time(start_time, desired_start_time_step, end_time) - start_time
        self.restart()

    # ---- DiscreteTime ---------------------------------------------------------------------------
    def restart(self):
        self._t_prev = self._t = self._t0
        self._t_next = _next_time(self._t0, self._k0, self._t_end)
        self._step = 0
        self._ratios = [1.0, 1.0]                       # (w, W)
        self._changed = {1: True, 2: True}
        self._alpha = {1: [1.0, -1.0] + [0.0] * (self._order - 1),
                       2: [1.0, -2.0, 1.0] + [0.0] * (self._order - 1)}

    current_time = property(lambda self: self._t)
    next_time = property(lambda self: self._t_next)
    previous_time = property(lambda self: self._t_prev)
    start_time = property(lambda self: self._t0)
    end_time = property(lambda self: self._t_end)
    step_number = property(lambda self: self._step)
    n_substeps = property(lambda self: 1)

    def is_at_start(self):
        return self._step == 0

    def is_at_end(self):
        return self._t == self._t_end

    def get_next_step_size(self):
        return self._t_next - self._t

    def get_previous_step_size(self):
        return self._t - self._t_prev

    def set_desired_next_step_size(self, next_step_size):
        assert isinstance(next_step_size, float) and next_step_size > 0.0
        self._t_next = _next_time(self._t, next_step_size, self._t_end)

    def set_end_time(self, new_end_time):
        assert isinstance(new_end_time, float) and new_end_time > self._t0 and new_end_time > self._t
        self._t_end = new_end_time
        k = self._k0 if self._step == 0 else self.get_previous_step_size()
        self._t_next = _next_time(self._t, k, self._t_end)

    def advance_time(self):
        assert self._t_next > self._t
        k = self.get_next_step_size()
        self._t_prev, self._t = self._t, self._t_next
        self._step += 1
        self._t_next = _next_time(self._t, k, self._t_end)

    def __str__(self):
        return "step number {0:8d}, current time {1:10.2e}, next step size {2:10.2e}".format(
            self._step, self._t, self.get_next_step_size())

    # ---- BDF coefficients -------------------------------------------------------------------------
    def update_coefficients(self):
        """Nothing changes in the very first step (BDF-1 start-up)."""
        if self._step == 0:
            return
        w = self.get_next_step_size() / self.get_previous_step_size()
        W = self._ratios[0]
        assert math.isfinite(w) and w > 0.0 and W > 0.0
        a1, a2 = self._alpha[1], self._alpha[2]
        same_w = self._ratios[0] == w and self._step > 1
        if self._order == 1:
            if same_w:
                self._changed = {1: False, 2: False}
                return
            self._ratios = [w, W]
            a1[0], a1[1] = 1.0, -1.0
            a2[0], a2[1], a2[2] = 2.0 * w / (1.0 + w), -2.0 * w, 2.0 * w * w / (1.0 + w)
            self._changed = {1: False, 2: True}
            return
        if same_w and self._ratios[1] == W:
            self._changed = {1: False, 2: False}
            return
        if same_w:
            self._ratios[1] = W
            self._changed[1] = False
        else:
            self._ratios = [w, W]
            a1[0], a1[1], a1[2] = (1.0 + 2.0 * w) / (1.0 + w), -(1.0 + w), w * w / (1.0 + w)
            self._changed[1] = True
        a2[0] = 2.0 * w * (1.0 + (2.0 + 3.0 * w) * W) / ((1.0 + w) * (1.0 + W + w * W))
        a2[1] = -2.0 * w * (1.0 + 2.0 * (1.0 + w) * W) / (1.0 + W)
        a2[2] = 2.0 * w ** 2 * (1.0 + W + 2.0 * w * W) / (1.0 + w)
        a2[3] = -2.0 * w ** 2 * (1.0 + 2.0 * w) * W ** 3 / ((1.0 + W) * (1.0 + W + w * W))
        self._changed[2] = True

    def coefficients(self, derivative=1):
        assert derivative in (1, 2)
        return tuple(self._alpha[derivative])

    def coefficients_changed(self, derivative=1):
        assert derivative in (1, 2)
        return self._changed[derivative]

    def n_levels(self, derivative=1):
        assert derivative in (1, 2)
        return len(self._alpha[derivative]) - 1
