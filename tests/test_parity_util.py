"""The comparison helpers themselves (CPU): the angle comparison must reject a 2*pi offset away from the cut."""
import math

import pytest
import torch

import parity_util as U


def test_angle_close_accepts_flip_at_the_cut_only():
    want = torch.tensor([0.3, -2.0, U.PI_F32, -U.PI_F32, 3.141592])
    U.angle_close(want.clone(), want)
    # representative flips within 1e-4 of the cut are legitimate (1-ulp atan2 differences)
    got = torch.tensor([0.3, -2.0, -U.PI_F32, U.PI_F32, -3.141592])
    U.angle_close(got, want)
    # ... anywhere else a 2*pi offset is an error, even though it is 0 modulo 2*pi
    with pytest.raises(AssertionError, match="outside"):
        U.angle_close(torch.tensor([0.3 + 2 * math.pi]), torch.tensor([0.3]))
    with pytest.raises(AssertionError, match="outside"):
        U.angle_close(want + 2 * math.pi, want)
    # in range but flipped away from the cut cannot happen for a 2*pi period (|got| <= pi and |want| <= pi - 1e-4 give
    # |got - want| < 2*pi - 1e-4): the range check is what closes the loophole; with an open range the flip check does
    with pytest.raises(AssertionError, match="away from the cut"):
        U.angle_close(torch.tensor([2.0 - 2 * math.pi]), torch.tensor([2.0]), bound=5.0)
    with pytest.raises(AssertionError, match="angular error"):
        U.angle_close(torch.tensor([0.30002]), torch.tensor([0.3]))


def test_close_and_exact_report_the_first_mismatch():
    a = torch.zeros(4, 3)
    b = a.clone()
    b[2, 1] = 1e-3
    with pytest.raises(AssertionError, match=r"first at \[2, 1\]"):
        U.close(a, b, what="x")
    with pytest.raises(AssertionError, match="1 mismatches"):
        U.exact(a, b, "x")
    U.close(a, a.clone())
    U.exact(a, a.clone())
