"""Minimal stand-in for the `gym` package (TEST INFRASTRUCTURE ONLY).

The reference imports `gym.spaces` at `vmas/simulator/environment/environment.py:13`
and subclasses `gym.Env` at `vmas/simulator/environment/gym/gym.py:13`; `gym` is not
installed in this image and cannot be fetched (no network).  This stub provides just
the container classes those two sites touch so the UNMODIFIED reference physics can
run as the parity oracle.  Nothing on the product path imports it.
"""
from . import spaces  # noqa: F401


class Env:
    metadata = {}
    observation_space = None
    action_space = None

    def step(self, action):
        raise NotImplementedError

    def reset(self, **kwargs):
        raise NotImplementedError

    def render(self, *args, **kwargs):
        raise NotImplementedError

    def close(self):
        pass
