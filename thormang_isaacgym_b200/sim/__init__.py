from .fake_sim import FakeSim  # noqa: F401
