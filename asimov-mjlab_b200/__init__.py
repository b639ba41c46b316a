"""B200-native batched rigid-body step backend (drop-in beneath mjlab's Simulation).

Public surface mirrors the five names mjlab uses from its physics dependency
(reference src/mjlab/sim/sim.py:116,120,143,146): `put_model`, `put_data`,
`step`, `forward`, `Model`/`Data`, plus the `Simulation` boundary class.
"""
from .hostmodel import HostData, HostModel  # noqa: F401
from .mjcf import Spec  # noqa: F401
