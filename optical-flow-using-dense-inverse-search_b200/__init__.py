"""B200-native Dense Inverse Search optical flow (hot path of nejcgalof/Optical-Flow-using-Dense-Inverse-Search).

The directory name is fixed by the project layout and is not a valid Python identifier; import it as
``import dis_b200`` (a root-level alias module) or via importlib.import_module.

  csrc/            hand-written sm_100a kernels + the C ABI (include/dis_b200.h) -> libdis_b200.so
  engine.py        ctypes binding of that ABI (no CPU path)
  synth.py         seeded synthetic frame pairs / sequences with known flow (inputs only)
"""
from . import engine, parallel, synth  # noqa: F401
from .engine import Engine, Pool, DisError, default_params, load_library, read_flo, write_flo  # noqa: F401
