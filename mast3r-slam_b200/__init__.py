"""m3s-b200: B200-native (sm_100a) matching + Gauss-Newton backend behind the operator API of
Aedelon/MASt3R-SLAM's ``backends/mpsgraph/kernels.py`` / ``matching.py`` / ``global_opt.py``.

Only the hot path lives here: ``csrc/`` (CUDA kernels + C ABI), ``_lib`` (ctypes loader),
``kernels`` (L1 operator API), ``matching`` / ``global_opt`` (L2 host entry points), ``dist``
(pair / edge sharding over NCCL), ``synth`` (seeded synthetic inputs).
"""
__version__ = "0.1.0"
