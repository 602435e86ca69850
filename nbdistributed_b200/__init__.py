"""nbdistributed_b200 — B200-native collective backend for nbdistributed notebooks.

The package holds only the tensor data path that nbdistributed's worker
processes drive (ref:src/nbdistributed/worker.py:135-151 selects the backend,
cell code then calls dist.all_reduce / broadcast / all_gather / reduce_scatter):

* ``csrc/``      hand-written sm_100a kernels + the C ABI (include/nbdist_b200.h)
                 + the c10d::Backend glue torch.distributed loads
* ``capi``       ctypes view of the C ABI (emulated-rank mode for 1-GPU tests/bench)
* ``backend``    ``register()`` -> ``dist.init_process_group("b200")``
* ``worker_init``mirror of worker.py's backend-selection seam
* ``topology``   NVLink peer discovery for process_manager.py

Importing the package is cheap and does not need a GPU; using any collective
without the built native libraries raises (there is no CPU/PyTorch fallback).
"""
__version__ = "0.1.0"
BACKEND_NAME = "b200"
