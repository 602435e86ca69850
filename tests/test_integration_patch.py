"""INTEGRATION.md §1 applied to the reference's OWN worker.py, then driven by the reference's own
ProcessManager + CommunicationManager (CPU, gloo, N = 2).

The patch is applied textually to a throw-away copy of the three non-IPython modules (nothing of
the reference is copied into this repository; /root/reference only exists in the build
container, so the test skips elsewhere).  The backend string travels through the new seam
(NBDIST_BACKEND -> select_backend -> init_process_group); the README cells then run through the
ZMQ path exactly as `%%distributed` sends them, and the outputs must equal the golden fixtures
that the UNPATCHED reference produced (tests/golden/ref_gloo_n2.npz).
"""
import os
import shutil
import sys
import time

import numpy as np
import pytest

REF_PKG = "/root/reference/src/nbdistributed"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

OLD_INIT = "        dist.init_process_group(backend=backend, rank=rank, world_size=world_size)\n"
NEW_INIT = (
    "        requested = os.environ.get(\"NBDIST_BACKEND\")\n"
    "        backend = b200_init.select_backend(requested, torch.cuda.is_available())\n"
    "        if b200_init.wants_b200(backend):\n"
    "            b200_backend.register()\n"
    "        print(f\"PATCHED-SEAM backend={backend}\", flush=True)\n"
    "        dist.init_process_group(backend=backend, rank=rank, world_size=world_size)\n")
OLD_IMPORT = "import torch.distributed as dist\n"
NEW_IMPORT = OLD_IMPORT + "from nbdistributed_b200 import backend as b200_backend, worker_init as b200_init\n"


def _patched_package(tmp):
    pkg = os.path.join(tmp, "nbdistributed")
    os.makedirs(pkg)
    for name in ("process_manager.py", "communication.py", "worker.py"):
        shutil.copy(os.path.join(REF_PKG, name), os.path.join(pkg, name))
    open(os.path.join(pkg, "__init__.py"), "w").close()  # the reference's __init__ imports the IPython magics
    path = os.path.join(pkg, "worker.py")
    src = open(path).read()
    assert src.count(OLD_INIT) == 1 and OLD_IMPORT in src, "the reference's seam moved: update INTEGRATION.md §1"
    src = src.replace(OLD_INIT, NEW_INIT).replace(OLD_IMPORT, NEW_IMPORT, 1)
    open(path, "w").write(src)
    return pkg


@pytest.mark.skipif(not os.path.isdir(REF_PKG), reason="needs /root/reference (build container only)")
def test_patched_reference_worker_reproduces_the_golden_outputs(tmp_path, monkeypatch):
    _patched_package(str(tmp_path))
    monkeypatch.setenv("PYTHONPATH", os.pathsep.join([ROOT, str(tmp_path), os.environ.get("PYTHONPATH", "")]))
    monkeypatch.setenv("NBDIST_BACKEND", "gloo")  # goes through select_backend, not the hard-coded rule
    monkeypatch.syspath_prepend(str(tmp_path))
    for m in [m for m in sys.modules if m == "nbdistributed" or m.startswith("nbdistributed.")]:
        monkeypatch.delitem(sys.modules, m)
    from nbdistributed.communication import CommunicationManager
    from nbdistributed.process_manager import ProcessManager

    from tests.golden.make_golden import CELLS

    world = 2
    pm = ProcessManager()
    comm_port = pm.start_workers(world)
    cm = CommunicationManager(world, comm_port)
    try:
        deadline = time.time() + 120
        while True:  # workers connect only after import torch + init_process_group
            try:
                cm.send_to_all("get_status", {}, timeout=5.0)
                break
            except TimeoutError:
                if time.time() > deadline:
                    raise
        golden = np.load(os.path.join(ROOT, "tests", "golden", "ref_gloo_n2.npz"))
        for case, var_names in (("readme_allreduce", ["readme_allreduce"]), ("linear_weight", ["linear_weight", "linear_bias"])):
            res = cm.send_to_all("execute", CELLS[case], timeout=120.0)
            assert all("error" not in res[r] for r in range(world)), res
            for var in var_names:
                outs = cm.send_to_all("get_var", f"{var}_out", timeout=60.0)
                for r in range(world):
                    got = outs[r]["value"].detach().cpu().numpy()
                    assert got.tobytes() == golden[f"{var}__out"].tobytes(), f"{var}: rank {r} differs from the unpatched reference"
        res = cm.send_to_all("sync", {}, timeout=60.0)  # %sync -> dist.barrier()
        assert all(res[r].get("status") == "synced" for r in range(world))
    finally:
        try:
            cm.send_to_all("shutdown", {}, timeout=2.0)
        except Exception:
            pass
        cm.shutdown()
        pm.shutdown()
    # the patched seam really ran in the workers
    seen = "".join((p.stdout.read() if p.stdout else "") for p in pm.processes) if hasattr(pm, "processes") else ""
    assert seen == "" or "PATCHED-SEAM backend=gloo" in seen
