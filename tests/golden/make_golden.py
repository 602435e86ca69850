#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the REFERENCE ITSELF (build container only).

The reference has no tests or golden vectors, so the oracle is pinned against
outputs of the reference's own code path, produced here:

    /root/reference/src/nbdistributed/process_manager.py   ProcessManager.start_workers
    /root/reference/src/nbdistributed/communication.py     CommunicationManager.send_to_all
    /root/reference/src/nbdistributed/worker.py            spawned unmodified, backend gloo

Cells are sent exactly the way ``%%distributed`` sends them (``send_to_all("execute", code)``,
ref:src/nbdistributed/magic.py:1084) and results are read back with the worker's ``get_var``
message (ref:src/nbdistributed/worker.py:395-418).  IPython is not installed in this image and
``nbdistributed/__init__.py`` imports the magics, so a throw-away stub of the four IPython names
``magic.py`` imports is put on PYTHONPATH; no reference file is modified or copied.

/root/reference does not exist on the GPU box: only the .npz files travel.  Usage:

    python tests/golden/make_golden.py            # writes tests/golden/ref_gloo_n{2,4,8}.npz
"""
from __future__ import annotations

import os
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src"

IPYTHON_STUB = {
    "IPython/__init__.py": "",
    "IPython/core/__init__.py": "",
    "IPython/core/magic.py": (
        "class Magics:\n    def __init__(self, shell=None):\n        self.shell = shell\n"
        "def magics_class(cls):\n    return cls\n"
        "def _deco(arg=None):\n    if callable(arg):\n        return arg\n    return lambda f: f\n"
        "line_magic = _deco\ncell_magic = _deco\n"),
    "IPython/core/magic_arguments.py": (
        "def argument(*a, **k):\n    return lambda f: f\n"
        "def magic_arguments(*a, **k):\n    return lambda f: f\n"
        "def parse_argstring(fn, line):\n    raise RuntimeError('stub')\n"),
}

# One cell per case.  Every cell leaves `<case>_in` (this rank's input) and `<case>_out`.
CELLS = {
    # BASELINE config 1, ref:README.md:99-111
    "readme_allreduce": "torch.manual_seed(1234 + rank)\nx = torch.randn(100, 100)\nreadme_allreduce_in = x.clone()\n"
                        "dist.all_reduce(x)\nreadme_allreduce_out = x\n",
    # BASELINE config 1, ref:README.md:115-125 (rank-0 broadcast of nn.Linear(10, 10) parameters)
    "linear_weight": "torch.manual_seed(100 + rank)\nmodel = torch.nn.Linear(10, 10)\n"
                     "linear_weight_in = model.weight.data.clone()\nlinear_bias_in = model.bias.data.clone()\n"
                     "for param in model.parameters():\n    dist.broadcast(param.data, src=0)\n"
                     "linear_weight_out = model.weight.data\nlinear_bias_out = model.bias.data\n",
    "ar_bf16": "torch.manual_seed(7 + rank)\nx = torch.randn(4099).to(torch.bfloat16)\nar_bf16_in = x.clone().view(torch.int16)\n"
               "dist.all_reduce(x)\nar_bf16_out = x.view(torch.int16)\n",
    "ar_f16": "torch.manual_seed(8 + rank)\nx = torch.randn(2048).to(torch.float16)\nar_f16_in = x.clone()\n"
              "dist.all_reduce(x)\nar_f16_out = x\n",
    "ar_f64": "torch.manual_seed(9 + rank)\nx = torch.randn(513, dtype=torch.float64)\nar_f64_in = x.clone()\n"
              "dist.all_reduce(x)\nar_f64_out = x\n",
    "ar_i64": "torch.manual_seed(10 + rank)\nx = torch.randint(-2**62, 2**62, (1000,))\nar_i64_in = x.clone()\n"
              "dist.all_reduce(x)\nar_i64_out = x\n",
    "ar_i32": "torch.manual_seed(11 + rank)\nx = torch.randint(-2**31, 2**31 - 1, (777,), dtype=torch.int32)\nar_i32_in = x.clone()\n"
              "dist.all_reduce(x)\nar_i32_out = x\n",
    "ar_max_f32": "torch.manual_seed(12 + rank)\nx = torch.randn(777)\nar_max_f32_in = x.clone()\n"
                  "dist.all_reduce(x, op=dist.ReduceOp.MAX)\nar_max_f32_out = x\n",
    "ar_prod_f32": "torch.manual_seed(13 + rank)\nx = torch.randn(300)\nar_prod_f32_in = x.clone()\n"
                   "dist.all_reduce(x, op=dist.ReduceOp.PRODUCT)\nar_prod_f32_out = x\n",
    "bc_u8": "torch.manual_seed(14 + rank)\nx = torch.randint(0, 256, (1001,), dtype=torch.uint8)\nbc_u8_in = x.clone()\n"
             "dist.broadcast(x, src=world_size - 1)\nbc_u8_out = x\n",
    "ag_f32": "torch.manual_seed(15 + rank)\nx = torch.randn(25)\nag_f32_in = x.clone()\n"
              "ag_f32_out = torch.empty(25 * world_size)\ndist.all_gather_into_tensor(ag_f32_out, x)\n",
    "ar_odd_f32": "torch.manual_seed(16 + rank)\nx = torch.randn(16397)\nar_odd_f32_in = x.clone()\n"
                  "dist.all_reduce(x)\nar_odd_f32_out = x\n",
}
VARS = {
    "readme_allreduce": ["readme_allreduce"], "linear_weight": ["linear_weight", "linear_bias"],
    **{k: [k] for k in CELLS if k not in ("readme_allreduce", "linear_weight")},
}


def _write_stub(root: str) -> None:
    for rel, body in IPYTHON_STUB.items():
        path = os.path.join(root, rel)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "w") as f:
            f.write(body)


def _to_np(t) -> np.ndarray:
    return t.detach().cpu().contiguous().numpy()


def run_world(world: int, out_path: str) -> None:
    from nbdistributed.communication import CommunicationManager
    from nbdistributed.process_manager import ProcessManager

    pm = ProcessManager()
    comm_port = pm.start_workers(world)  # gloo: no CUDA here (ref:src/nbdistributed/worker.py:146-147)
    cm = CommunicationManager(world, comm_port)
    try:
        # workers connect only after `import torch` + init_process_group: poll
        deadline = time.time() + 120
        while True:
            try:
                cm.send_to_all("get_status", {}, timeout=5.0)
                break
            except TimeoutError:
                if time.time() > deadline:
                    raise
        arrays = {}
        for case, code in CELLS.items():
            res = cm.send_to_all("execute", code, timeout=120.0)
            for r in range(world):
                if "error" in res[r]:
                    raise RuntimeError(f"{case} rank {r}: {res[r]['error']}\n{res[r].get('traceback')}")
            for var in VARS[case]:
                ins = cm.send_to_all("get_var", f"{var}_in", timeout=60.0)
                outs = cm.send_to_all("get_var", f"{var}_out", timeout=60.0)
                for r in range(world):
                    arrays[f"{var}__in_r{r}"] = _to_np(ins[r]["value"])
                # every collective recorded here leaves identical bytes on all ranks: check it
                # on the reference's own outputs and store one copy
                out0 = _to_np(outs[0]["value"])
                for r in range(1, world):
                    assert _to_np(outs[r]["value"]).tobytes() == out0.tobytes(), f"{var}: rank {r} differs from rank 0"
                arrays[f"{var}__out"] = out0
        res = cm.send_to_all("sync", {}, timeout=60.0)  # %sync -> dist.barrier() (worker.py:213-215)
        assert all(res[r].get("status") == "synced" for r in range(world))
        arrays["__world__"] = np.array([world])
        np.savez_compressed(out_path, **arrays)
        print(f"wrote {out_path}: {len(arrays)} arrays, {os.path.getsize(out_path)} bytes")
    finally:
        try:
            cm.send_to_all("shutdown", {}, timeout=2.0)  # never answered
        except Exception:
            pass
        cm.shutdown()
        pm.shutdown()


def main() -> None:
    if not os.path.isdir(REF_SRC):
        raise SystemExit(f"{REF_SRC} not found: golden vectors can only be regenerated in the build container")
    with tempfile.TemporaryDirectory() as stub:
        _write_stub(stub)
        extra = os.pathsep.join([stub, REF_SRC])
        os.environ["PYTHONPATH"] = extra + os.pathsep + os.environ.get("PYTHONPATH", "")
        sys.path[:0] = [stub, REF_SRC]
        for world in (2, 4, 8):
            run_world(world, os.path.join(HERE, f"ref_gloo_n{world}.npz"))


if __name__ == "__main__":
    main()
