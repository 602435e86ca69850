"""Guards on the built library that need no GPU: it holds sm_100a code only, and the register /
spill budget of the hot kernels stays where it was when they were measured (a refactor that
pushed k_allreduce_twoshot from 200 to 328 bytes of spill stack cost 18 % at 8xB200)."""
import re
import shutil
import subprocess

import pytest

CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"


def _run(*args):
    try:
        return subprocess.run([CUOBJDUMP, *args], capture_output=True, text=True, check=True).stdout
    except (FileNotFoundError, subprocess.CalledProcessError) as e:  # pragma: no cover
        pytest.skip(f"cuobjdump unavailable: {e}")


def test_library_is_sm_100a_only(native_lib):
    from nbdistributed_b200 import capi

    elfs = _run("--list-elf", capi.LIB_PATH)
    archs = set(re.findall(r"sm_(\d+a?)", elfs))
    assert archs == {"100a"}, archs


def test_hot_kernels_keep_their_register_and_spill_budget(native_lib):
    from nbdistributed_b200 import capi

    usage = _run("--dump-resource-usage", capi.LIB_PATH)
    found = {}
    for name, regs, stack in re.findall(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+)", usage):
        found[name] = (int(regs), int(stack))
    budgets = {  # mangled-name fragment -> (max registers, max stack bytes)
        "k_allreduce_twoshotIfLb1ELi1E": (128, 256),
        "k_allreduce_twoshotI13__nv_bfloat16Lb1ELi1E": (128, 256),
        "k_allreduce_oneshotIfLb1ELi1E": (128, 64),
        "k_reduce_scatterI13__nv_bfloat16Lb1ELi1E": (128, 64),
        "k_allreduce_zcIfLb1ELi1E": (128, 64),
        "k_allreduce_zcIfLb1ELi8E": (128, 64),
        "k_reduce_scatter_zcIfLb1ELi1E": (128, 64),
        "k_all_gatherILi1E": (128, 32),
        "k_all_gather_zcILi1E": (128, 32),
        "k_broadcastILi1E": (128, 32),
    }
    for frag, (max_regs, max_stack) in budgets.items():
        hits = [v for k, v in found.items() if frag in k]
        assert hits, f"kernel {frag} not found in the library"
        for regs, stack in hits:
            # 128 registers = one 512-thread CTA per SM, which the grid cap (SM count) relies on
            assert regs <= max_regs, (frag, regs)
            assert stack <= max_stack, f"{frag}: {stack} bytes of local stack (budget {max_stack}): spills in a hot loop"


def test_product_library_holds_tma_and_multimem_code(native_lib):
    """The tile pipeline's copy phases are bulk async copies (TMA: UBLKCP in SASS) and its
    NVSwitch transport is multimem.ld_reduce (LDGMC in SASS) - in the shipped library, not in a
    probe tool."""
    from nbdistributed_b200 import capi

    sass = _run("-sass", capi.LIB_PATH)
    assert sass.count("UBLKCP") >= 50, sass.count("UBLKCP")
    assert "LDGMC.E.ADD.F32x4" in sass and "LDGMC.E.HPADD.BF16x8" in sass
    assert "SYNCS.ARRIVE.TRANS64" in sass  # mbarrier expect_tx of the TMA rings
