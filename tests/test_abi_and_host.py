"""CPU-side checks: the C-ABI library loads and exports every symbol that
include/b200ctc.h declares; host-side helpers; loud failure without a GPU."""
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b200ctc.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ctc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from espnet1_b200 import _lib
    from espnet1_b200.build import build_library

    build_library()
    h = _lib.lib()
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(h, n), f"{n} declared in b200ctc.h but not exported"
    assert sorted(_lib.SIGNATURES) == names  # ctypes table mirrors the header 1:1
    assert h.b200ctc_abi_version() == 6
    assert h.b200ctc_loss_workspace_bytes(64, 500, 5000, 150) > 64 * 500 * 301 * 8


def test_beam_position_argument_checks_need_no_gpu():
    """`b200ctc_beam_position` validates its plan before it touches the device: the error paths
    (and the ctypes mirror of `b200ctc_beam_plan`, whose size the loader checks) run anywhere."""
    import ctypes

    from espnet1_b200 import _lib

    h = _lib.lib()
    assert h.b200ctc_beam_plan_bytes() == ctypes.sizeof(_lib.BeamPlan)
    assert h.b200ctc_beam_position(None, None, None, 0, 1, None) == -1
    assert b"null plan" in h.b200ctc_last_error()
    plan = _lib.BeamPlan()
    plan.U, plan.W, plan.V, plan.T, plan.K_full = 2, 3, 50, 40, 9
    assert h.b200ctc_beam_position(ctypes.byref(plan), None, None, 0, 1, None) == -1
    assert b"K_full" in h.b200ctc_last_error()
    plan.K_full, plan.snum = 0, 5  # a pre-beam without a full scorer to take it from
    assert h.b200ctc_beam_position(ctypes.byref(plan), None, None, 0, 1, None) == -1
    assert b"pre-beam" in h.b200ctc_last_error()
    plan.snum, plan.yseq_stride = 0, 8
    assert h.b200ctc_beam_position(ctypes.byref(plan), None, None, 0, 8, None) == -1
    assert b"prefix buffers" in h.b200ctc_last_error()
    assert h.b200ctc_beam_position(ctypes.byref(plan), None, None, 0, 1, None) == -1
    assert b"null buffer" in h.b200ctc_last_error()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU error path")
def test_no_gpu_means_error_not_fallback():
    from espnet1_b200 import _lib, functional as F
    from espnet1_b200.espnet2.asr.ctc import CTC

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        F.log_softmax(torch.randn(2, 5))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        CTC(5, 4)(torch.randn(1, 3, 4), torch.tensor([3]), torch.tensor([[1]]), torch.tensor([1]))
    assert _lib.lib().b200ctc_device_info(None, None, None) != 0
    assert b"no CUDA device" in _lib.lib().b200ctc_last_error()


def test_product_never_imports_oracle():
    bad = []
    for dp, _, fs in os.walk(os.path.join(ROOT, "espnet1_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M):
                    bad.append(f)
    assert not bad, bad


def test_shard_range_partitions():
    from espnet1_b200.sharded import shard_range

    for n, w in ((1024, 8), (10, 4), (3, 8)):
        spans = [shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


def test_sharded_loss_world2_gloo(tmp_path):
    """world_size 2 over gloo: reduced loss == single-process loss over the union."""
    script = tmp_path / "w.py"
    script.write_text(
        "import os, sys, torch, torch.distributed as dist\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "from espnet1_b200.sharded import reduce_sharded_loss, shard_range\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "g = torch.Generator().manual_seed(0)\n"
        "nll = torch.rand(11, generator=g) * 100\n"
        "lo, hi = shard_range(11, r, w)\n"
        "loss, wt = reduce_sharded_loss(nll[lo:hi])\n"
        "ref = nll.sum() / 11\n"
        "assert abs(float(loss) - float(ref)) < 1e-6 * float(ref), (float(loss), float(ref))\n"
        "assert float(wt) == 11\n"
        "print('rank', r, 'ok')\n")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
         "--master-addr", "127.0.0.1", "--master-port", "29613", str(script)],
        capture_output=True, text=True, timeout=240, env=env)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("ok") == 2
