"""CPU: the C-ABI shared library builds for sm_100a, loads without a GPU, and exports exactly the
entry points include/hipe_b200.h declares; the ctypes table mirrors the header."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "hipe_b200.h")


def declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"HIPE_API\s+[\w\s\*]+?\b(hipe_\w+)\s*\(", src)


def test_header_declares_the_path():
    names = declared()
    for need in ("hipe_softargmax_fwd", "hipe_softargmax_bwd", "hipe_l1_loss_fwd", "hipe_joint_loss_fused",
                 "hipe_volume_softmax", "hipe_procrustes_batched", "hipe_joint_location_loss_host",
                 "hipe_procrustes_host", "hipe_version", "hipe_strerror"):
        assert need in names


def test_library_exports_every_declared_symbol():
    from hipe_b200 import _build, _lib
    path = _build.build()
    lib = ctypes.CDLL(path)
    for name in declared():
        assert hasattr(lib, name), "libhipe_b200.so does not export %s" % name
    assert sorted(_lib.SIGNATURES) == sorted(declared())
    assert _lib.load().hipe_version() >= 100
    assert _lib.load().hipe_strerror(-5).decode().startswith("cluster")


def test_library_is_sm100a_only_and_uses_bulk_copies():
    from hipe_b200 import _build
    path = _build.build()
    out = subprocess.run(["cuobjdump", "-lelf", path], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs
    sass = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass            # cp.async.bulk (TMA 1-D) in the streaming kernels
    assert "MUFU.EX2" in sass


def test_no_cpu_path():
    import torch
    from hipe_b200 import ops
    from nets import loss as L
    x = torch.zeros(1, 21 * 4, 4, 4)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.softargmax_forward(x, 21, 4, 4, 4)
    with pytest.raises(RuntimeError, match="no CPU path"):
        L.JointLocationLoss()(x, torch.zeros(1, 63), torch.ones(1, 63))


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "hand-integral-pose-estimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f


def test_launch_plan_host_logic():
    """ops.fused_plan is pure host logic (dtype, geometry, the library's tiling query): the plan
    table of DESIGN.md 3.2 without a GPU."""
    import torch
    from hipe_b200 import ops

    def plan(dtype, n_vol, d):
        x = torch.empty(0, dtype=dtype)
        return ops._fused_plan(x, n_vol, d, d, d)

    assert plan(torch.float32, 1344, 64) == "cluster"       # specialised loops: 4096-byte visits, H | SR
    assert plan(torch.float32, 42, 64) == "cluster"
    assert plan(torch.float32, 2016, 56) == "cluster"       # 3136-byte visits
    assert plan(torch.float32, 672, 96) == "cluster"        # 3.5 MB: fits no cluster, streamed shares on 16 CTAs
    assert plan(torch.float32, 672, 128) == "stream"        # 8 MB: the L2 could not hold one volume per cluster
    assert plan(torch.bfloat16, 672, 96) == "stream"
    assert plan(torch.float32, 2016, 48) == "stream"        # fits, but the general loops lose to the stream launch
    assert plan(torch.float32, 1344, 32) == "cluster"       # 128 KB volumes
    assert plan(torch.float32, 1344, 16) is None            # 16 KB volumes: forward + recompute backward
    assert plan(torch.bfloat16, 2688, 64) == "cluster"
    assert plan(torch.bfloat16, 4032, 56) == "stream"        # the one specialised geometry the stream launch keeps
    assert plan(torch.bfloat16, 2688, 32) == "cluster"     # 64 KB volumes: still ahead of the pair (0.162 vs 0.248 ms)
    assert plan(torch.bfloat16, 2688, 16) is None
    assert plan(torch.float64, 1344, 64) is None
    # the plan query behind "cluster": cluster size, ring slots, stages per share, bytes per stage
    # (cluster, slots, share, stage bytes, lead): fp32 64^3 streams its shares on 8 CTAs, bf16 64^3 and
    # fp32 56^3 stay resident, fp32 96^3 (fits no cluster) streams wrapped 12 KiB stages on 16 CTAs
    assert ops.cluster_plan(torch.empty(0), 1344, 64, 64, 64) == (8, 14, 8, 16384, 21)
    assert ops.cluster_plan(torch.empty(0, dtype=torch.bfloat16), 1344, 64, 64, 64) == (4, 14, 8, 16384, 0)
    assert ops.cluster_plan(torch.empty(0), 672, 56, 56, 56)[3:] == (12544, 0)
    assert ops.cluster_plan(torch.empty(0), 672, 96, 96, 96) == (16, 18, 18, 12288, 30)
    # the tiling query behind "resident": supported geometries report a workspace size, others 0
    lib = __import__("hipe_b200._lib", fromlist=["load"]).load()
    assert lib.hipe_resident_workspace_bytes(0, 1344, 64, 64, 64) > 0
    assert lib.hipe_resident_workspace_bytes(0, 1344, 5, 6, 7) == 0


def test_cluster_plan_invariants_over_a_sweep_of_geometries():
    """hipe_cluster_plan is host-only arithmetic: for every geometry it accepts, the plan must be one
    the kernel's pipeline logic is safe on -- shares cover the volume, the ring holds a share with
    look-ahead (resident form) or the lead exceeds the share (streamed form: the no-cycle condition
    of the first producer's throttle), 16-CTA clusters only where the kernel has 16 exchange slots."""
    import torch
    from hipe_b200 import ops
    seen = {"resident": 0, "streamed": 0, "none": 0}
    for dt in (torch.float32, torch.bfloat16):
        esz = 4 if dt == torch.float32 else 2
        x = torch.empty(0, dtype=dt)
        for D, H, W in [(d, d, d) for d in (8, 16, 24, 32, 40, 48, 56, 64, 72, 80, 96, 112, 128)] + \
                       [(16, 128, 64), (12, 20, 24), (32, 96, 96), (64, 32, 32), (5, 6, 7), (3, 9, 100), (96, 96, 64)]:
            plan = ops.cluster_plan(x, 1344, D, H, W)
            if plan is None:
                seen["none"] += 1
                continue
            C, slots, share, stage_bytes, lead = plan
            row_bytes = W * esz
            assert row_bytes % 16 == 0 and stage_bytes % row_bytes == 0 and 0 < stage_bytes <= 16384, (dt, D, H, W, plan)
            nsv = D * H * W * esz // stage_bytes
            assert nsv * stage_bytes == D * H * W * esz                     # whole stages
            assert 1 <= C <= 16 and share == -(-nsv // C) and nsv >= C, (dt, D, H, W, plan)
            assert slots * stage_bytes <= 227 * 1024
            if lead == 0:
                seen["resident"] += 1
                assert C <= 8 and share + 3 <= slots, (dt, D, H, W, plan)
            else:
                seen["streamed"] += 1
                assert lead > share and slots >= 6, (dt, D, H, W, plan)
                if C > 8:
                    assert C == 16 and dt == torch.float32 and row_bytes == 384, (dt, D, H, W, plan)
            # what the module would launch: the cluster plan only where the L2 bound of the streamed form holds
            if ops._fused_plan(torch.empty(1, dtype=dt), 1344, D, H, W) == "cluster" and lead > 0:
                sms = {8: 120, 16: 112}.get(C, 148)
                assert lead * stage_bytes * sms <= 46e6, (dt, D, H, W, plan)
    assert seen["resident"] >= 8 and seen["streamed"] >= 4 and seen["none"] >= 2, seen
