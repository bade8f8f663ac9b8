"""
dist.PeerFrames (fvsrn_peer_buffer_*, include/fvsrn_b200.h): fused render + gather over peer memory. Needs two GPUs on the box:
one process per GPU under torch.distributed.run (tools/check_peer_frames.py), the frames every rank's DVR kernel wrote into
rank 0's buffer over NVLink are bit-identical to the NCCL gather of the same views / bands and to rank 0 rendering alone.
On a single-GPU box the buffer life cycle is checked within one process (create, open is refused for the creating process by
CUDA, free). No counterpart in the reference (single GPU).
"""
import ctypes as C
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_buffer_life_cycle(built_lib):
    import torch
    assert torch.cuda.is_available()
    L = built_lib
    p = C.c_void_p(); h = C.create_string_buffer(64)
    assert L.fvsrn_peer_buffer_create(1 << 20, C.byref(p), h) == 0 and p.value
    assert any(h.raw)                                           # an IPC handle was written
    from fvsrn_b200.dist import _DevicePointer
    t = torch.as_tensor(_DevicePointer(p.value, (1 << 18,), "<i4"), device="cuda")
    assert not t.any()                                          # created zero-filled
    t.fill_(7); torch.cuda.synchronize()
    assert int(t.sum()) == 7 << 18
    del t
    assert L.fvsrn_peer_buffer_free(p) == 0
    assert L.fvsrn_peer_buffer_create(0, C.byref(p), h) != 0   # size 0 is refused
    assert L.fvsrn_peer_buffer_create(16, None, h) != 0
    assert L.fvsrn_peer_buffer_open(None, C.byref(p)) != 0


def test_peer_frames_equal_nccl_gather_on_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on the box (gpurun --gpus 2)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tools", "check_peer_frames.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert "peer frames: OK" in r.stdout
