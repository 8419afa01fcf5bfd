"""One process per GPU (the layout bench.py --gpus N uses): the sharded sweep over
CUDA IPC + NVLink must equal the unsharded sweep bit for bit.  Skips itself on a box
with fewer than 2 GPUs; tests/test_sharded_gpu.py covers the same protocol with all
ranks in one process."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 4, 8])
def test_one_process_per_gpu_equals_unsharded(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {torch.cuda.device_count()}")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "mp_shard_check.py"), str(1 << 18), "12"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MP_SHARD_CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_ipmcmc_over_ranks_with_the_c_abi_exchange():
    """`infer :ipmcmc` with whole nodes per GPU (anglican_b200/ipmcmc.py): the node
    log-marginals cross ranks through apop_ipmcmc_exchange_logz (NCCL allgather inside the
    C ABI, ipmcmc.cljc:183-187), retained paths point to point; 2 processes, 8 nodes."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip(f"needs 2 GPUs, this box has {torch.cuda.device_count()}")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tools", "bench_configs.py"), "c4dist", "--nodes", "8",
           "--particles", str(1 << 14), "--T", "30", "--steps", "2", "--warmup", "1", "--width", "2"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    import json
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert "apop_ipmcmc_exchange_logz" in d["log_z_exchange"]
    zs = d["node_log_z"]
    assert len(zs) == 2 and len(zs[0]) == 8 and all(abs(z) < 1e6 for z in zs[0])
