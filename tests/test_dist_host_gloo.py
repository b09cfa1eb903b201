"""CPU tests of the N > 1 host logic over torch.distributed `gloo` (world sizes 2 and 4, no GPU): the block-row partition,
the generator's process grid, the receive plan computed by the library and the request / send-list handshake of
build_partition_and_halo, checked by running one halo exchange + distributed SpMV + all-reduce against the global result."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from oracle_lib import Oracle

HERE = os.path.dirname(os.path.abspath(__file__))


def run_gloo(world, n, stencil):
    port = 29600 + (os.getpid() % 300) + world
    with tempfile.TemporaryDirectory() as tmp:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
               "--master-port", str(port), os.path.join(HERE, "gloo_worker.py"), tmp, str(n), str(stencil)]
        env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
        assert p.returncode == 0, p.stdout[-4000:]
        return [dict(np.load(os.path.join(tmp, f"rank{r}.npz"))) for r in range(world)]


@pytest.mark.parametrize("world,n,stencil", [(2, 6, 7), (2, 5, 27), (4, 4, 27)])
def test_halo_plan_and_distributed_spmv_over_gloo(world, n, stencil):
    res = run_gloo(world, n, stencil)
    grid = tuple(int(v) for v in res[0]["grid"])
    assert grid == {2: (2, 1, 1), 4: (2, 2, 1)}[world]
    o = Oracle()
    o.gen_laplace((n, n, n), p=grid, stencil=stencil)
    A = o.scipy(0, "A")
    x = np.random.default_rng(99).standard_normal(A.shape[0])
    y = np.concatenate([d["y"] for d in res])
    np.testing.assert_allclose(y, A @ x, rtol=1e-13, atol=1e-13)
    assert all(abs(float(d["dot"]) - float(x @ x)) <= 1e-12 * float(x @ x) for d in res)
    # the plans are symmetric for this symmetric operator, and the halo is one face (7-pt) / face + edges (27-pt) per neighbour
    for r, d in enumerate(res):
        assert sorted(d["recv_rank"].tolist()) == sorted(d["send_rank"].tolist())
        assert r not in d["recv_rank"].tolist()
    if world == 2:
        assert all(int(d["n_halo"]) == n * n for d in res)
