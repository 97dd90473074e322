"""Multi-GPU parity (-m gpu, needs >= 2 GPUs on the box; skipped otherwise): the partitioned NCCL
path against the oracle, launched the way the driver launches bench.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_partitioned_search_matches_the_oracle(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29610 + world),
           os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert f"MULTI_GPU_OK world={world}" in r.stdout


@pytest.mark.parametrize("world", [2, 4])
def test_cli_gpus_flag_is_the_mpirun_form(world, tmp_path):
    """`b2bfs <src> <file> --gpus N` = the reference's `mpirun -np 2 ./osu_bw <src> <file>` (README.md:19-22):
    N ranks, rank 0 prints the reference's lines and checks the depths against a host BFS."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import importlib
    import numpy as np
    sys.path.insert(0, ROOT)
    import oracle
    b2 = importlib.import_module("distributed-cuda-bfs_b200")
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, 14, 16, 1)
    path = str(tmp_path / "rmat14.txt")
    oracle.write_edge_file(path, n, s, d)
    dump = str(tmp_path / "depth.txt")
    r = subprocess.run([b2.CLI_PATH, "2", path, "--gpus", str(world), "--dump-depth", dump], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Output OK!" in r.stdout and r.stdout.count("Starting queue parallel bfs.") == 1
    assert f"Number of vertices {n}" in r.stdout and f"Number of edges {2 * len(s)}" in r.stdout
    want = oracle.bfs_serial(oracle.build_csr(n, s, d), 2)
    assert np.array_equal(np.loadtxt(dump, dtype=np.int64), want.astype(np.int64))
