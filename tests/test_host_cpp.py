"""The C++ host mirror of the reference's calling sequence (4c_b200/host/fourc_b200.hpp) compiled as a
stand-alone program against the C-ABI library: builds here (CPU), fails loudly without a device, and on the GPU
reproduces what the ctypes binding computes."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "_build", "host_example")


def build_example():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    lib = os.path.join(ROOT, "4c_b200", "lib")
    cmd = ["g++", "-std=c++17", "-O1", os.path.join(ROOT, "examples", "host_example.cpp"), "-I" + ROOT, "-I" + cuda + "/include",
           "-L" + lib, "-lfourc_b200", "-L" + cuda + "/lib64", "-lcudart", "-Wl,-rpath," + lib, "-Wl,-rpath," + cuda + "/lib64", "-o", EXE]
    subprocess.run(cmd, check=True, capture_output=True, text=True)


def test_host_mirror_compiles_and_links():
    build_example()
    assert os.path.exists(EXE)


def test_host_mirror_fails_loudly_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    build_example()
    r = subprocess.run([EXE, "3"], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_host_mirror_matches_ctypes_binding(fc):
    build_example()
    n = 5
    r = subprocess.run([EXE, str(n)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"ok=1 nnz=(\d+) sum_abs_K=(\S+) sum_abs_f=(\S+)", r.stdout)
    assert m, r.stdout
    assert "missing-state: exception code -1" in r.stdout
    g = fc.Grid(fc.FC_HEX8, (n, n, n))
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
    s.set_material(fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25)
    n_row, n_col, nnz = s.build_graph()
    X = g.X
    u = 0.01 * np.stack([X[:, 0] * X[:, 1], X[:, 1] * X[:, 2], X[:, 2] * X[:, 0]], axis=1).reshape(-1)
    K, f = np.zeros(nnz), np.zeros(n_row)
    s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST)
    assert int(m.group(1)) == nnz
    assert abs(float(m.group(2)) - 2 * np.abs(K).sum()) <= 1e-11 * np.abs(K).sum()
    assert abs(float(m.group(3)) - 2 * np.abs(f).sum()) <= 1e-11 * np.abs(f).sum() + 1e-13 * np.abs(K).sum()
    s.close()


HALO_EXE = os.path.join(ROOT, "examples", "_build", "halo_example")


def build_halo_example():
    os.makedirs(os.path.dirname(HALO_EXE), exist_ok=True)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    lib = os.path.join(ROOT, "4c_b200", "lib")
    cmd = ["g++", "-std=c++17", "-O1", os.path.join(ROOT, "examples", "halo_example.cpp"), "-I" + ROOT, "-I" + cuda + "/include",
           "-L" + lib, "-lfourc_b200", "-L" + cuda + "/lib64", "-lcudart", "-lnccl", "-lpthread", "-Wl,-rpath," + lib,
           "-Wl,-rpath," + cuda + "/lib64", "-o", HALO_EXE]
    subprocess.run(cmd, check=True, capture_output=True, text=True)


def test_nccl_halo_example_compiles_and_links():
    build_halo_example()
    assert os.path.exists(HALO_EXE)


@pytest.mark.gpu
def test_nccl_halo_exchange_two_gpus():
    """fc_solid_halo_exchange with a caller-made ncclComm_t (2 ranks = 2 GPUs of one box, one thread per rank):
    ghosts receive the owners' values bit for bit, both ranks evaluate, the assembled forces are self-equilibrated."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    build_halo_example()
    r = subprocess.run([HALO_EXE], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "HALO OK" in r.stdout, r.stdout + r.stderr


SOLVE_EXE = os.path.join(ROOT, "examples", "_build", "solve_example")


def build_solve_example():
    os.makedirs(os.path.dirname(SOLVE_EXE), exist_ok=True)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    lib = os.path.join(ROOT, "4c_b200", "lib")
    cmd = ["g++", "-std=c++17", "-O1", os.path.join(ROOT, "examples", "solve_example.cpp"), "-I" + ROOT, "-I" + cuda + "/include",
           "-L" + lib, "-lfourc_b200", "-L" + cuda + "/lib64", "-lcudart", "-lnccl", "-lpthread", "-Wl,-rpath," + lib,
           "-Wl,-rpath," + cuda + "/lib64", "-o", SOLVE_EXE]
    subprocess.run(cmd, check=True, capture_output=True, text=True)


def test_two_rank_solve_example_compiles_and_links():
    build_solve_example()
    assert os.path.exists(SOLVE_EXE)


@pytest.mark.gpu
def test_linear_solve_two_gpus_matches_one_rank():
    """fc_solid_linear_solve on a system partitioned over 2 ranks = 2 GPUs (halo import of the search direction, NCCL all-reduce
    per dot product) gives the solution of the same system held by one rank, CG and BiCGStab"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    build_solve_example()
    r = subprocess.run([SOLVE_EXE], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "SOLVE OK" in r.stdout, r.stdout + r.stderr
