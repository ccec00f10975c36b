"""The host-side C++ mirror of the reference's operator interface (intercession_b200/host/*.hpp):
compiles as plain C++17 against the C ABI, fails loudly without a GPU, and -- on the GPU box --
reproduces the case-study golden fixture through submit / run_relays / reset_relays."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host", "test_host_dynamo.cpp")
EXE = os.path.join(ROOT, "build", "test_host_dynamo")


@pytest.fixture(scope="module")
def exe(built):
    libdir = os.path.join(ROOT, "intercession_b200")
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", SRC, "-o", EXE, "-L", libdir, "-licx",
                    f"-Wl,-rpath,{libdir}"], check=True)
    return EXE


def test_host_mirror_compiles_and_refuses_without_gpu(exe):
    from intercession_b200 import abi
    if abi.load().icx_device_count() > 0:
        pytest.skip("a CUDA device is present")
    r = subprocess.run([exe, "3"], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_host_mirror_reproduces_case_study(exe):
    r = subprocess.run([exe, "100"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    g = np.load(os.path.join(ROOT, "tests", "golden", "verbatim_case_study.npz"))
    seen = 0
    for line in r.stdout.splitlines():
        w = line.split()
        if w[0] == "tick":
            t = int(w[1])
            eT, eP = g[f"push1_t{t}_T"], g[f"push1_t{t}_P"]
            assert np.abs(np.array(w[3:6], np.float32) - eT["origin"][2]).max() < 1e-4, line
            assert np.abs(np.array(w[7:10], np.float32) - eT["origin"][3]).max() < 1e-4, line
            assert np.abs(np.array(w[11:14], np.float32) - eP["velocity"][3]).max() < 1e-4, line
            seen += 1
        elif w[0] == "callbacks":
            assert int(w[1]) > 10 and int(w[3]) == 0          # the legs' ray collider asks for on_collision; nobody is forked
            assert abs(float(w[6])) > 0.9                     # ground normal is +-y
            assert int(w[-1]) == 1
            seen += 1
        elif w[0] == "projectile":                            # 299.99 + 1/60 > maxLifetime 300: condemned by the first update
            assert float(w[2]) == 299.99 + 1.0 / 60 and int(w[4]) == 1
            seen += 1
        elif w[0] == "lifecycle":                             # ENTITY_CREATED record -> recorded components -> create on the device -> replay -> remove
            assert [int(w[2]), int(w[4]), int(w[6]), int(w[8])] == [1, 1, 1, 1], line
            seen += 1
    assert seen == 6


NCCL_SRC = os.path.join(ROOT, "tests", "host", "test_host_nccl.cpp")
NCCL_EXE = os.path.join(ROOT, "build", "test_host_nccl")


@pytest.fixture(scope="module")
def nccl_exe(built):
    libdir = os.path.join(ROOT, "intercession_b200")
    os.makedirs(os.path.dirname(NCCL_EXE), exist_ok=True)
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I/usr/local/cuda/include", NCCL_SRC, "-o", NCCL_EXE,
                    "-L", libdir, "-licx", "-L/usr/local/cuda/lib64", "-lcudart", "-ldl", "-lpthread", f"-Wl,-rpath,{libdir}"], check=True)
    return NCCL_EXE


def test_nccl_host_compiles_and_refuses_without_gpu(nccl_exe):
    from intercession_b200 import abi
    if abi.load().icx_device_count() > 0:
        pytest.skip("a CUDA device is present")
    r = subprocess.run([nccl_exe], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_cpp_server_runs_config4_and_config5_exchanges_over_nccl(nccl_exe):
    """include/icx_nccl.h from a plain C++ host with its own ncclComm_t (world size 2 when the box has two GPUs, else 1)."""
    r = subprocess.run([nccl_exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "config4 allgather ok" in r.stdout and "config5 send/recv ok applied 1 id_mismatches 0" in r.stdout, r.stdout


POOLS_SRC = os.path.join(ROOT, "tests", "host", "test_host_pools.cpp")
POOLS_EXE = os.path.join(ROOT, "build", "test_host_pools")


@pytest.fixture(scope="module")
def pools_exe(built):
    libdir = os.path.join(ROOT, "intercession_b200")
    os.makedirs(os.path.dirname(POOLS_EXE), exist_ok=True)
    subprocess.run(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", POOLS_SRC, "-o", POOLS_EXE, "-L", libdir, "-licx",
                    f"-Wl,-rpath,{libdir}"], check=True)
    return POOLS_EXE


def test_component_pools_compile_and_refuse_without_gpu(pools_exe):
    from intercession_b200 import abi
    if abi.load().icx_device_count() > 0:
        pytest.skip("a CUDA device is present")
    r = subprocess.run([pools_exe, "64", "2"], capture_output=True, text=True)
    assert r.returncode == 2 and "error" in r.stdout            # pinned rows cannot be allocated: no CPU fallback


@pytest.mark.gpu
def test_component_pools_equal_the_packet_path(pools_exe):
    """Device-shared pools (only edited rows cross PCIe; create / remove on the device) and the packet path that resubmits
    every component every tick give byte-identical Transform / Physics rows after every tick."""
    r = subprocess.run([pools_exe, "4096", "12"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    w = r.stdout.splitlines()[0].split()
    assert w[0] == "pools" and int(w[2]) == int(w[4]) == 4097 and int(w[6]) == 0            # + 1 created - 1 removed + the floor
    assert int(w[8]) == 1                                                                   # one full upload: the first tick
    assert 4097 <= int(w[10]) < 4097 + 12 * 8                                               # afterwards a handful of rows per tick
    assert r.stdout.splitlines()[2].split()[1] == "1"
