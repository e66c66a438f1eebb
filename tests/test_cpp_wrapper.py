"""include/AuctionB200.h: the C++ mirror of Auction<S>::solve / AuctionMT<S>::solve.
CPU: it compiles (also against the reference's own headers and Eigen when they are present) and
fails loudly without a device.  GPU: its edges equal the reference-made golden vectors."""
import os
import subprocess

import numpy as np
import pytest

import parity

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "wrapper_demo.cpp")
PKG = os.path.join(ROOT, "auctionalgorithm_b200")
REF = "/root/reference"


def _compile(out, extra=()):
    from auctionalgorithm_b200 import build
    build.build()
    cmd = ["g++", "-std=c++11", "-O1", "-w", f"-I{ROOT}/include", *extra, SRC, "-o", out, f"-L{PKG}", "-lb200lsap",
           "-lpthread", f"-Wl,-rpath,{PKG}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


@pytest.fixture(scope="module")
def demo(tmp_path_factory):
    return _compile(str(tmp_path_factory.mktemp("cpp") / "wrapper_demo"))


def test_wrapper_compiles_and_fails_loudly_without_gpu(demo):
    import torch
    r = subprocess.run([demo, "8", "12", "1"], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0
    else:
        assert r.returncode == 1 and "no CUDA device" in r.stdout


def test_wrapper_compiles_against_reference_headers(tmp_path):
    if not os.path.exists(os.path.join(REF, "AuctionAlgorithm", "Auction.h")):
        pytest.skip("reference tree not present on this machine")
    _compile(str(tmp_path / "wrapper_demo_eigen"), ("-DWITH_REFERENCE_EIGEN", f"-I{REF}", f"-I{REF}/external"))


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["auction_8x12_s1", "auction_200x300_s2", "auction_256x256_s2"])
@pytest.mark.parametrize("rowmajor", [False, True])
def test_wrapper_edges_match_golden(demo, case, rowmajor):
    g = np.load(os.path.join(parity.GOLDEN_DIR, case + ".npz"))
    args = [demo, str(int(g["rows"])), str(int(g["cols"])), str(int(g["seed"]))] + (["rowmajor"] if rowmajor else [])
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    edges = np.array([[float(t) for t in l.split()] for l in lines[:-1]])
    assert np.array_equal(edges[:, 0], np.arange(int(g["rows"])))
    assert np.array_equal(edges[:, 1].astype(np.int64), parity.golden_row_to_col(g))
    assert abs(float(lines[-1].split()[1]) - float(g["objective"])) < 1e-9


@pytest.mark.gpu
def test_wrapper_throws_where_the_reference_hangs(demo):
    r = subprocess.run([demo, "9", "5", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "rows <= cols" in r.stdout


# ---- examples/lsap_b200.cpp: the reference's driver (LSAP.cpp:21-63) on the B200 path ----------------
DRIVER = os.path.join(ROOT, "examples", "lsap_b200.cpp")


def _compile_driver(out, extra=()):
    from auctionalgorithm_b200 import build
    build.build()
    cmd = ["g++", "-std=c++11", "-O1", "-w", f"-I{ROOT}/include", *extra, DRIVER, "-o", out, f"-L{PKG}", "-lb200lsap",
           "-lpthread", f"-Wl,-rpath,{PKG}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def test_driver_compiles_standalone_and_against_reference_headers(tmp_path):
    exe = _compile_driver(str(tmp_path / "lsap_b200"))
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([exe, "8", "12"], capture_output=True, text=True)
        assert r.returncode == 1 and "no CUDA device" in r.stdout
    if os.path.exists(os.path.join(REF, "AuctionAlgorithm", "Auction.h")):
        _compile_driver(str(tmp_path / "lsap_b200_eigen"), ("-DWITH_REFERENCE_EIGEN", f"-I{REF}", f"-I{REF}/external"))


@pytest.mark.gpu
def test_driver_prints_the_reference_lines_and_four_equal_objectives(tmp_path):
    exe = _compile_driver(str(tmp_path / "lsap_b200"))
    r = subprocess.run([exe, "300", "300", "7"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    assert len(lines) == 8
    assert all(l.startswith("command  <solution = Auction") and l.rstrip().endswith("us") for l in lines[0::2])
    objs = [float(l.split(":")[1]) for l in lines[1::2]]
    assert all(l.startswith("objective function value:") for l in lines[1::2])
    # dense, sparse, "MT" dense, "MT" sparse: the same optimum (the matrix has no zero, so the sparse
    # matrix holds every coefficient, as in LSAP.cpp:36-38)
    assert max(objs) - min(objs) <= 300 * (1e-5 / 300) and 290 < objs[0] <= 300


# ---- several GPUs (or several ranks on one GPU) from C++: no Python, no NCCL on the data path -----------------------
@pytest.fixture(scope="module")
def multi_demo(tmp_path_factory):
    from auctionalgorithm_b200 import build
    build.build()
    out = str(tmp_path_factory.mktemp("cpp") / "multi_demo")
    cmd = ["g++", "-std=c++11", "-O1", "-w", f"-I{ROOT}/include", os.path.join(ROOT, "tests", "cpp", "multi_demo.cpp"), "-o", out,
           f"-L{PKG}", "-lb200lsap", "-lpthread", f"-Wl,-rpath,{PKG}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def test_multi_demo_compiles(multi_demo):
    assert os.path.exists(multi_demo)


@pytest.mark.gpu
@pytest.mark.parametrize("case,mode,devices", [("auction_200x300_s2", "sharded", [0, 0]), ("auction_512x768_s1", "sharded_rowmajor", [0, 0, 0]),
                                               ("auction_256x256_s2", "sharded", [0, 0]), ("auction_2048x3072_s1", "sharded", [0, 0, 0, 0]),
                                               ("auction_2048x3072_s1", "sharded", [0, 1]), ("auction_1024x1536_s1", "sharded", [0])])
def test_cpp_sharded_solve_matches_golden_and_single_gpu(multi_demo, case, mode, devices):
    import torch
    if max(devices) >= torch.cuda.device_count():
        pytest.skip("needs more GPUs")
    g = np.load(os.path.join(parity.GOLDEN_DIR, case + ".npz"))
    args = [multi_demo, str(int(g["rows"])), str(int(g["cols"])), str(int(g["seed"])), mode] + [str(d) for d in devices]
    r = subprocess.run(args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr
    lines = r.stdout.strip().splitlines()
    assert lines[-1] == "same_as_single 1"
    edges = np.array([[float(t) for t in l.split()] for l in lines[:-2]])
    assert np.array_equal(edges[:, 1].astype(np.int64), parity.golden_row_to_col(g))
    assert abs(float(lines[-2].split()[1]) - float(g["objective"])) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("devices", [[0], [0, 0], [0, 1]])
def test_cpp_batch_split_over_devices(multi_demo, devices, oracle_lib):
    import torch
    if max(devices) >= torch.cuda.device_count():
        pytest.skip("needs more GPUs")
    batch = 3000
    r = subprocess.run([multi_demo, str(batch), "0", "1", "batch"] + [str(d) for d in devices], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr
    lines = r.stdout.strip().splitlines()
    assert lines[-1] == "same_as_single 1"
    got = np.array([[int(t) for t in l.split()] for l in lines[:-1]], np.int32)
    a = np.stack([oracle_lib.generate(32, 48, 1, problem=b) for b in range(batch)])
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0 and np.array_equal(got, ref)
