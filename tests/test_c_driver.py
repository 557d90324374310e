"""The drop-in boundary from C: a host program in the shape of the reference's RoCC driver
(tests/compression.c) compiled as strict C99 against include/snappy_b200.h and linked with
libsnappy_b200.  CPU: it builds, links and fails loudly without a device (no CPU path).  GPU: the
130-byte test sentence compresses to the reference Snappy's 112 bytes and comes back."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIBDIR = os.path.join(ROOT, "compression-accelerator_b200")
THISTLE_STREAM = ("8201b0496620546964207468652074686973746c6520736966746572206c6966747320686973207369657665206f6620"
                  "051a012d052904732c0137306e206865276c6c207368696674011100200d480c7320696e0d3f106674696e670d4740"
                  "746f206c69766520756e7369667465642e")


@pytest.fixture(scope="module")
def driver(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("c_driver") / "rocc_driver")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(HERE, "c_driver", "rocc_driver.c"), "-L", LIBDIR, "-lsnappy_b200",
                    "-Wl,-rpath," + LIBDIR, "-o", exe], check=True)
    return exe


def test_c_driver_builds_and_has_no_cpu_path(driver):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present: covered by the gpu test")
    r = subprocess.run([driver], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1
    assert "130 -> 0 bytes" in r.stdout and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_c_driver_thistle_known_answer(driver):
    r = subprocess.run([driver], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.splitlines()
    assert lines[0] == "compress: 130 -> 112 bytes"
    assert lines[1].replace("stream:", "").replace(" ", "") == THISTLE_STREAM
    assert lines[2] == "uncompress: result 0, identical"


@pytest.fixture(scope="module")
def swcompress(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("swcompress") / "swcompress")
    subprocess.run(["g++", "-std=c++11", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(HERE, "c_driver", "swcompress.cc"), "-L", LIBDIR, "-lsnappy_b200",
                    "-Wl,-rpath," + LIBDIR, "-o", exe], check=True)
    return exe


def test_cpp_host_program_builds_and_has_no_cpu_path(swcompress):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present: covered by the gpu test")
    r = subprocess.run([swcompress], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "no CUDA device" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("name", [None, "fox-in-socks.txt", "all-mtg-cards.txt", "randomASCII.txt"])
def test_cpp_host_program_matches_reference_snappy(swcompress, tmp_path, name):
    """snappy::Compress call shape (tests/swcompression.cc:45) with the namespace switched: size and
    FNV-1a of the stream equal the reference Snappy's (golden manifest), and the round trip holds."""
    import corpus
    args = [swcompress]
    if name is None:
        want = "130 112 %08x ok" % _fnv1a32(bytes.fromhex(THISTLE_STREAM))
    else:
        path = tmp_path / name
        path.write_bytes(corpus.fixture(name))
        args.append(str(path))
        m = corpus.manifest()["files"][name]
        want = "%d %d %s ok" % (m["n"], m["compressed_n"], m["fnv1a32"])
    r = subprocess.run(args, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.strip() == want


def _fnv1a32(data: bytes) -> int:
    h = 2166136261
    for b in data:
        h = ((h ^ b) * 16777619) & 0xFFFFFFFF
    return h
