"""In-tree builds: libeqpcount.so (nvcc, sm_100a), the C++ host CLI, and the test infrastructure (oracle, worldgen, emu).

Everything is rebuilt only when a source is newer than its output.  Built files are git-ignored but travel to the
GPU box with the gpurun snapshot.
"""
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libeqpcount.so")
CLI_PATH = os.path.join(LIB_DIR, "eqp_compute_counts")
ORACLE_BIN = os.path.join(ROOT, "oracle", "_build", "eqp_oracle")
WORLDGEN_BIN = os.path.join(ROOT, "tools", "_build", "worldgen")
EMU_BIN = os.path.join(ROOT, "tools", "_build", "eqp_emu")
INFLATE_EMU_BIN = os.path.join(ROOT, "tools", "_build", "inflate_emu")
PGZIP_BIN = os.path.join(ROOT, "tools", "_build", "pgzip")

NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "--extended-lambda",
              "-Xcompiler", "-fPIC", "-shared", "-diag-suppress", "550"]


def _newer(sources, out):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(s) > t for s in sources)


def _glob(d, exts):
    return [os.path.join(d, f) for f in sorted(os.listdir(d)) if f.endswith(exts)]


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s" % (" ".join(cmd), r.stdout[-4000:]))
    return r.stdout


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found: libeqpcount.so cannot be built (there is no CPU fallback)")
    return p


def build_library(force=False):
    srcs = _glob(CSRC, (".cu", ".cuh", ".h", ".hpp")) + [os.path.join(ROOT, "include", "eqp_count.h")]
    if force or _newer(srcs, LIB_PATH):
        os.makedirs(LIB_DIR, exist_ok=True)
        _run([nvcc_path()] + NVCC_FLAGS + ["-o", LIB_PATH, os.path.join(CSRC, "eqp_capi.cu"), "-lz", "-lpthread"])
    return LIB_PATH


def build_cli(force=False):
    src = os.path.join(PKG_DIR, "host", "compute_counts_main.cpp")
    if not os.path.exists(src):
        return None
    srcs = _glob(os.path.join(PKG_DIR, "host"), (".cpp", ".hpp")) + [os.path.join(ROOT, "include", "eqp_count.h")]
    if force or _newer(srcs, CLI_PATH):
        os.makedirs(LIB_DIR, exist_ok=True)
        _run(["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"), "-o", CLI_PATH, src, "-ldl", "-lz", "-lpthread"])
    return CLI_PATH


def build_oracle(force=False):
    d = os.path.join(ROOT, "oracle")
    if force or _newer(_glob(d, (".cpp", ".hpp")), ORACLE_BIN):
        os.makedirs(os.path.dirname(ORACLE_BIN), exist_ok=True)
        _run(["g++", "-std=c++17", "-O2", "-o", ORACLE_BIN, os.path.join(d, "eqp_oracle.cpp")])
    return ORACLE_BIN


def build_tools(force=False):
    d = os.path.join(ROOT, "tools")
    os.makedirs(os.path.dirname(WORLDGEN_BIN), exist_ok=True)
    if force or _newer([os.path.join(d, "worldgen.cpp")], WORLDGEN_BIN):
        _run(["g++", "-std=c++17", "-O2", "-o", WORLDGEN_BIN, os.path.join(d, "worldgen.cpp")])
    emu_src = os.path.join(ROOT, "tests", "emu", "eqp_emu.cpp")
    if force or _newer(_glob(CSRC, (".h", ".hpp")) + [emu_src], EMU_BIN):
        _run(["g++", "-std=c++17", "-O2", "-o", EMU_BIN, emu_src, "-lz"])
        _run(["g++", "-std=c++17", "-O2", "-DEQP_FLAT_SCAN=0", "-o", EMU_BIN + "_chain", emu_src, "-lz"])   # duplicate-chain walk for every fragment
        _run(["g++", "-std=c++17", "-O2", "-DEQP_FS_SMALL=2", "-DEQP_FT_SMALL=3", "-o", EMU_BIN + "_retry", emu_src, "-lz"])   # tiny small tables: most fragments take the large-table retry
    zf_src = os.path.join(ROOT, "tests", "emu", "inflate_emu.cpp")
    if force or _newer([zf_src, os.path.join(CSRC, "eqp_inflate.h"), os.path.join(CSRC, "eqp_types.h")], INFLATE_EMU_BIN):
        _run(["g++", "-std=c++17", "-O2", "-o", INFLATE_EMU_BIN, zf_src])
    if force or _newer([os.path.join(d, "pgzip.cpp")], PGZIP_BIN):
        _run(["g++", "-std=c++17", "-O2", "-o", PGZIP_BIN, os.path.join(d, "pgzip.cpp"), "-lz", "-lpthread"])
    return WORLDGEN_BIN, EMU_BIN


def build_all(force=False):
    build_library(force)
    build_cli(force)
    build_oracle(force)
    build_tools(force)
