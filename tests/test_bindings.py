"""The boundary as a non-Python host sees it: a plain C program driving libeqpcount.so through include/eqp_count.h alone
(tests/abi/abi_smoke.c: create -> load -> streamed feed -> count -> reduce -> rows -> destroy), and the JNI glue / Java sources that
ship under bindings/ (no JDK in the image: the glue is type-checked against a stub jni.h, the Java files are checked for the
native methods the glue exports)."""
import os
import re
import shutil
import subprocess

import pytest

import eqp_cluster_b200 as eqp
from conftest import ROOT, assert_rows_match, read_rows, read_stats


def build_abi_smoke(tmp_path):
    lib = eqp.build.build_library()
    exe = os.path.join(str(tmp_path), "abi_smoke")
    r = subprocess.run(["gcc", "-std=c11", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "abi", "abi_smoke.c"), "-o", exe,
                        "-L", os.path.dirname(lib), "-leqpcount", "-Wl,-rpath," + os.path.dirname(lib)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return exe


def chunk_args(world, c):
    sfx = "pe" if c["layout"] == "pe" else "sr"
    return [world.anno] + [os.path.join(c["dir"], "%s-%s.sam" % (k, sfx)) for k in ("transcript", "junction", "genome")]


def test_c_host_fails_loudly_without_gpu(world, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    exe = build_abi_smoke(tmp_path)
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    r = subprocess.run([exe] + chunk_args(world, c) + [str(tmp_path)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 3 and "ERROR" in r.stderr and "no CPU path" in r.stderr and "ABI_SMOKE=ok" not in r.stdout


@pytest.mark.gpu
def test_c_host_through_the_header_alone(world, tmp_path):
    exe = build_abi_smoke(tmp_path)
    c = world.chunk("pe_small", 1003, 1, 3000, "pe")
    r = subprocess.run([exe] + chunk_args(world, c) + [str(tmp_path)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr
    got = dict(l.split("=", 1) for l in r.stdout.splitlines() if "=" in l)
    st = read_stats(os.path.join(c["oracle"], "stats.txt"))
    for k in ("NUMBER_MAPPED_READS", "NUMBER_EXPRESSED_READS", "NUM_WEIGHT_OBJECTS"):
        assert got[k] == st[k], k
    assert got["ABI_SMOKE"] == "ok" and got["NCCL"] == "ok" and int(got["KERNEL_LAUNCHES"]) > 0
    for m in ("gene", "exon", "junction"):
        assert_rows_match(read_rows(os.path.join(str(tmp_path), m + ".rows")), read_rows(os.path.join(c["oracle"], m + ".cnt")), what=m)


def test_jni_glue_type_checks_and_matches_the_java_class(tmp_path):
    inc = os.path.join(str(tmp_path), "inc")
    os.makedirs(inc)
    shutil.copy(os.path.join(ROOT, "tests", "abi", "jni_stub.h"), os.path.join(inc, "jni.h"))
    glue = os.path.join(ROOT, "bindings", "jni", "eqpcount_jni.c")
    r = subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-fsyntax-only", "-DEQP_HAVE_JNI", "-I", inc, "-I", os.path.join(ROOT, "include"), glue],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    r = subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-c", "-o", os.path.join(str(tmp_path), "empty.o"), glue], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout                                   # without -DEQP_HAVE_JNI the translation unit is empty
    exported = set(re.findall(r"Java_EqpCountNative_(\w+)\(", open(glue).read()))
    java = open(os.path.join(ROOT, "bindings", "java", "EqpCountNative.java")).read()
    declared = set(re.findall(r"public static native [\w\[\]]+ (\w+)\(", java))
    assert exported == declared and len(declared) >= 14
    main = open(os.path.join(ROOT, "bindings", "java", "ComputeCountsGpu.java")).read()
    for call in re.findall(r"EqpCountNative\.(\w+)\(", main):
        assert call in declared, call
    header = open(os.path.join(ROOT, "include", "eqp_count.h")).read()
    panama = open(os.path.join(ROOT, "bindings", "java", "EqpCountPanama.java")).read()
    for sym in re.findall(r'fn\("(eqp_\w+)"', panama):
        assert re.search(r"\b%s\s*\(" % sym, header), sym
