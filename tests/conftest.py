import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import eqp_cluster_b200 as eqp  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def run(cmd, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, **kw)
    assert r.returncode == 0, "%s\n%s\n%s" % (" ".join(cmd), r.stdout[-2000:], r.stderr[-2000:])
    return r.stdout


@pytest.fixture(scope="session")
def tools():
    """Paths of the test-infrastructure binaries (oracle, worldgen, emu), built on demand."""
    oracle = eqp.build.build_oracle()
    worldgen, emu = eqp.build.build_tools()
    return {"oracle": oracle, "worldgen": worldgen, "emu": emu}


class World:
    """A seeded synthetic world: annotation dir + chunks, each with the oracle's outputs."""

    def __init__(self, tools, root, anno_seed=76, n_genes=300):
        self.tools, self.root = tools, str(root)
        self.anno = os.path.join(self.root, "anno")
        os.makedirs(self.anno, exist_ok=True)
        run([tools["worldgen"], "anno", self.anno, str(anno_seed), str(n_genes)])
        self.chunks = {}

    def chunk(self, name, seed, first, n, layout, spliced=None):
        if name in self.chunks:
            return self.chunks[name]
        d = os.path.join(self.root, name)
        o = os.path.join(d, "oracle")
        os.makedirs(o, exist_ok=True)
        cmd = [self.tools["worldgen"], "reads", self.anno, d, str(seed), str(first), str(n), layout]
        if spliced is not None:
            cmd.append(str(spliced))
        run(cmd)
        run([self.tools["oracle"], "run", self.anno, d, layout, o, "--quiet"])
        self.chunks[name] = {"dir": d, "oracle": o, "layout": layout, "n": n}
        return self.chunks[name]


    def oracle_run(self, name, flags):
        """Oracle outputs of a chunk under extra switches (--unweighted, --unambig, --containment, --nonsplice)."""
        c = self.chunks[name]
        o = os.path.join(c["dir"], "oracle" + "".join(flags).replace("--", "_"))
        if not os.path.exists(os.path.join(o, "stats.txt")):
            os.makedirs(o, exist_ok=True)
            run([self.tools["oracle"], "run", self.anno, c["dir"], c["layout"], o, "--quiet"] + list(flags))
        return o

    def emu_run(self, name, flags=(), variant=""):
        """TEST-ONLY host emulation of the device functions (tests/emu): id -> value per mode + stats."""
        c = self.chunks[name]
        o = os.path.join(c["dir"], "emu" + variant + "".join(flags).replace("--", "_"))
        os.makedirs(o, exist_ok=True)
        run([self.tools["emu"] + variant, self.anno, c["dir"], "pe" if c["layout"] == "pe" else "se", o] + list(flags))
        out = {}
        for m in ("gene", "exon", "junction"):
            out[m] = {}
            for line in open(os.path.join(o, m + ".tsv")):
                k, fx, v = line.rstrip("\n").split("\t")
                out[m][k] = float(v)
        out["stats"] = read_stats(os.path.join(o, "stats.txt"))
        return out


@pytest.fixture(scope="session")
def world(tools, tmp_path_factory):
    return World(tools, tmp_path_factory.mktemp("world"))


def read_sams(chunk_dir, layout):
    sfx = "pe" if layout == "pe" else "sr"
    return [open(os.path.join(chunk_dir, "%s-%s.sam" % (k, sfx)), "rb").read() for k in ("transcript", "junction", "genome")]


def read_cnt(path):
    out = {}
    with open(path) as f:
        for line in f:
            k, v = line.rstrip("\n").split("\t")
            out[k] = float(v)
    return out


def read_stats(path):
    return dict(line.strip().split("=") for line in open(path) if "=" in line)


def assert_counts_match(got, want, rel=1e-9, exact=False, what=""):
    """got/want: id -> value.  exact=True (integer counts: unweighted runs) demands bit-equality; otherwise fractional
    counts must agree within the north-star tolerance of 1e-9 relative (BASELINE.json)."""
    bad = []
    for k in set(got) | set(want):
        a, b = got.get(k, 0.0), want.get(k, 0.0)
        if a == b:
            continue
        if exact or abs(a - b) > rel * max(abs(a), abs(b)):
            bad.append((k, a, b))
    assert not bad, "%s: %d mismatches, first %s" % (what, len(bad), bad[:5])
