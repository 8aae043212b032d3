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

    def __init__(self, tools, root, anno_seed=76, n_genes=300, knobs=()):
        """knobs: worldgen annotation-shape settings ("ld=0.2", "tr=0.2", ...; tools/worldgen.cpp); the defaults are the SURVEY 8d mix."""
        self.tools, self.root = tools, str(root)
        self.anno = os.path.join(self.root, "anno")
        os.makedirs(self.anno, exist_ok=True)
        run([tools["worldgen"], "anno", self.anno, str(anno_seed), str(n_genes)] + list(knobs))
        self.chunks = {}

    def chunk(self, name, seed, first, n, layout, spliced=None, flags=()):
        """flags: oracle switches the base run of this chunk uses (e.g. --stranded=forward)."""
        if name in self.chunks:
            return self.chunks[name]
        d = os.path.join(self.root, name)
        o = os.path.join(d, "oracle")
        os.makedirs(o, exist_ok=True)
        cmd = [self.tools["worldgen"], "reads", self.anno, d, str(seed), str(first), str(n), layout]
        if spliced is not None:
            cmd.append(str(spliced))
        run(cmd)
        run([self.tools["oracle"], "run", self.anno, d, layout, o, "--quiet"] + list(flags))
        self.chunks[name] = {"dir": d, "oracle": o, "layout": layout, "n": n, "flags": tuple(flags)}
        return self.chunks[name]


    def genomic_chunk(self, name, seed, first, n, layout, flags=()):
        """One SAM of a spliced genome aligner (worldgen ... genomic) + the oracle's compute-genomic-counts.sh restatement."""
        if name in self.chunks:
            return self.chunks[name]
        d = os.path.join(self.root, name)
        o = os.path.join(d, "oracle")
        os.makedirs(o, exist_ok=True)
        run([self.tools["worldgen"], "reads", self.anno, d, str(seed), str(first), str(n), layout, "genomic"])
        sam = os.path.join(d, "genomic-%s.sam" % ("pe" if layout == "pe" else "sr"))
        run([self.tools["oracle"], "genomic", self.anno, sam, o, "--quiet"] + list(flags))
        self.chunks[name] = {"dir": d, "oracle": o, "layout": layout, "n": n, "flags": tuple(flags), "sam": sam}
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
            out[m + "_rows"] = [(l.split("\t")[0], float(l.rstrip("\n").split("\t")[1])) for l in open(os.path.join(o, m + ".rows"))]
        out["stats"] = read_stats(os.path.join(o, "stats.txt"))
        out["weights"] = {l.split()[0]: (int(l.split()[1]), int(l.split()[2])) for l in open(os.path.join(o, "weights.tsv"))}
        return out


@pytest.fixture(scope="session")
def world(tools, tmp_path_factory):
    return World(tools, tmp_path_factory.mktemp("world"))


SHAPES = {                                  # one dense world per annotation shape of SURVEY 8d, and the mix
    "ld": ("ld=0.3", "tr=0", "inv=0", "ver2=0"),                       # transcript exon longer / shorter than its genome exon (a13, a17)
    "tr": ("tr=0.3", "ld=0", "inv=0", "ver2=0"),                       # tr- filler exons (a12, a16, a18, a20)
    "trlong": ("tr=0.3", "trlong=1", "ld=0", "inv=0", "ver2=0"),       # fillers longer than a read: halves and alignments of tr- intervals only
    "inv": ("inv=0.3", "ld=0", "tr=0", "ver2=0"),                      # .inv-strand references (a12)
    "ver2": ("ver2=0.5", "ld=0", "tr=0", "inv=0"),                     # transcripts listed under two versioned ids (a8, a9)
    "map3": ("map3=1",),                                               # 3-column junction maps (a24, a26)
    "mixed": ("ld=0.2", "tr=0.2", "trlong=0.3", "inv=0.1", "ver2=0.2"),
}


@pytest.fixture(scope="session")
def shape_worlds(tools, tmp_path_factory):
    cache = {}

    def get(shape):
        if shape not in cache:
            cache[shape] = World(tools, tmp_path_factory.mktemp("world_" + shape), knobs=SHAPES[shape])
        return cache[shape]
    return get


def read_rows(path):
    return [(l.split("\t")[0], float(l.rstrip("\n").split("\t")[1])) for l in open(path)]


def assert_rows_match(got, want, rel=1e-9, exact=False, what=""):
    """Count-file rows in print order: same ids in the same order, values equal (exact) or within 1e-9 relative."""
    assert [r[0] for r in got] == [r[0] for r in want], "%s: row ids differ (%d vs %d rows)" % (what, len(got), len(want))
    bad = [(g, w) for g, w in zip(got, want) if g[1] != w[1] and (exact or abs(g[1] - w[1]) > rel * max(abs(g[1]), abs(w[1])))]
    assert not bad, "%s: %d mismatches, first %s" % (what, len(bad), bad[:5])


def read_weights(oracle_dir):
    """{fragment: (a20 of the weight pass, n of final.wgt)} from the oracle's bed.wgt / final.wgt."""
    bed = {l.split("\t")[0]: int(l.split("\t")[1]) for l in open(os.path.join(oracle_dir, "bed.wgt"))}
    fin = {l.split("\t")[0]: int(l.split("\t")[1]) for l in open(os.path.join(oracle_dir, "final.wgt"))}
    return {f: (n, fin[f]) for f, n in bed.items()}


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
