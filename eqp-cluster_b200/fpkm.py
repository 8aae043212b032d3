"""FPKM of the project gene matrix (second half): host-side mirror of `bin/util-lib/computeFpkm.py` as
`process-samples.sh:1827-1832` calls it (`-c <project gene .cnt> -g <GTF> -n <aligned reads file> -o <.fpkm> -l <gene lengths>`).

fpkm[id][s] = count / (length / 1000.0) / (alignedReads[s] / 1e6); lengths come from the gene-length file, else from the GTF
(union length of all rows of a gene_id / transcript_id per chromosome, the longest chromosome wins); ids without a length are
left out; rows sorted bytewise; values printed as Python 2 `str(float)`.
"""
import argparse
import sys

from .combine import py2_float_str

_TEXT = dict(encoding="utf-8", errors="surrogateescape", newline="\n")
TRANSCRIPT_PREFIXES = ("NM_", "XM_", "NR_", "XR_", "ENST", "ENSMUST", "ENSRNOT")        # computeFpkm.py:358-359


def read_count_matrix(path):
    """readCountFile (computeFpkm.py:90-114): {id: [values]}, column headers of the `Id` row.  Blank lines become the id ""
    with no values (sic), which only matters when the read numbers are derived from the matrix."""
    counts, headers = {}, [""]
    with open(path, "r", **_TEXT) as f:
        for line in f:
            fields = line.rstrip().split("\t")
            if fields[0] != "Id":
                counts[fields[0]] = [float(v) for v in fields[1:]]
            else:
                headers = fields[1:]
    return counts, headers


def read_num_reads(path):
    """readNumReadsFile (computeFpkm.py:122-147): `sample\\t[chunk\\t]alignedReads`, summed per sample; rows whose sample field
    contains "Sample" are taken for headers and skipped (sic)."""
    reads = {}
    with open(path, "r", **_TEXT) as f:
        for line in f:
            fields = line.rstrip().split("\t")
            if len(fields) >= 3:
                sample, _chunk, aligned = fields              # more than three fields: ValueError, as in the reference
            else:
                sample, aligned = fields
            if "Sample" not in sample:
                reads[sample] = reads.get(sample, 0) + int(aligned)
    return reads


def read_gene_lengths(path):
    """readGeneLengthFile (computeFpkm.py:155-176): `id\\tlength`, header id `Gene Id` skipped."""
    lengths = {}
    with open(path, "r", **_TEXT) as f:
        for line in f:
            fields = line.rstrip().split("\t")
            if fields[0] != "Gene Id":
                lengths[fields[0]] = int(fields[1])
    return lengths


def union_length(intervals):
    """computeLength (computeFpkm.py:239-258): closed intervals sorted by start; an interval is merged into the running one
    unless it starts after the running end (adjacent-but-not-touching intervals stay separate); sum of end - start + 1."""
    total, cur = 0, None
    for s, e in sorted(intervals, key=lambda iv: iv[0]):
        if cur is not None and not cur[1] < s:
            cur[1] = max(cur[1], e)
        else:
            if cur is not None:
                total += cur[1] - cur[0] + 1
            cur = [s, e]
    if cur is not None:
        total += cur[1] - cur[0] + 1
    return total


def read_gtf_lengths(path):
    """readGtfFile + computeCountObjectLengths (computeFpkm.py:184-231, 266-281): every GTF row adds its interval to its
    gene_id and to its transcript_id, per chromosome; an id's length is the largest per-chromosome union length."""
    objects = {}
    with open(path, "r", **_TEXT) as f:
        for line in f:
            g = line.rstrip().split("\t")
            chrom, interval = g[0], (int(g[3]), int(g[4]))
            for entry in g[8].split(";"):
                if entry != "":
                    kind, value = entry.strip().split(" ")    # attributes with spaces in the value: ValueError, as in the reference
                    if kind in ("gene_id", "transcript_id"):
                        objects.setdefault(value.strip('"'), {}).setdefault(chrom, []).append(interval)
    return {k: max([0] + [union_length(iv) for iv in per_chr.values()]) for k, per_chr in objects.items()}


def compute_fpkm(counts, samples, lengths, num_reads):
    """Main program (computeFpkm.py:322-377).  Returns ({id: [fpkm per sample]}, genes without length, transcripts without length)."""
    fpkm, genes_missing, transcripts_missing = {}, set(), set()
    for key, vals in counts.items():
        if key not in lengths:
            (transcripts_missing if key.startswith(TRANSCRIPT_PREFIXES) else genes_missing).add(key)
            continue
        row = [c / (lengths[key] / 1000.0) for c in vals]
        for i, s in enumerate(samples):
            row[i] = row[i] / (num_reads[s] / (1000.0 * 1000.0))
        fpkm[key] = row
    return fpkm, genes_missing, transcripts_missing


def write_fpkm(path, samples, fpkm, compute_mean=False, control_samples=""):
    """outputFpkmFile (computeFpkm.py:288-315).  The mean column is printed like the other values (12 significant digits)."""
    keys = sorted(fpkm, key=lambda k: k.encode("utf-8", "surrogateescape"))
    with open(path, "w", **_TEXT) as out:
        if compute_mean:
            idx = [samples.index(c) for c in control_samples.split(":")] if control_samples != "" else list(range(len(samples)))
            out.write("Gene Id\tMean control FPKM\n")
            for k in keys:
                out.write(k + "\t" + py2_float_str(sum(fpkm[k][i] for i in idx) / len(idx)) + "\n")
        else:
            out.write("\t".join(["Gene Id"] + samples) + "\n")
            for k in keys:
                out.write(k + "\t" + "\t".join(py2_float_str(v) for v in fpkm[k]) + "\n")


def run(count_file, fpkm_file, gene_length_file, read_num_file="", gtf_file="", compute_mean=False, control_samples=""):
    counts, samples = read_count_matrix(count_file)
    lengths = read_gene_lengths(gene_length_file)
    if read_num_file != "":
        num_reads = read_num_reads(read_num_file)
    else:                                                     # normalise by the column sums (computeFpkm.py:329-333)
        num_reads = {s: sum(c[i] for c in counts.values()) for i, s in enumerate(samples)}
    if gtf_file != "":
        gtf_lengths = read_gtf_lengths(gtf_file)
        for key in counts:
            if key not in lengths and key in gtf_lengths:
                lengths[key] = gtf_lengths[key]
    fpkm, genes_missing, transcripts_missing = compute_fpkm(counts, samples, lengths, num_reads)
    if genes_missing:
        sys.stderr.write("There are %d genes without length.\n" % len(genes_missing))
    if transcripts_missing:
        sys.stderr.write("There are %d transcripts without length.\n" % len(transcripts_missing))
    write_fpkm(fpkm_file, samples, fpkm, compute_mean, control_samples)
    return fpkm


def main(argv=None):
    ap = argparse.ArgumentParser(description="FPKM values of a count matrix (computeFpkm.py options)")
    ap.add_argument("-d", type=int, default=0, dest="debug_level")
    ap.add_argument("-c", dest="count_file", required=True)
    ap.add_argument("-n", dest="read_num_file", default="")
    ap.add_argument("-g", dest="gtf_file", default="")
    ap.add_argument("-o", dest="fpkm_file", required=True)
    ap.add_argument("-l", dest="gene_length_file", required=True)
    ap.add_argument("-m", dest="compute_mean", action="store_true")
    ap.add_argument("-C", dest="control_samples", default="")
    a = ap.parse_args(argv)
    run(a.count_file, a.fpkm_file, a.gene_length_file, a.read_num_file, a.gtf_file, a.compute_mean, a.control_samples)


if __name__ == "__main__":
    main()
