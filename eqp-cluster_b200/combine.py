"""combine-counts: chunk `.cnt` files -> sample `.cnt` (sum) and sample files -> project matrix.

Host-side mirror of `bin/util-lib/combineCounts.py` as `process-samples.sh` calls it:
  * `combineCounts.py -A -i <chunk .cnt ...> -o <sample .cnt>`                  (process-samples.sh:1730,1739,1748)
  * `combineCounts.py -i <sample .cnt ...> -o <project .cnt> -s <suffix>`       (process-samples.sh:1796,1806,1821)
Same options, same text: ids sorted bytewise, values printed as Python 2 `str(float)` (12 significant digits), the `Id` header row
only for the matrix, zero-fill of ids missing from a file.  `write_counts()` is the short cut after the NCCL reduce: it writes
the file `-A` would have produced from the chunk files, straight from the reduced count vector.

Behaviours of the reference that look like accidents are kept and marked "(sic)", because downstream files depend on them.
"""
import argparse
import gzip
import os
import re
import sys

# fgrep patterns of createFilterFile (combineCounts.py:96-113): a row survives if any value starts with a non-zero digit pattern
FILTER_PATTERNS = ["Id\t"] + ["\t0.0%d" % i for i in range(10)] + ["\t0.%d" % (i + 1) for i in range(9)] + ["\t%d" % (i + 1) for i in range(9)]


def py2_float_str(v):
    """Python 2 `str(float)`: '%.12g', with '.0' appended when the result reads like an integer (floatobject.c, PyFloat_STR_PRECISION 12)."""
    v = float(v)
    s = "%.12g" % v
    if s.lstrip("-").isdigit():
        s += ".0"
    return s


def read_count_file(path, use_filter=False):
    """getCounts (combineCounts.py:122-170): returns (max number of value columns, column headers, {id: [values]}).
    Lines are right-stripped as a whole, split at tabs; an empty id is skipped; the `Id` row supplies the headers."""
    headers = [""]
    counts = {}
    width = 0
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rt", encoding="utf-8", errors="surrogateescape", newline="") as f:   # bytes in, bytes out; no newline translation (Python 2 file semantics)
        lines = f.read().split("\n")
    if use_filter:                                            # `fgrep -f <filter> <file>` then split("\n"): the last element is ""
        lines = [ln for ln in lines if any(p in ln for p in FILTER_PATTERNS)] + [""]
    elif lines and lines[-1] == "":
        lines.pop()                                           # file iteration does not yield an empty last line
    for line in lines:
        fields = line.rstrip().split("\t")
        width = max(width, len(fields) - 1)
        key = fields[0]
        if key == "":
            continue
        if key != "Id":
            counts[key] = [float(v) for v in fields[1:]]
        else:
            headers = fields[1:]
    return width, headers, counts


def combine_counts(input_files, output, add_up=False, multi_column=False, file_suffix="", use_filter=False, filter_file="", id_file="",
                   output_zeroes=True):
    """Main program of combineCounts.py (:178-299).  `output_zeroes` (-n) is accepted and, as in the reference, has no effect:
    the test there compares an array with 0, which is always unequal (sic, :206,:212)."""
    del output_zeroes
    combined = {}
    col_headers = []
    filtering = bool(use_filter or filter_file)               # the reference creates a filter file only when one was asked for (:181-184, :127)
    use_filter_now = bool(use_filter)
    last_key = None
    for n_file, path in enumerate(input_files):
        width, headers, count = read_count_file(path, use_filter_now and filtering)
        if width > 0:
            if len(headers) == 1 and headers[0] == "":
                col_headers.append(re.sub(file_suffix, "", os.path.basename(path)).rstrip(" -_"))
            else:
                col_headers = col_headers + headers
        for key, vals in count.items():
            last_key = key
            if add_up:
                if key in combined:
                    if len(combined[key]) != len(vals):
                        raise ValueError("Length of combined counts %d does not match the length of the added counts %d for %s" % (len(combined[key]), len(vals), key))
                    combined[key] = [0 + a + b for a, b in zip(combined[key], vals)]      # sum((a, b)) starts from integer 0
                else:
                    combined[key] = list(vals)
            else:
                if key in combined:
                    combined[key] = combined[key] + vals
                elif not multi_column:
                    combined[key] = [0.0] * n_file + vals
                else:
                    raise ValueError("Count object %s occurs in file %s but not in the other files and option -M is set." % (key, path))
        if last_key is None:
            raise ValueError("%s has no count rows (the reference stops with a NameError here, combineCounts.py:219)" % path)
        if len(col_headers) != len(combined[last_key]):
            sys.stderr.write("Differing number of columns for header %d and counts %d\n" % (len(col_headers), len(combined[last_key])))
        missing = [k for k in combined if k not in count]
        if missing and multi_column:
            raise ValueError("There are %d count objects missing in file %s and option -M is set." % (len(missing), path))
        if not add_up:
            for k in missing:
                combined[k].append(0.0)
        if not multi_column:
            use_filter_now = True                             # later files may be read through the filter (:235-236), if there is one
    with open(output, "w", encoding="utf-8", errors="surrogateescape", newline="\n") as out:
        if not add_up:
            out.write("Id\t" + "\t".join(col_headers) + "\n")
        if id_file:
            opener = gzip.open if id_file.endswith(".gz") else open
            n_col = len(next(iter(combined.values())))
            with opener(id_file, "rt", encoding="utf-8", errors="surrogateescape", newline="\n") as f:
                for line in f:
                    key = line.split("\t")[0]                 # (sic) an id file with one column keeps its newline in the id (:282)
                    if key in combined:
                        out.write(key + "\t" + "\t".join(py2_float_str(v) for v in combined[key]) + "\n")
                    else:
                        out.write(key + "\t" + "\t".join(["0.0"] * n_col) + "\n")
        else:
            for key in sorted(combined, key=lambda k: k.encode("utf-8", "surrogateescape")):
                out.write(key + "\t" + "\t".join(py2_float_str(v) for v in combined[key]) + "\n")
    return len(combined)


def write_counts(feature_ids, values, path, nonzero_only=False):
    """The sample `.cnt` that `combineCounts.py -A` makes of a sample's chunk files, from the reduced vector (values already in
    reads, i.e. fixed point / 2^38).  Gene and exon chunk files list every feature; junction chunk files list only non-zero
    rows when ComputeCounts runs with its non-zero option (ComputeCounts.java:877-897, compute-counts.sh:1034-1037), so their sum
    lists the junctions that are non-zero in any chunk: pass nonzero_only=True for those."""
    rows = [(i.encode("utf-8", "surrogateescape"), float(v)) for i, v in zip(feature_ids, values) if not (nonzero_only and v == 0)]
    rows.sort(key=lambda r: r[0])
    with open(path, "wb") as out:
        out.write(b"".join(k + b"\t" + py2_float_str(v).encode() + b"\n" for k, v in rows))
    return len(rows)


def write_sample_files(feature_ids, vector, prefix, suffix=".cnt", shift=38, nonzero_junctions=True):
    """Writer-rank step after the sample reduce: `vector` is the reduced gene||exon||junction vector of one sample, either WIDE
    (2n entries: fraction limbs in 2^-shift reads, then integer limbs; include/eqp_count.h "sample level") or a plain chunk vector
    (n entries, 2^-shift reads); `feature_ids` = {mode: [ids]} in the vector's order (CountContext.feature_ids).
    Writes <prefix>-gene<suffix>, -exon, -junction as process-samples.sh:1730-1748 would via combineCounts.py -A."""
    import numpy as np
    n = sum(len(feature_ids[m]) for m in ("gene", "exon", "junction"))
    raw = np.ascontiguousarray(vector).view(np.uint64)
    if raw.size == 2 * n and n:
        v = raw[n:].astype(np.float64) + raw[:n].astype(np.float64) / float(1 << shift)
    else:
        v = raw.astype(np.float64) / float(1 << shift)
    off, paths = 0, {}
    for mode in ("gene", "exon", "junction"):
        ids = feature_ids[mode]
        paths[mode] = "%s-%s%s" % (prefix, mode, suffix)
        write_counts(ids, v[off:off + len(ids)], paths[mode], nonzero_only=(mode == "junction" and nonzero_junctions))
        off += len(ids)
    if off != len(v):
        raise ValueError("vector has %d entries, the annotation has %d features" % (len(v), off))
    return paths


def main(argv=None):
    ap = argparse.ArgumentParser(description="Combine the count files from different chunks or samples (combineCounts.py options)")
    ap.add_argument("-d", type=int, default=0, dest="debug_level")
    ap.add_argument("-n", dest="output_zeroes", action="store_false", default=True)
    ap.add_argument("-A", dest="add_up", action="store_true", default=False)
    ap.add_argument("-M", dest="multi_column", action="store_true", default=False)
    ap.add_argument("-s", dest="file_suffix", default="")
    ap.add_argument("-i", dest="input_files", nargs="+", required=True)
    ap.add_argument("-f", dest="filter_file", default="")
    ap.add_argument("-F", dest="use_filter", action="store_true", default=False)
    ap.add_argument("-I", dest="id_file", default="")
    ap.add_argument("-o", dest="output", required=True)
    a = ap.parse_args(argv)
    combine_counts(a.input_files, a.output, add_up=a.add_up, multi_column=a.multi_column, file_suffix=a.file_suffix, use_filter=a.use_filter,
                   filter_file=a.filter_file, id_file=a.id_file, output_zeroes=a.output_zeroes)
    sys.stderr.write("combineCounts done.\n")


if __name__ == "__main__":
    main()
