"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export by CUDA source line: where the instructions and stall samples go."""
import collections
import csv
import sys


def main(path, top=40):
    inst = collections.Counter(); samp = collections.Counter(); text = {}
    cur_file = None
    total_i = total_s = 0
    with open(path, newline="") as fh:
        rd = csv.reader(fh)
        col = None
        for r in rd:
            if not r:
                continue
            if r[0] == "File Path":
                cur_file = r[1].split("/")[-1]; continue
            if r[0] == "Function Name":
                continue
            if r[0] == "Line No":
                col = {h: i for i, h in enumerate(r)}
                # duplicate "Source" header: first = CUDA text, second = SASS
                continue
            if col is None or len(r) < 8:
                continue
            try:
                line = int(r[0])
            except ValueError:
                continue
            # rows for a source line carry totals of its SASS; SASS rows follow with empty line no in some versions
            try:
                n = int(r[col["Instructions Executed"]]); s = int(r[col["# Samples"]])
            except (ValueError, KeyError):
                continue
            key = (cur_file, line)
            inst[key] += n; samp[key] += s; text.setdefault(key, r[1].strip()[:110])
            total_i += n; total_s += s
    print("total warp instructions %d, stall samples %d" % (total_i, total_s))
    print("---- by instructions executed")
    for k, v in inst.most_common(int(top)):
        print("%6.2f%% inst %6.2f%% samp  %s:%d  %s" % (100.0 * v / max(1, total_i), 100.0 * samp[k] / max(1, total_s), k[0], k[1], text[k]))
    byfile = collections.Counter(); sf = collections.Counter()
    for k, v in inst.items():
        byfile[k[0]] += v; sf[k[0]] += samp[k]
    print("---- by file")
    for k, v in byfile.most_common():
        print("%6.2f%% inst %6.2f%% samp  %s" % (100.0 * v / max(1, total_i), 100.0 * sf[k] / max(1, total_s), k))


if __name__ == "__main__":
    main(*sys.argv[1:])
