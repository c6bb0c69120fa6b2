"""Hot spots of an `ncu --page source --csv --print-source sass` dump: instructions with the most stall samples."""
import csv
import sys


def main(path, ntop=50):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r]
    hdr = rows[hi[0]]
    end = hi[1] - 1 if len(hi) > 1 else len(rows)
    data = [r for r in rows[hi[0] + 1:end] if len(r) == len(hdr)]
    ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    tot = sum(int(r[isamp]) for r in data)
    print("total samples", tot, "instructions", len(data))
    top = sorted(range(len(data)), key=lambda i: -int(data[i][isamp]))[:ntop]
    for i in sorted(top):
        print("%5d %7s %5.1f%% %9s  %s" % (i, data[i][isamp], 100 * int(data[i][isamp]) / tot, data[i][iex], data[i][ia].strip()[:110]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 50)
