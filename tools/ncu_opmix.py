"""Opcode mix (weighted by executed warp instructions) and stall samples of an `ncu --page source --csv` export."""
import collections, csv, re, sys
for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if "Source" in r][0]
    hdr = rows[hi]
    si, ei, sa = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    agg, smp, tot, stot = collections.Counter(), collections.Counter(), 0, 0
    for r in rows[hi + 1:]:
        try:
            n, s = int(r[ei]), int(r[sa])
        except (ValueError, IndexError):
            continue
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[si])
        op = m.group(2).split(".")[0] if m else "?"
        agg[op] += n; smp[op] += s; tot += n; stot += s
    print("== %s | %s : %d warp instructions, %d samples" % (path, rows[0][1][:60], tot, stot))
    for k, v in agg.most_common(22):
        print("   %-10s %11d %5.1f%%   samples %5.1f%%" % (k, v, 100 * v / tot, 100 * smp[k] / max(stot, 1)))
