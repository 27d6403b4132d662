"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` launch list."""
import collections, csv, re, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
iK, iM, iV, iID = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('ID')
per = collections.OrderedDict()
for r in rows[1:]:
    per.setdefault(r[iID], {'k': r[iK]})[r[iM]] = float(r[iV].replace(',', ''))
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for d in per.values():
    n = re.sub(r'\(.*', '', d['k'])[:64]
    a = agg[n]
    a[0] += 1; a[1] += d.get('gpu__time_duration.sum', 0); a[2] += d.get('dram__bytes_read.sum', 0); a[3] += d.get('dram__bytes_write.sum', 0)
tot = sum(a[1] for a in agg.values())
print("%-66s %5s %10s %7s %9s %9s %8s" % ("kernel", "n", "us/launch", "share", "rd MB", "wr MB", "GB/s"))
for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    t = a[1] / a[0] / 1e3
    print("%-66s %5d %10.1f %6.1f%% %9.1f %9.1f %8.0f" % (n, a[0], t, 100 * a[1] / tot, a[2] / a[0] / 1e6, a[3] / a[0] / 1e6,
                                                         (a[2] + a[3]) / a[0] / 1e3 / t if t else 0))
